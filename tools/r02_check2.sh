# round 2, second pass: whole GPU suite, launch list of the timed proofs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -22 gpurun_out/pytest_gpu.log
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
