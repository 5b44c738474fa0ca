mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-verify > gpurun_out/ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 300 python bench.py --workload groth16 --logn 20 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_g16_20.json 2> gpurun_out/bench_g16_20.err; echo "g16 2^20 rc=$?"
timeout 300 python sweep.py --msm 16,18,20,22,24,26 --ntt 16,18,20,22,24,26 > gpurun_out/sweep_g1.jsonl 2> gpurun_out/sweep.err; echo "sweep rc=$?"; tail -c 200 gpurun_out/sweep.err
tail -c 300 gpurun_out/bench_full.err; cut -c1-200 gpurun_out/bench_full.json
