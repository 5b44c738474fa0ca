mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q > gpurun_out/pytest_gpu4.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu4.log
timeout 300 python bench.py --steps 3 --warmup 3 --shard --workload groth16 --logn 20 --no-cpu > gpurun_out/bench_shard1.json 2> gpurun_out/bench_shard1.err; echo "rc=$?"; tail -c 600 gpurun_out/bench_shard1.err; cut -c1-300 gpurun_out/bench_shard1.json
