# balanced shards: sharded / MSM / prover tests + the 2^22 UltraGroth full-size parity test, then the per-shard phase times of an
# 8-way sharded 2^22 UltraGroth prover with all shards on one GPU
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_msm.py tests/test_gpu_prover.py tests/test_gpu_multi_device.py "tests/test_gpu_fullsize.py::test_config2_ultragroth_2pow22_bit_exact" -x -q > gpurun_out/pytest_gpu6.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu6.log
timeout 900 python tools/shard_phase_times.py --world 8 --logn 22 > gpurun_out/shard_phase_times_w8.txt 2>&1; echo "phase rc=$?"; cat gpurun_out/shard_phase_times_w8.txt | tail -12
