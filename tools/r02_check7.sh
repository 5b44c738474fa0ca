# no-alloc MSM entry point alone in a fresh process, then the whole GPU suite (it collects last), then per-shard phase times (8-way, one GPU)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -15 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu7.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu7.log
timeout 600 python tools/shard_phase_times.py --world 8 --logn 22 > gpurun_out/shard_phase_times_w8.txt 2>&1; echo "phase rc=$?"; tail -12 gpurun_out/shard_phase_times_w8.txt
