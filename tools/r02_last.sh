# last check of the committed tree on one GPU: whole GPU suite (the no-alloc test collects last), smoke()
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_last.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_last.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
