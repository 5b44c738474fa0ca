# two GPUs, final build: the in-process two-device tests (skipped on one-GPU boxes), the sharded tests, and the driver's N=2 bench line
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
timeout 600 python -m pytest tests/test_gpu_multi_device.py tests/test_gpu_sharded.py -x -q -m gpu -rs > gpurun_out/pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2_final.json 2> gpurun_out/bench_n2_final.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_n2_final.err
python -c "
import json
for line in open('gpurun_out/bench_n2_final.json'):
    if line.startswith('{'):
        d=json.loads(line); print(d['value'], d['e2e']['value'], d['sharded'])"
