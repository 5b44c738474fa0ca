# four GPUs, final build: the driver's N=4 bench line (replicas + sharded sub-record)
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/bench_n4_final.json 2> gpurun_out/bench_n4_final.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_n4_final.err
python -c "
import json
for line in open('gpurun_out/bench_n4_final.json'):
    if line.startswith('{'):
        d=json.loads(line); print(d['value'], d['e2e']['value'], d['sharded'])"
