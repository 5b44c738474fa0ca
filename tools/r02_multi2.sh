mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"; tail -c 1500 gpurun_out/bench_n2.err
python -c "
import json;d=json.load(open('gpurun_out/bench_n2.json'));print(d['value'],d['e2e']['value'],d.get('sharded'))"
