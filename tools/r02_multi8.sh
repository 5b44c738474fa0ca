mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?"; tail -c 800 gpurun_out/bench_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --shard --workload groth16 --logn 24 --no-cpu > gpurun_out/bench_shard_g16_24_n$N.json 2> gpurun_out/bench_shard_g16_24_n$N.err; echo "rc=$?"; tail -c 800 gpurun_out/bench_shard_g16_24_n$N.err
python - <<PY
import json
for f in ("gpurun_out/bench_n$N.json", "gpurun_out/bench_shard_g16_24_n$N.json"):
    for line in open(f):
        if line.startswith("{"):
            d = json.loads(line)
            print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("sharded"), d.get("create_s"), d.get("setup_s"))
PY
