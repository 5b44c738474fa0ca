# round 2 multi-GPU evidence at N GPUs: the driver's bench line (replicas + sharded sub-record) and the configs[4] sweep
mkdir -p gpurun_out
export MASTER_ADDR=127.0.0.1
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_final_n$N.json 2> gpurun_out/bench_final_n$N.err; echo "bench rc=$?"; tail -c 400 gpurun_out/bench_final_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 sweep.py --msm 16,18,20,22,24,26 --ntt 16,18,20,22,24,26 --reps 5 > gpurun_out/sweep_n$N.jsonl 2> gpurun_out/sweep_n$N.err; echo "sweep rc=$?"; tail -c 400 gpurun_out/sweep_n$N.err
python - <<PY
import json
for line in open("gpurun_out/bench_final_n$N.json"):
    if line.startswith("{"):
        d = json.loads(line)
        print(d["value"], d["e2e"]["value"], d.get("sharded"))
for line in open("gpurun_out/sweep_n$N.jsonl"):
    if line.startswith("{"):
        d = json.loads(line)
        print({k: d[k] for k in ("op", "log2n", "n_gpus", "ms_device", "ms_end_to_end", "points_per_s", "ms_median", "elements_per_s") if k in d})
PY
