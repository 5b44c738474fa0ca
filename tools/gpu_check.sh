mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.err
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
timeout 300 python bench.py --steps 3 --warmup 3 --shard --workload groth16 --logn 22 > gpurun_out/bench_shard1.json 2> gpurun_out/bench_shard1.err; tail -c 600 gpurun_out/bench_shard1.err; cat gpurun_out/bench_shard1.json | cut -c1-400
