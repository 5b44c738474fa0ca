mkdir -p gpurun_out
UGB_BATCH_TRACE=1 timeout 500 python bench.py --steps 4 --warmup 3 --no-cpu --no-verify > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err; echo "bench rc=$?"
grep "\[batch\]" gpurun_out/bench_trace.err | head -60
python -c "
import json;d=json.load(open('gpurun_out/bench_trace.json'));print(d['value'],d['e2e']['ms_per_proof'],d['prove_latency_ms'])"
