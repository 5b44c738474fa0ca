mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prover.py tests/test_gpu_sharded.py -x -q > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu5.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench5.json 2> gpurun_out/bench5.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench5.err
python -c "
import json;d=json.load(open('gpurun_out/bench5.json'));print(d['value'],d['e2e'],d['prove_latency_ms'])"
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu --workload groth16 --logn 20 > gpurun_out/bench5_g16.json 2> gpurun_out/bench5_g16.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench5_g16.json'));print(d['value'],d['e2e'],d['prove_latency_ms'])"
