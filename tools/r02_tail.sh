# host tail of a proof with its independent scalar multiplications on helper threads: parity tests and the bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_prover.py tests/test_gpu_sharded.py tests/test_gpu_synth.py tests/test_dropin.py -x -q -m gpu > gpurun_out/pytest_gpu_tail.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_tail.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_tail.json 2> gpurun_out/bench_tail.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_tail.json'));print(d['value'],d['e2e']['value'],d['e2e']['sequential'],d['stage_ms']['total'])"
