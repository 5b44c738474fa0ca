# round 2, first pass: the whole GPU suite (incl. the full-size bit-exact parity tests), both bench arms
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
timeout 500 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench.err
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'],d['e2e']['value'],d['stage_ms'],d['create_ms'],d['roofline'],d.get('cpu_baseline'))"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 300 gpurun_out/bench_ref.err; cut -c1-600 gpurun_out/bench_ref.json
