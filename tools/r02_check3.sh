mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_prover.py tests/test_gpu_synth.py tests/test_gpu_sharded.py -x -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu3.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench3.json 2> gpurun_out/bench3.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench3.err
python -c "
import json;d=json.load(open('gpurun_out/bench3.json'));print(d['value'],d['e2e']['value'],d['stage_ms'],d['create_ms'],d['inf_points'],d['roofline']['frac'], d['kernel_share'])"
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu --workload groth16 --logn 20 > gpurun_out/bench3_g16.json 2> gpurun_out/bench3_g16.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench3_g16.json'));print(d['value'],d['e2e']['value'],d['stage_ms'],d['create_ms'],d['inf_points'])"
