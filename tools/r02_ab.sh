# A/B of two library builds on the whole proof: $1 = tag of the variant library (ultragroth_b200/libultragroth_b200_$1.so)
mkdir -p gpurun_out
echo "== default library"
timeout 600 python -m pytest tests/test_gpu_msm.py tests/test_gpu_prover.py "tests/test_gpu_fullsize.py::test_config1_groth16_2pow20_bit_exact" -x -q 2>&1 | tail -2
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_def.json 2> gpurun_out/bench_def.err; python -c "
import json;d=json.load(open('gpurun_out/bench_def.json'));print(d['value'],d['e2e']['value'],d['stage_ms'],d['roofline']['frac'])"
echo "== variant $1"
UGB_LIB=$PWD/ultragroth_b200/libultragroth_b200_$1.so timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err; python -c "
import json;d=json.load(open('gpurun_out/bench_$1.json'));print(d['value'],d['e2e']['value'],d['stage_ms'],d['roofline']['frac'])"
