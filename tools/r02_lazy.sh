mkdir -p gpurun_out
python - <<'PY' 2>&1 | tee gpurun_out/fq2_micro.txt
import os, sys
sys.path.insert(0, os.getcwd())
import ultragroth_b200 as ug
ctx = ug.Context(0)
for k, name in ((2, "Fq product (CIOS)"), (8, "Fq2 product, 3 complete Fq products"), (9, "Fq2 product, lazy reduction")):
    v = max(ctx.bench_int_pipe(k) for _ in range(3))
    print("%-42s %.3f G/s" % (name, v / 1e9))
PY
echo "== default library"
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_def.json 2> gpurun_out/bench_def.err; python -c "
import json;d=json.load(open('gpurun_out/bench_def.json'));print(d['value'],d['e2e']['value'],d['stage_ms']['msm_b2'],d['stage_ms']['total'],d['create_ms'])"
echo "== UG_FQ2_LAZY library"
UGB_LIB=$PWD/ultragroth_b200/libultragroth_b200_lazy.so timeout 600 python -m pytest tests/test_gpu_msm.py tests/test_gpu_prover.py "tests/test_gpu_fullsize.py::test_config1_groth16_2pow20_bit_exact" -x -q 2>&1 | tail -3
UGB_LIB=$PWD/ultragroth_b200/libultragroth_b200_lazy.so timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_lazy.json 2> gpurun_out/bench_lazy.err; python -c "
import json;d=json.load(open('gpurun_out/bench_lazy.json'));print(d['value'],d['e2e']['value'],d['stage_ms']['msm_b2'],d['stage_ms']['total'],d['create_ms'])"
