mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_field.py -x -q 2>&1 | tail -5
python - <<'PY' 2>&1 | tee gpurun_out/dfma_micro.txt
import ultragroth_b200 as ug
ctx = ug.Context(0)
names = {0: "IMAD", 1: "IMAD.WIDE plain", 3: "IMAD.WIDE carry chain", 4: "IMAD.WIDE carry out", 2: "Fq modmul CIOS (fp_mul)", 5: "Fq modmul Karatsuba (fp_mul_ks)", 6: "DFMA (fma.rz.f64)", 7: "Fq modmul FP64-assisted (fp_mul_dfma)"}
for k in (0, 1, 3, 4, 6, 2, 5, 7):
    v = max(ctx.bench_int_pipe(k) for _ in range(3))
    print("%-42s %.3f G/s" % (names[k], v / 1e9))
PY
