mkdir -p gpurun_out /tmp/p
echo "== correctness with UGB_NTT_TMA=1"
UGB_NTT_TMA=1 timeout 900 python -m pytest tests/test_gpu_ntt.py "tests/test_gpu_fullsize.py::test_ntt_full_size_vs_cpu_oracle" tests/test_gpu_prover.py -x -q 2>&1 | tail -4
echo "== default"
timeout 600 python sweep.py --msm "" --ntt 18,20,22,24 --reps 9 2>/dev/null | grep ntt_fr | cut -c1-200 | tee gpurun_out/ntt_default.jsonl
for th in 256 512 1024; do
echo "== TMA contiguous passes, $th threads"
UGB_NTT_TMA=1 UGB_NTT_TMA_THREADS=$th timeout 600 python sweep.py --msm "" --ntt 18,20,22,24 --reps 9 2>/dev/null | grep ntt_fr | cut -c1-200 | tee gpurun_out/ntt_tma_$th.jsonl
done
cat > gpurun_out/ntt_one.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import ultragroth_b200 as ug
ctx = ug.Context(0)
rng = np.random.default_rng(1)
v = rng.integers(0, 256, size=(1 << 22, 32), dtype=np.uint8); v[:, 31] &= 0x1f
ds = [ctx.upload(v) for _ in range(3)]
for _ in range(2):
    ctx.coset_transform(ds, 22)
ctx.sync()
PY
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_ntt_pass" --launch-skip 4 -c 4 -o /tmp/p/ntt_default python gpurun_out/ntt_one.py > gpurun_out/ncu_ntt_default.log 2>&1
UGB_NTT_TMA=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_ntt_pass" --launch-skip 4 -c 4 -o /tmp/p/ntt_tma python gpurun_out/ntt_one.py > gpurun_out/ncu_ntt_tma.log 2>&1
for f in ntt_default ntt_tma; do ncu -i /tmp/p/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv; done
python - <<'PY'
import csv
for f in ("ntt_default", "ntt_tma"):
    rows = list(csv.reader(open(f"gpurun_out/{f}_raw.csv")))
    hdr = rows[0]
    keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]
    idx = [hdr.index(k) for k in keys if k in hdr]
    print(f, [hdr[i] for i in idx])
    for r in rows[2:]:
        print("   ", [r[i][:40] for i in idx])
PY
