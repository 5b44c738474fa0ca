mkdir -p gpurun_out /tmp/p
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_ntt_pass_bulk" --launch-skip 2 -c 2 -o /tmp/p/ntt_tma python tools/ntt_one.py tma > gpurun_out/ncu_ntt_tma.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_ntt_pass" --launch-skip 8 -c 2 -o /tmp/p/ntt_def python tools/ntt_one.py > gpurun_out/ncu_ntt_def.log 2>&1
ncu -i /tmp/p/ntt_tma.ncu-rep --page raw --csv > gpurun_out/ntt_tma_raw.csv
ncu -i /tmp/p/ntt_def.ncu-rep --page raw --csv > gpurun_out/ntt_def_raw.csv
python - <<'PY'
import csv
for f in ("ntt_def", "ntt_tma"):
    rows = list(csv.reader(open(f"gpurun_out/{f}_raw.csv")))
    hdr = rows[0]
    keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "launch__registers_per_thread"]
    idx = [hdr.index(k) for k in keys if k in hdr]
    print(f, [hdr[i] for i in idx])
    for r in rows[1:]:
        print("   ", [r[i][:44] for i in idx])
PY
