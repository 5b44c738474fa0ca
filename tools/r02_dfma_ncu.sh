mkdir -p gpurun_out /tmp/p

timeout 600 ncu --set full --clock-control none -k regex:"k_bench_modmul$|k_bench_modmul_dfma|k_bench_dfma" --launch-skip 2 -c 9 -o /tmp/p/dfma python tools/micro_pipes.py > gpurun_out/ncu_dfma.log 2>&1
ncu -i /tmp/p/dfma.ncu-rep --page raw --csv > gpurun_out/dfma_raw.csv
python - <<'PY'
import csv
rows = list(csv.reader(open('gpurun_out/dfma_raw.csv')))
hdr = rows[0]
want = [i for i, h in enumerate(hdr) if any(s in h for s in ("Kernel Name", "gpu__time_duration.sum", "sm__inst_executed_pipe_fmaheavy", "sm__inst_executed_pipe_fp64", "sm__inst_executed_pipe_alu", "sm__inst_executed_pipe_fma.", "sm__pipe_fmaheavy_cycles_active", "sm__pipe_fp64_cycles_active", "sm__pipe_alu_cycles_active", "sm__pipe_fma_cycles_active", "smsp__issue_active.avg.pct", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__inst_executed.sum", "sm__throughput.avg.pct"))]
for r in rows[:2] + rows[2:]:
    print(" | ".join(r[i][:60] for i in want))
PY
