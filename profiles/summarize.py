#!/usr/bin/env python
"""Turns the ncu outputs a gpurun call brought back (gpurun_out/) into the small text summaries kept under
profiles/ (the .ncu-rep files themselves are tens of MiB and stay in scratch).

    python profiles/summarize.py launches gpurun_out/launches.csv           > profiles/rNN_launches.txt
    python profiles/summarize.py full     gpurun_out/prof_msm.ncu-rep       > profiles/rNN_ncu_msm.txt
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "smsp__sass_inst_executed_op_local_ld.sum",
    "smsp__sass_inst_executed_op_local_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]
STALL = "smsp__average_warps_issue_stalled_"


def short(name):
    return re.sub(r"\(.*", "", name).replace("void ", "")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 14]
    hdr, rows = rows[0], rows[1:]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows:
        a = agg.setdefault(short(r[ki]), [0, 0])
        a[0] += 1
        a[1] += int(float(r[vi]))
    tot = sum(v[1] for v in agg.values())
    print(f"# {len(rows)} launches, {tot / 1e6:.3f} ms of kernel time (gpu__time_duration.sum, cold-cache, serialised)")
    print(f"{'kernel':40s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {v[0]:8d} {v[1] / 1e6:10.3f} {v[1] / tot:7.4f}")


def full(path):
    # a .ncu-rep, or the `ncu -i x.ncu-rep --page raw --csv` text exported on the GPU box (reports are > 64 MiB)
    if path.endswith(".csv"):
        out = open(path).read()
    else:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, rows = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    for r in rows:
        print("== " + short(r[ki]))
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:75s} {r[i]:>18s} {units[i]}")
        st = []
        for i, h in enumerate(hdr):
            if h.startswith(STALL) and h.endswith("_per_issue_active.ratio"):
                st.append((float(r[i]), h[len(STALL):-len("_per_issue_active.ratio")]))
        print("  warp stalls per issued instruction: " + ", ".join(f"{n} {v:.2f}" for v, n in sorted(st, reverse=True)[:7]))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
