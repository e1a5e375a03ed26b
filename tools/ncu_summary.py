#!/usr/bin/env python
"""Markdown table of the per-kernel metrics of an ncu --set full report (one column per kernel launch).
usage: python tools_ncu_summary.py report.ncu-rep"""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[0], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
M = [("gpu__time_duration (us, under ncu)", "gpu__time_duration.sum", 1.0, "%.1f"),
     ("dram__bytes_read.sum (MB)", "dram__bytes_read.sum", 1.0, "%.1f"),
     ("dram__bytes_write.sum (MB)", "dram__bytes_write.sum", 1.0, "%.1f"),
     ("smsp__inst_executed.sum (M warp instr)", "smsp__inst_executed.sum", 1e-6, "%.1f"),
     ("issue active (% of peak)", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0, "%.1f"),
     ("warps active (% of 64/SM)", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0, "%.1f"),
     ("registers / thread", "launch__registers_per_thread", 1.0, "%.0f"),
     ("grid", "launch__grid_size", 1.0, "%.0f"),
     ("waves / SM", "launch__waves_per_multiprocessor", 1.0, "%.2f"),
     ("shared-memory wavefronts (M)", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 1e-6, "%.1f"),
     ("shared wavefronts, % of peak", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", 1.0, "%.1f"),
     ("FMA pipe active (%)", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", 1.0, "%.1f"),
     ("long-scoreboard stall / issue", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", 1.0, "%.2f")]
units = rows[1]
names = [r[col["Kernel Name"]].split("(")[0][:28] for r in data]
print("| metric | " + " | ".join(names) + " |")
print("|---|" + "---|" * len(names))
for label, key, scale, fmt in M:
    if key not in col:
        continue
    vals = []
    for r in data:
        try:
            v = float(r[col[key]].replace(",", "")) * scale
            u = units[col[key]]
            if key.startswith("dram__bytes") and u == "Gbyte":
                v *= 1000.0
            if key.startswith("dram__bytes") and u == "Kbyte":
                v /= 1000.0
            if key.startswith("dram__bytes") and u == "byte":
                v /= 1e6
            if key.startswith("gpu__time") and u in ("ns", "nsecond"):
                v /= 1000.0
            if key.startswith("gpu__time") and u in ("ms", "msecond"):
                v *= 1000.0
            vals.append(fmt % v)
        except ValueError:
            vals.append(r[col[key]])
    print("| " + label + " | " + " | ".join(vals) + " |")
