#!/usr/bin/env python
"""Per-CUDA-line hot spots of one kernel from an .ncu-rep (needs -lineinfo and --import-source on).
usage: python tools_srcprof.py report.ncu-rep kernel_regex [top]"""
import csv
import os
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + kern, "--launch-count", "1"], capture_output=True, text=True).stdout
fname, hdr, items = "", None, []
for r in csv.reader(out.splitlines()):
    if len(r) >= 2 and r[0] == "File Path":
        fname = os.path.basename(r[1])
    elif len(r) > 8 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) >= len(hdr) - 2 and r[2] == "-":
        try:
            smp, ins = float(r[6]), float(r[7])
        except ValueError:
            continue
        stalls = {hdr[i]: float(r[i]) for i in range(32, min(49, len(r))) if r[i] not in ("", "0", "-")}
        items.append((smp, ins, fname, r[0], r[1].strip()[:100], stalls))
ts, ti = sum(x[0] for x in items), sum(x[1] for x in items)
print("total warp-instr %.0f  samples %.0f" % (ti, ts))
items.sort(key=lambda x: -x[0])
for smp, ins, f, ln, src, st in items[:top]:
    s3 = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print("%5.1f%% smp %5.1f%% ins %-18s:%-4s %-100s %s" % (100 * smp / max(ts, 1), 100 * ins / max(ti, 1), f[:18], ln, src,
                                                        " ".join("%s=%d" % (k.replace("stall_", ""), v) for k, v in s3)))
