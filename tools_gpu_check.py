#!/usr/bin/env python
"""Diagnostic sweep (not a test): every case x {fused, generic}, all tensors, no early exit.
Writes gpurun_out/parity_report.txt."""
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from tests import rigs  # noqa: E402
from tests.common import compare_case  # noqa: E402

os.makedirs("gpurun_out", exist_ok=True)
only = sys.argv[1:] or None
with open("gpurun_out/parity_report.txt", "w") as fh:
    worst = 0.0
    for case in rigs.CASES:
        if only and not any(o in case["name"] for o in only):
            continue
        for fg in (False, True):
            tag = "%s[%s]" % (case["name"], "generic" if fg else "fused")
            try:
                rows = compare_case(case, force_generic=fg)
                for n, e, g in rows:
                    flag = "" if (e <= 1e-5 and (g is None or g <= 2e-5)) else "   <<<<<< FAIL"
                    fh.write("%-34s %-22s oracle64 %.3e  golden %s%s\n" % (tag, n, e, "%.3e" % g if g is not None else "-", flag))
                    worst = max(worst, e if e == e else 1e9)
            except Exception:
                fh.write("%-34s EXCEPTION\n%s\n" % (tag, traceback.format_exc()))
                worst = 1e9
            fh.flush()
    fh.write("WORST %.3e\n" % worst)
print(open("gpurun_out/parity_report.txt").read()[-6000:])
