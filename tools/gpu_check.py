#!/usr/bin/env python
"""Diagnostic sweep (not a test): every case x {fused, generic}, all tensors, no early exit.
Writes gpurun_out/parity_report.txt.  A sticky CUDA error (misaligned address, illegal access)
poisons the whole process, so the sweep runs in a worker that is restarted after the combo that
killed it."""
import os
import subprocess
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REPORT = os.path.join(ROOT, "gpurun_out", "parity_report.txt")


def combos(only):
    from tests import rigs

    out = []
    for case in rigs.CASES:
        if only and not any(o in case["name"] for o in only):
            continue
        for fg in (False, True):
            out.append((case, fg))
    return out


def worker(start, only):
    from tests.common import compare_case

    todo = combos(only)
    with open(REPORT, "a") as fh:
        for i in range(start, len(todo)):
            case, fg = todo[i]
            tag = "%s[%s]" % (case["name"], "generic" if fg else "fused")
            try:
                rows = compare_case(case, force_generic=fg)
                for n, e, g in rows:
                    flag = "" if (e <= 1e-5 and (g is None or g <= 2e-5)) else "   <<<<<< FAIL"
                    fh.write("%-34s %-22s oracle64 %.3e  golden %s%s\n" % (
                        tag, n, e, "%.3e" % g if g is not None else "-", flag))
            except Exception:
                fh.write("%-34s EXCEPTION   <<<<<< FAIL\n%s\n" % (tag, traceback.format_exc()))
                fh.flush()
                print("DIED %d" % i, flush=True)
                os._exit(3)
            fh.flush()
            print("DONE %d" % i, flush=True)


def main():
    os.makedirs(os.path.dirname(REPORT), exist_ok=True)
    args = sys.argv[1:]
    if args and args[0] == "--worker":
        worker(int(args[1]), args[2:] or None)
        return
    only = args or None
    open(REPORT, "w").close()
    n = len(combos(only))
    start = 0
    while start < n:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", str(start)] + (only or []),
                           capture_output=True, text=True, timeout=900)
        last = -1
        for line in r.stdout.splitlines():
            if line.startswith(("DONE", "DIED")):
                last = int(line.split()[1])
        if r.returncode != 0 and last < start and "DIED" not in r.stdout:
            with open(REPORT, "a") as fh:
                fh.write("combo %d: worker crashed rc=%d   <<<<<< FAIL\n%s\n" % (start, r.returncode, r.stderr[-2000:]))
            last = start
        start = last + 1
    txt = open(REPORT).read()
    nfail = txt.count("<<<<<< FAIL")
    with open(REPORT, "a") as fh:
        fh.write("FAILURES %d\n" % nfail)
    print(txt[-5000:])
    print("FAILURES %d" % nfail)


if __name__ == "__main__":
    main()
