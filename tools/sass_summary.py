#!/usr/bin/env python
"""SASS mnemonic counts and resource usage per kernel of the built library (static counts, not executed ones).
usage: python tools/sass_summary.py [lib.so] > profiles/rNN_sass_summary.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "smilify_b200", "libsmil_b200.so")
KEEP = [r"pose_fwd_ls_kernelILi8", r"skin_fwd_kernelILi5", r"skin_bwd_vm_kernelILi5", r"pose_bwd_ls_kernelILi8", r"pose_fwd_kernel",
        r"pose_bwd_lean_kernel", r"skin_bwd_kernelILi5", r"fitter_prior", r"adam_step",
        r"slot_reduce", r"skin_fwd_kernelILi3", r"skin_bwd_vm_kernelILi3", r"keypoint|proj_", r"tri_fwd|tri_bwd", r"rot6d"]
COLS = ["UBLKCP", "SYNCS", "FFMA", "FFMA2", "FMUL2|FADD2", "LDS", "STS", "LDG", "STG", "REDG|RED|ATOMG|ATOM", "BAR", "SHFL", "MUFU", "DFMA|DADD|DMUL"]

res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
usage = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"REG:(\d+).*?SHARED:(\d+).*?LOCAL:(\d+)", line)
    if m and cur:
        usage[cur] = tuple(int(x) for x in m.groups())
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
counts = collections.defaultdict(collections.Counter)
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        counts[cur][m.group(1)] += 1
        counts[cur]["_total"] += 1


def demangle(n):
    m = re.match(r"_ZN4smil(\d+)", n)
    if not m:
        return n
    k = int(m.group(1))
    base = n[len(m.group(0)):len(m.group(0)) + k]
    t = re.findall(r"Li(\d+)E", n[len(m.group(0)) + k:])
    return base + ("<" + ",".join(t) + ">" if t else "")


print("# SASS summary of `%s` (sm_100a, static instruction counts)\n" % os.path.basename(lib))
print("Made by `python tools/sass_summary.py`.  `UBLKCP` = `cp.async.bulk` (TMA bulk copy), `SYNCS` = mbarrier operations,")
print("`FFMA2` / `FMUL2|FADD2` = packed two-lane fp32 arithmetic (`fma.rn.f32x2`), `RED` = `red.global.add(.v4).f32`,")
print("`DFMA..` = fp64 (projection losses / triangulation only).  LOCAL = bytes of local memory (0 = no spills).\n")
print("| kernel | regs | local B | SASS instr | " + " | ".join(COLS) + " |")
print("|---|---|---|---|" + "---|" * len(COLS))
seen = set()
for pat in KEEP:
    for fn in sorted(counts):
        if re.search(pat, fn) and fn not in seen:
            seen.add(fn)
            c = counts[fn]
            u = usage.get(fn, (0, 0, 0))
            vals = [sum(v for k, v in c.items() if re.fullmatch(col, k)) for col in COLS]
            print("| `%s` | %d | %d | %d | %s |" % (demangle(fn), u[0], u[2], c["_total"], " | ".join(str(v) for v in vals)))
print("\nAll %d kernels of the library: %d use local memory (spills)." % (len(counts), sum(1 for f in counts if usage.get(f, (0, 0, 0))[2] > 0)))
