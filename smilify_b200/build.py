"""Build ``libsmil_b200.so`` in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m smilify_b200.build [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "smil_b200.cu")
OUT = os.path.join(HERE, "libsmil_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC", "--use_fast_math=false"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def sources():
    d = os.path.join(HERE, "csrc")
    inc = os.path.join(os.path.dirname(HERE), "include", "smil_b200.h")
    return [os.path.join(d, f) for f in sorted(os.listdir(d))] + [inc]


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    return all(os.path.getmtime(s) <= t for s in sources())


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    flags = [f for f in NVCC_FLAGS if not f.startswith("--use_fast_math")]
    if os.environ.get("SMIL_SKIN_BWD_G"):   # tuning: samples per warp pass of the fused backward (8 | 16 | 32)
        flags.append("-DSMIL_SKIN_BWD_G=%d" % int(os.environ["SMIL_SKIN_BWD_G"]))
    if os.environ.get("SMIL_FAST_BUILD"):    # development: only the nB = 3 / 5 instances of the fused kernels
        flags.append("-DSMIL_FAST_BUILD")
    cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building %s" % OUT)
    if verbose:
        sys.stderr.write(r.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
