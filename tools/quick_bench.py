#!/usr/bin/env python
"""Quick device timing of SMAL fwd / fwd+bwd (not the contract bench; see bench.py)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smilify_b200  # noqa: E402
from smilify_b200.smal_torch import SMAL  # noqa: E402
from tests import rigs  # noqa: E402

rig = sys.argv[1] if len(sys.argv) > 1 else "stick"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
smilify_b200.configure(rigs.REAL_RIGS[rig])
dev = torch.device("cuda:0")
model = SMAL(dev)
dd = rigs.get_rig(rig)
inp = rigs.make_inputs(dd, B, seed=0, with_scale=("scale" in sys.argv), with_trans_off=("scale" in sys.argv))
t = {k: torch.from_numpy(v).to(dev).requires_grad_(True) for k, v in inp.items()}
V, J = model.size[0], model._tables["J"]
gv = torch.randn(B, V, 3, device=dev)
gj = torch.randn(B, J, 3, device=dev)


def step(bwd=True):
    verts, joints, Rs, vs = model(t["beta"], t["theta"], trans=t["trans"], betas_logscale=t.get("betas_logscale"),
                                  betas_trans=t.get("betas_trans"))
    if bwd:
        torch.autograd.backward([verts, joints], [gv, gj])


for bwd in (False, True):
    for _ in range(5):
        step(bwd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 30
    t0 = time.time()
    e0.record()
    for _ in range(n):
        step(bwd)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("%s B=%d %s: %.3f ms/iter (wall %.3f)  -> %.2f M meshes/s" % (rig, B, "fwd+bwd" if bwd else "fwd", ms,
          (time.time() - t0) / n * 1e3, B / ms / 1e3))
