#!/usr/bin/env python
"""Per-kernel device times (library event pairs, eager) + graph-replayed step time of one fused fwd+bwd.
Development tool (tuning sweeps); the contract numbers come from bench.py.
usage: python tools_gpu/kbench.py [rig] [B] [tag]      (tuning switches are read from the environment)"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smilify_b200  # noqa: E402
from smilify_b200 import _lib  # noqa: E402
from smilify_b200.graphed import GraphedStep  # noqa: E402
from smilify_b200.smal_torch import SMAL  # noqa: E402
from tests import rigs  # noqa: E402

rig = sys.argv[1] if len(sys.argv) > 1 else "stick"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
tag = sys.argv[3] if len(sys.argv) > 3 else ""
smilify_b200.configure(rigs.REAL_RIGS[rig])
dev = torch.device("cuda:0")
model = SMAL(dev)
V, J, nB = model._tables["V"], model._tables["J"], model.num_betas
rs = np.random.RandomState(0)
x = dict(beta=rs.standard_normal((B, nB)) * 0.5, theta=rs.standard_normal((B, J, 3)) * 0.3, trans=rs.standard_normal((B, 3)),
         g_verts=rs.standard_normal((B, V, 3)), g_joints=rs.standard_normal((B, J, 3)))
x = {k: torch.from_numpy(v.astype(np.float32)).to(dev) for k, v in x.items()}
beta, theta, trans = (x[k].clone().requires_grad_(True) for k in ("beta", "theta", "trans"))


def eager():
    verts, joints, Rs, vs = model(beta, theta, trans=trans)
    return torch.autograd.grad([verts, joints], [beta, theta, trans], [x["g_verts"], x["g_joints"]])


for _ in range(5):
    eager()
torch.cuda.synchronize()
n = 30
_lib.profile_enable(16 * n)
for _ in range(n):
    eager()
torch.cuda.synchronize()
prof = _lib.profile_collect()
_lib.profile_enable(0)
gs = GraphedStep(model, B)
with torch.no_grad():
    for k in ("beta", "theta", "trans", "g_verts", "g_joints"):
        getattr(gs, k).copy_(x[k])
for _ in range(10):
    gs.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 100
e0.record()
for _ in range(K):
    gs.replay()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
n_in = nB + 3 * J + 3
bytes_mesh = 4 * (n_in + 6 * V + 12 * J) + 4 * (3 * V + 3 * J + 2 * n_in)
out = {"tag": tag, "rig": rig, "B": B, "ms_step": round(ms, 4), "Mmesh_s": round(B / ms / 1e3, 3),
       "GBps_step": round(bytes_mesh * B / ms / 1e6, 1), "slots": int(model._dm.info.num_ga_slots) if hasattr(model, "_dm") else None,
       "kernels_us": {k: round(v[1] / v[0] * 1e3, 2) for k, v in prof.items()},
       "env": {k: v for k, v in os.environ.items() if k.startswith("SMIL_")}}
print(json.dumps(out), flush=True)
