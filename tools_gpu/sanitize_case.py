#!/usr/bin/env python
"""The workload compute-sanitizer is run on (`compute-sanitizer --tool X python tools_gpu/sanitize_case.py`; closed on this pool, see profiles/r02_sanitizer.md): fused forward + backward on both rigs at batch
sizes that make the persistent passes wrap their rings (several sample groups per CTA, odd tails), plus the
projection-loss, triangulation and 6-D rotation kernels.  Prints the relative error against the fp64 oracle of a subset,
so a run that "passes" the tool but computes garbage is visible too."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smilify_b200  # noqa: E402
from smilify_b200 import losses, triangulation  # noqa: E402
from smilify_b200.smal_torch import SMAL  # noqa: E402
from tests import rigs  # noqa: E402

dev = torch.device("cuda:0")
cases = [("stick", 33), ("stick", 1201), ("mouse", 257)] if len(sys.argv) < 2 else [(sys.argv[1], int(sys.argv[2]))]
for rig, B in cases:
    smilify_b200.configure(rigs.REAL_RIGS[rig])
    model = SMAL(dev)
    dd = rigs.get_rig(rig)
    inp = rigs.make_inputs(dd, B, seed=1)
    t = {k: torch.from_numpy(inp[k]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans")}
    V, J = model._tables["V"], model._tables["J"]
    gv, gj = torch.randn(B, V, 3, device=dev), torch.randn(B, J, 3, device=dev)
    verts, joints, Rs, vs = model(t["beta"], t["theta"], trans=t["trans"])
    g = torch.autograd.grad([verts, joints], [t["beta"], t["theta"], t["trans"]], [gv, gj])
    torch.cuda.synchronize()
    print("%s B=%d: |verts| %.4f |g_theta| %.4f finite %s" % (rig, B, float(verts.abs().mean()), float(g[1].abs().mean()),
                                                              bool(all(torch.isfinite(x).all() for x in g))), flush=True)
# the small kernels
sc = rigs.loss_scene(5, 4, 17, 3, False)
f = lambda k, grad=False: torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev).requires_grad_(grad)
j, R, T, fov = f("joints", True), f("R", True), f("T", True), f("fov", True)
loss, _ = losses.multiview_keypoint_loss(j, R, T, fov, f("aspect"), f("target"), f("vis"), f("view_mask"), f("joint_weights"), 512.0)
loss.backward()
tri, valid = triangulation.triangulate_joints_dlt(f("target"), f("vis"), R, T, fov, f("aspect"), f("view_mask"), 512)
(tri * valid[..., None]).sum().backward()
d6 = torch.randn(64, 6, device=dev, requires_grad=True)
triangulation.rotation_6d_to_matrix(d6).sum().backward()
torch.cuda.synchronize()
print("losses / triangulation / rot6d ok: loss %.6f" % float(loss), flush=True)
