#!/usr/bin/env python
"""BASELINE configs[2] / configs[4] timed on real GPUs: the neuralSMIL multi-view training-loss path
(SMAL forward inside the regressor loss -> root-centred world placement -> B x V cameras -> row-10 loss, SURVEY.md
section 8a) forward + backward to the per-frame parameters, frames sharded over the ranks, NO collective (in the
reference the only exchange of this path is DDP's gradient all-reduce of the network weights, which is not part of it).

  configs[2]: python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
                  tools_gpu/multiview_bench.py --rig mouse --static 1 --views 18 --frames 8192
  configs[4]: python tools_gpu/multiview_bench.py --rig stick --views 5 --frames 1024 --pca 1 --propagate 1

Parity of exactly this composition against the oracle is tests/test_gpu_fullsize.py::test_multiview_training_loss_fullsize;
this script only times it (CUDA events, max over ranks) -- eagerly and replayed from one CUDA graph."""
import argparse
import json
import math
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smilify_b200  # noqa: E402
from smilify_b200 import _lib, losses  # noqa: E402
from smilify_b200.smal_torch import SMAL  # noqa: E402
from tests import rigs  # noqa: E402


def ring_cameras(nviews, radius=3.0, elev_deg=15.0):
    """look_at_view_transform(radius, elev, azim) for cameras on a ring (tests/test_triangulation_consistency.py:73-107),
    PyTorch3D row-vector convention: R = [x y z] as columns, T = -R^T C."""
    Rs, Ts = [], []
    e = math.radians(elev_deg)
    for i in range(nviews):
        a = 2 * math.pi * i / nviews
        C = np.array([radius * math.cos(e) * math.sin(a), radius * math.sin(e), radius * math.cos(e) * math.cos(a)])
        z = -C / np.linalg.norm(C)
        x = np.cross(np.array([0.0, 1.0, 0.0]), z)
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        R = np.stack([x, y, z], axis=1)
        Rs.append(R)
        Ts.append(-R.T @ C)
    return np.stack(Rs).astype(np.float32), np.stack(Ts).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rig", default="mouse")
    ap.add_argument("--static", type=int, default=1)
    ap.add_argument("--views", type=int, default=18)
    ap.add_argument("--frames", type=int, default=8192, help="total over all ranks")
    ap.add_argument("--pca", type=int, default=0)
    ap.add_argument("--propagate", type=int, default=0)
    ap.add_argument("--get-skin", type=int, default=1, help="1: the SMAL call also poses the mesh (the reference's default)")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = smilify_b200.configure(rigs.REAL_RIGS[args.rig])
    cfg.ALLOW_LIMB_SCALING = True
    cfg.ignore_hardcoded_body = True
    cfg.STATIC_JOINT_LOCATIONS = bool(args.static)
    model = SMAL(dev)
    dd = rigs.get_rig(args.rig)
    J, nB, V = model._tables["J"], model.num_betas, model._tables["V"]
    n = args.frames // world
    rs = np.random.RandomState(100 + rank)
    f32 = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32)).to(dev)
    t = dict(beta=f32(rs.normal(0, 0.5, (n, nB))), theta=f32(rs.normal(0, 0.3, (n, J, 3))), trans=f32(rs.normal(0, 0.05, (n, 3))))
    if args.pca:
        t["scale_w"] = f32(rs.normal(0, 0.5, (n, nB)))
        t["trans_w"] = f32(rs.normal(0, 0.5, (n, nB)))
        sd = f32(np.asarray(dd["scaledirs"])).permute(1, 2, 0).contiguous()
        td = f32(np.asarray(dd["transdirs"])).permute(1, 2, 0).contiguous()
    for v in t.values():
        v.requires_grad_(True)
    R, T = ring_cameras(args.views)
    cR = f32(R)[None].expand(n, -1, -1, -1).contiguous()
    cT = f32(T)[None].expand(n, -1, -1).contiguous()
    fov = torch.full((n, args.views), 60.0, device=dev)
    asp = torch.ones(n, args.views, device=dev)
    tgt = f32(rs.uniform(0.2, 0.8, (n, args.views, J, 2)))
    vis = f32(rs.uniform(size=(n, args.views, J)) < 0.8)

    def step():
        for v in t.values():
            v.grad = None
        ls = bt = None
        if args.pca:
            ls, bt = losses.pca_joint_values(t["scale_w"], t["trans_w"], sd, td)
        out = model(t["beta"], t["theta"], betas_logscale=ls, betas_trans=bt, get_skin=bool(args.get_skin),
                    propagate_scaling=bool(args.propagate))
        joints = out[1] if args.get_skin else out
        world_j = (joints - joints[:, :1]) * 10.0 * 0.1 + t["trans"][:, None]      # multiview_smil_regressor.py:1522-1604
        loss, _ = losses.multiview_keypoint_loss(world_j, cR, cT, fov, asp, tgt, vis, None, None, 512.0)
        loss.backward()
        return loss.detach()

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        if world > 1:
            x = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(x, op=dist.ReduceOp.MAX)
            ms = float(x.item())
        return ms

    c0 = _lib.launch_count()
    loss = step()
    per_step = _lib.launch_count() - c0
    ms_eager = timed(step)
    # fresh leaves for the captured step: the AccumulateGrad nodes of the eager leaves live on the default stream,
    # which a capture must not touch
    import gc
    gc.collect()
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        for k in list(t):
            t[k] = t[k].detach().clone().requires_grad_(True)
        for _ in range(3):
            step()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        loss_g = step()
    ms_graph = timed(g.replay)
    finite = bool(torch.isfinite(loss_g).item()) and all(bool(torch.isfinite(v.grad).all().item()) for v in t.values() if v.grad is not None)
    # bytes the SMAL call moves by contract + the projection-loss inputs, per frame
    in_f = nB + 3 * J + 3 + (6 * J if args.pca else 0)
    smal = 4 * (in_f + (6 * V if args.get_skin else 0) + 3 * J + (9 * J if args.get_skin else 0)) + 4 * (3 * J + 2 * in_f)
    proj = 4 * (3 * J + args.views * 14 + args.views * J * 3) + 4 * (3 * J + args.views * 14)
    if rank == 0:
        print(json.dumps({
            "metric": "multi-view training-loss frames/s (SMAL fwd + %d-camera reprojection loss, fwd+bwd)" % args.views,
            "value": n * world / ms_graph * 1e3, "unit": "frames/s", "n_gpus": world, "frames_total": n * world,
            "frames_per_gpu": n, "ms_per_step": ms_graph, "ms_per_step_eager": ms_eager, "steps": args.steps, "warmup": args.warmup,
            "launch": "cuda-graph replay (eager beside it)", "rig": args.rig, "V": V, "J": J, "nB": nB, "views": args.views,
            "static_joints": bool(args.static), "pca_scale_trans": bool(args.pca), "propagate_scaling": bool(args.propagate),
            "get_skin": bool(args.get_skin), "collective": "none", "gpu_launches_per_step": per_step,
            "alg_bytes_per_frame": smal + proj,
            "achieved_GBps_per_gpu": (smal + proj) * n / ms_graph * 1e-6, "loss": float(loss_g), "finite": finite}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
