#!/usr/bin/env python
"""BASELINE configs[3] on real GPUs: SMALFitter-style joint fit, frames sharded over the ranks, ONE packed NCCL
all-reduce of the shared-parameter gradients (betas, fov) per step.

Launch (N = 2, 4, 8):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools_gpu/fitter_multigpu.py --frames 65536 --steps 10 --warmup 3
or, for the single-GPU line, ``python tools_gpu/fitter_multigpu.py``.

Every frame's parameters and targets are generated from the GLOBAL frame index, so every world size solves the same
problem.  After the timed region rank 0 re-evaluates ALL frames alone (8192-frame chunks, gradients accumulated) and
reports how far the all-reduced shared gradients and the loss of the sharded run are from that single-process result.
OmniAnt's pickle is not in the repository: the default rig is its stand-in "stick13" (STICK
geometry, J = 55, shape space padded to nB = 13 by tests/rigs.py::padded_rig).
Prints one JSON line on rank 0."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smilify_b200  # noqa: E402
from smilify_b200 import _lib  # noqa: E402
from smilify_b200 import fitter_step as FS  # noqa: E402
from smilify_b200.smal_torch import SMAL  # noqa: E402
from tests import rigs  # noqa: E402


def frame_block(lo, hi, J, block=4096):
    """Per-frame parameters / targets of global frames [lo, hi): seeded per 4096-frame block of the GLOBAL index."""
    out = {k: [] for k in ("global_rotation", "joint_rotations", "trans", "target", "vis")}
    b0 = (lo // block) * block
    while b0 < hi:
        rs = np.random.RandomState(1000 + b0 // block)
        gr = rs.normal(0, 0.3, (block, 3)).astype(np.float32)
        jr = rs.normal(0, 0.3, (block, J - 1, 3)).astype(np.float32)
        tr = rs.normal(0, 0.1, (block, 3)).astype(np.float32)
        tg = rs.uniform(0, 512, (block, J, 2)).astype(np.float32)
        vi = (rs.uniform(size=(block, J)) < 0.8).astype(np.float32)
        s, e = max(lo, b0) - b0, min(hi, b0 + block) - b0
        for k, a in zip(out, (gr, jr, tr, tg, vi)):
            out[k].append(a[s:e])
        b0 += block
    return {k: np.concatenate(v) for k, v in out.items()}


def prior_setup(J, nB, dev):
    """Synthetic constants of the parameter-space terms (fitter.py:298-350) for the full-step record: joint limits,
    the identity pose prior of fitter.py:25-52 (kernel default), a betas precision, masks; the weights of a late
    optimisation stage."""
    rs = np.random.RandomState(11)
    f = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32)).to(dev)
    consts = dict(max_limits=f(rs.uniform(0.2, 0.8, (J - 1, 3))), min_limits=f(-rs.uniform(0.2, 0.8, (J - 1, 3))),
                  betas_mean=f(np.zeros(nB)), betas_prec=f(np.eye(nB) + 0.05 * rs.standard_normal((nB, nB))),
                  rotation_mask=f(np.ones((J - 1, 3))), global_mask=f(np.ones((1, 3))))
    weights = dict(w_limit=1.0, w_pose=0.5, w_splay=0.01, w_betas=0.3, w_temp=0.2)
    return consts, weights


def run_fitter(frames, steps, warmup, rank, world, dev, rig="stick13", graph=True, get_skin=True, check=True, full_step=False):
    """One frame-sharded SMALFitter-style step timed on this rank's device (the process group, if any, is initialised
    by the caller).  Returns the result dict on rank 0 (None elsewhere).  ``rig``: "stick13" is the OmniAnt stand-in of
    SURVEY.md section 8(d) config 4 (STICK geometry, shape space padded to nB = 13)."""
    from tests.common import rig_file

    smilify_b200.configure(rig_file(rig))
    model = SMAL(dev)
    J, nB = model._tables["J"], model.num_betas
    # the fitter's default camera, look_at_view_transform(2.7, 0, 0) (p3d_renderer.py:34-37)
    R = torch.diag(torch.tensor([-1.0, 1.0, -1.0]))[None].to(dev)
    T = torch.tensor([[0.0, 0.0, 2.7]], device=dev)

    def build(lo, hi):
        d = frame_block(lo, hi, J)
        fp = {k: torch.from_numpy(d[k]).to(dev).requires_grad_(True) for k in ("global_rotation", "joint_rotations", "trans")}
        rs = np.random.RandomState(7)
        sp = dict(betas=torch.from_numpy(rs.normal(0, 0.5, (nB,)).astype(np.float32)).to(dev).requires_grad_(True),
                  fov=torch.tensor([60.0], device=dev, requires_grad=True))
        fn = FS.smal_joint_loss_fn(model, R, T, torch.from_numpy(d["target"]).to(dev), torch.from_numpy(d["vis"]).to(dev),
                                   get_skin=bool(get_skin))
        return fp, sp, fn

    lo, hi = FS.shard_range(frames, rank, world)
    fp, sp, fn = build(lo, hi)
    optimizer = None
    if full_step:
        # (f)2: + joint limits / pose prior / splay / betas prior / temporal smoothness (1-frame halos between the
        # shards) + the fused Adam of optimize_to_joints.py:117-126, all inside the captured step
        from smilify_b200.fitter_terms import FusedAdam

        consts, weights = prior_setup(J, nB, dev)
        fn = FS.with_prior_terms(fn, frames, lo, weights, window=10, **consts)
        optimizer = FusedAdam([{"params": [fp["global_rotation"], fp["joint_rotations"], fp["trans"], sp["betas"]], "lr": 5e-3},
                               {"params": [sp["fov"]], "lr": 1.0}], lr=5e-3)
    c0 = _lib.launch_count()
    loss0 = FS.ShardedFitterStep(fp, sp, fn, frames)()   # one eager step: counts the kernels a step launches
    per_step = _lib.launch_count() - c0
    g_first = {k: v.grad.detach().clone() for k, v in sp.items()}   # at the INITIAL parameters (the optimiser moves them)
    loss_first = float(loss0)
    kw = dict(optimizer=optimizer) if (graph and optimizer is not None) else {}
    step = (FS.GraphedShardedFitterStep if graph else FS.ShardedFitterStep)(fp, sp, fn, frames, **kw)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        loss = step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / steps
    # the collective alone (same packed buffer size), device-timed
    ar_us = None
    if world > 1:
        buf = torch.zeros(nB + 2, device=dev)
        for _ in range(5):
            dist.all_reduce(buf)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(20):
            dist.all_reduce(buf)
        a1.record()
        torch.cuda.synchronize()
        ar_us = a0.elapsed_time(a1) / 20 * 1e3
        t = torch.tensor([ms, ar_us], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ar_us = float(t[0].item()), float(t[1].item())
    g_sharded, loss_sharded = g_first, loss_first
    loss_last = float(loss)

    # single-process run over ALL frames on rank 0 (chunked accumulation): what the sharded gradients must equal
    chk = None
    if rank == 0 and world > 1 and check:
        acc = {k: torch.zeros_like(v) for k, v in sp.items()}
        tot = 0.0
        for c0 in range(0, frames, 8192):
            fpc, spc, fnc = build(c0, min(frames, c0 + 8192))
            part = fnc(fpc, spc) / frames
            part.backward()
            tot += float(part.detach())
            for k in acc:
                acc[k] += spc[k].grad
        if full_step:   # the parameter-space terms over the WHOLE sequence in one call (no halos needed)
            from smilify_b200.fitter_terms import fitter_prior_terms

            fpa, spa, _ = build(0, frames)
            pr = fitter_prior_terms(fpa["global_rotation"], fpa["joint_rotations"], fpa["trans"], spa["betas"], weights, window=10, **consts)
            pr.backward()
            tot += float(pr.detach())
            acc["betas"] += spa["betas"].grad
        chk = {k: float((g_sharded[k] - acc[k]).abs().max() / acc[k].abs().max().clamp_min(1e-30)) for k in acc}
        chk["loss"] = abs(loss_sharded - tot) / abs(tot)
    if rank != 0:
        return None
    in_graph = bool(getattr(step, "collective_in_graph", False))
    return {
        "metric": ("SMALFitter-style frames/s (SMAL fwd+bwd + joint reprojection loss + limits / pose prior / splay / betas "
                   "prior / temporal terms + fused Adam, shared betas/fov)") if full_step else
                  "SMALFitter-style frames/s (SMAL fwd+bwd + joint reprojection loss, shared betas/fov)",
        "full_step": bool(full_step), "optimizer_in_graph": bool(getattr(step, "_opt_in_graph", False)), "loss_last": loss_last,
        "value": frames / ms * 1e3, "frames_per_s": frames / ms * 1e3, "unit": "frames/s", "n_gpus": world, "ms_per_step": ms,
        "steps": steps, "warmup": warmup, "frames_total": frames, "frames_per_gpu": hi - lo, "scaling": "strong",
        "collective": "1 all_reduce(SUM) per step, %d bytes (g_betas, g_fov, loss)" % step.collective_bytes,
        "collective_bytes": step.collective_bytes, "allreduce_us": ar_us,
        "backend": "nccl" if world > 1 else "none", "rig": "%s (J = %d, nB = %d)" % (rig, J, nB), "get_skin": bool(get_skin),
        "launch": ("cuda-graph replay incl. the NCCL all_reduce" if in_graph else "cuda-graph replay + eager all_reduce")
                  if graph else "eager",
        "gpu_launches": per_step * steps, "loss": loss_sharded, "sharded_vs_single_rel_err": chk}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--graph", type=int, default=1, help="1: replay the step from a CUDA graph")
    ap.add_argument("--get-skin", type=int, default=1, help="1: pose the mesh too, as fitter.py:262 does")
    ap.add_argument("--rig", default="stick13", help="stick13 = OmniAnt stand-in (nB padded to 13), stick, mouse40, ...")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    out = run_fitter(args.frames, args.steps, args.warmup, rank, world, dev, rig=args.rig, graph=bool(args.graph),
                     get_skin=bool(args.get_skin))
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
