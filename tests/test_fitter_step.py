"""Frame-sharded shared-parameter step.

CPU (gloo, world size 2): the host logic -- contiguous shards, ONE packed all-reduce carrying the
shared gradients and the loss -- reproduces the single-process gradients, with the oracle standing
in for the CUDA call.  GPU: the same step driven by the CUDA path equals the oracle's gradients."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import projection_oracle as PO
from oracle import smal_oracle as SO
from smilify_b200 import fitter_step as FS
from smilify_b200 import losses as L
from tests import rigs

N_FRAMES = 11  # odd on purpose: ragged shards (6 + 5)


def _problem(dd, n, seed=3, dtype=torch.float64):
    rs = np.random.RandomState(seed)
    J = np.asarray(dd["weights"]).shape[1]
    nB = np.asarray(dd["shapedirs"]).shape[-1]
    t = lambda a: torch.from_numpy(np.asarray(a)).to(dtype)
    frame = dict(global_rotation=t(rs.normal(0, 0.3, (n, 3))), joint_rotations=t(rs.normal(0, 0.3, (n, J - 1, 3))),
                 trans=t(rs.normal(0, 0.1, (n, 3))))
    shared = dict(betas=t(rs.normal(0, 0.5, (nB,))), fov=t([60.0]))
    R, T = PO.look_at_default(2.7, dtype=dtype)
    target = t(rs.uniform(0, 512, (n, J, 2)))
    vis = t(rs.uniform(size=(n, J)) < 0.8)
    return frame, shared, R, T, target, vis


def _oracle_loss_fn(model, R, T, target, vis):
    def fn(fp, sp):
        n = fp["trans"].shape[0]
        theta = torch.cat([fp["global_rotation"].unsqueeze(1), fp["joint_rotations"]], dim=1)
        _, joints, _, _ = SO.smal_forward(model, sp["betas"].unsqueeze(0).expand(n, -1), theta)
        loss, _ = PO.fitter_joint_loss(joints, fp["trans"], R, T, sp["fov"].expand(n), target, vis, 1.0, 512.0)
        return loss * n
    return fn


def _single_process_reference(dd, dtype=torch.float64):
    frame, shared, R, T, target, vis = _problem(dd, N_FRAMES, dtype=dtype)
    model = SO.build_model(dd, dtype=dtype)
    fp = {k: v.clone().requires_grad_(True) for k, v in frame.items()}
    sp = {k: v.clone().requires_grad_(True) for k, v in shared.items()}
    step = FS.ShardedFitterStep(fp, sp, _oracle_loss_fn(model, R, T, target, vis), N_FRAMES)
    loss = step()
    return loss, {k: v.grad.clone() for k, v in fp.items()}, {k: v.grad.clone() for k, v in sp.items()}


def _worker(rank, world, port, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        dd = rigs.get_rig("syn_k3")
        frame, shared, R, T, target, vis = _problem(dd, N_FRAMES)
        lo, hi = FS.shard_range(N_FRAMES, rank, world)
        model = SO.build_model(dd, dtype=torch.float64)
        fp = {k: v[lo:hi].clone().requires_grad_(True) for k, v in frame.items()}
        sp = {k: v.clone().requires_grad_(True) for k, v in shared.items()}
        step = FS.ShardedFitterStep(fp, sp, _oracle_loss_fn(model, R, T, target[lo:hi], vis[lo:hi]), N_FRAMES)
        loss = step()
        out_q.put((rank, lo, hi, float(loss), {k: v.grad.numpy() for k, v in fp.items()},
                   {k: v.grad.numpy() for k, v in sp.items()}, step.collective_bytes))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_range_covers_everything():
    for n in (1, 7, 8, 65536):
        for w in (1, 2, 3, 8):
            spans = [FS.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        FS.shard_range(4, 2, 2)


def test_pack_roundtrip_and_none():
    like = [torch.zeros(5), torch.zeros(1), torch.zeros(1, 4, 3)]
    g = [torch.arange(5.0), None, torch.ones(1, 4, 3)]
    buf = FS.pack_grads(g, like)
    assert buf.shape == (18,) and buf.dtype == torch.float32
    out = FS.unpack_grads(buf, like)
    assert torch.equal(out[0], g[0]) and torch.equal(out[1], torch.zeros(1)) and out[2].shape == (1, 4, 3)


@pytest.mark.timeout(300)
def test_gloo_world2_matches_single_process():
    dd = rigs.get_rig("syn_k3")
    ref_loss, ref_fg, ref_sg = _single_process_reference(dd)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    nB = ref_sg["betas"].numel()
    for rank, lo, hi, loss, fg, sg, nbytes in res:
        assert nbytes == 4 * (nB + 1 + 1)            # betas + fov + the loss value: ONE small buffer
        assert abs(loss - float(ref_loss)) <= 1e-6 * abs(float(ref_loss))      # fp32 buffer
        for k in fg:                                                           # per-frame grads never left the rank
            np.testing.assert_allclose(fg[k], ref_fg[k][lo:hi].numpy(), rtol=1e-9, atol=1e-12)
        for k in sg:
            np.testing.assert_allclose(sg[k], ref_sg[k].numpy(), rtol=2e-6, atol=1e-9)


def test_pca_joint_values_matches_oracle():
    rs = np.random.RandomState(0)
    J, C, B = 9, 4, 6
    sd, td = torch.from_numpy(rs.normal(0, 0.3, (J, 3, C))), torch.from_numpy(rs.normal(0, 0.3, (J, 3, C)))
    sw, tw = torch.from_numpy(rs.normal(0, 2.0, (B, C))), torch.from_numpy(rs.normal(0, 1.0, (B, C)))
    a = L.pca_joint_values(sw, tw, sd, td)
    b = SO.pca_weights_to_joint_values(sd, td, sw, tw)
    assert torch.allclose(a[0], b[0]) and torch.allclose(a[1], b[1])
    assert (sd * 0 + 1).sum() > 0 and torch.isfinite(a[0]).all()   # the clamp kept log() finite
    z = L.pca_joint_values(sw, None, sd, td)
    assert torch.count_nonzero(z[1]) == 0
    # non-finite weights are sanitised as the reference does (smil_image_regressor.py:1166-1232: nan -> 0, +inf -> 0,
    # -inf -> -10 / 0), without a host-side isfinite() branch; finite entries and their gradients are untouched
    bad = sw.clone()
    bad[0, 0] = float("nan")
    bad[1, 1] = float("inf")
    tb = tw.clone()
    tb[2, 0] = float("-inf")
    ls, bt = L.pca_joint_values(bad, tb, sd, td)
    assert torch.isfinite(ls).all() and torch.isfinite(bt).all()
    assert torch.allclose(ls[2:], a[0][2:]) and torch.allclose(bt[3:], a[1][3:])
    g_in = sw.clone().requires_grad_(True)
    L.pca_joint_values(g_in, tw, sd, td)[0].sum().backward()
    g_ref = sw.clone().requires_grad_(True)
    SO.pca_weights_to_joint_values(sd, td, g_ref, tw)[0].sum().backward()
    assert torch.allclose(g_in.grad, g_ref.grad)


@pytest.mark.gpu
def test_cuda_step_matches_oracle():
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    dd = rigs.get_rig("stick")
    ref_loss, ref_fg, ref_sg = _single_process_reference(dd)
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    dev = torch.device("cuda:0")
    model = SMAL(dev)
    frame, shared, R, T, target, vis = _problem(dd, N_FRAMES, dtype=torch.float32)
    fp = {k: v.to(dev).requires_grad_(True) for k, v in frame.items()}
    sp = {k: v.to(dev).requires_grad_(True) for k, v in shared.items()}
    fn = FS.smal_joint_loss_fn(model, R.to(dev), T.to(dev), target.to(dev), vis.to(dev))
    loss = FS.ShardedFitterStep(fp, sp, fn, N_FRAMES)()
    tol = 1e-5
    assert abs(float(loss) - float(ref_loss)) <= tol * abs(float(ref_loss))
    for k in fp:
        r = ref_fg[k].float()
        assert float((fp[k].grad.cpu() - r).abs().max()) <= tol * float(r.abs().max()), k
    for k in sp:
        r = ref_sg[k].float()
        assert float((sp[k].grad.cpu() - r).abs().max()) <= tol * float(r.abs().max()), k


@pytest.mark.gpu
def test_graphed_step_equals_eager_step():
    """The CUDA-graph replay of the local part leaves the same loss and gradients as the eager step, and follows
    in-place updates of the leaves."""
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    dd = rigs.get_rig("stick")
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    dev = torch.device("cuda:0")
    model = SMAL(dev)
    n = 37
    frame, shared, R, T, target, vis = _problem(dd, n, dtype=torch.float32)
    fp = {k: v.to(dev).requires_grad_(True) for k, v in frame.items()}
    sp = {k: v.to(dev).requires_grad_(True) for k, v in shared.items()}
    fn = FS.smal_joint_loss_fn(model, R.to(dev), T.to(dev), target.to(dev), vis.to(dev), get_skin=True)
    graphed = FS.GraphedShardedFitterStep(fp, sp, fn, n)
    for trial in range(2):
        loss_g = float(graphed())
        got = {k: v.grad.clone() for k, v in list(fp.items()) + list(sp.items())}
        loss_e = float(FS.ShardedFitterStep(fp, sp, fn, n)())
        assert abs(loss_g - loss_e) <= 1e-6 * abs(loss_e)
        for k, v in list(fp.items()) + list(sp.items()):
            assert float((got[k] - v.grad).abs().max()) <= 1e-6 * float(v.grad.abs().max()), (trial, k)
        with torch.no_grad():   # an optimiser step in place: the next replay must see it
            for v in list(fp.values()) + list(sp.values()):
                v.add_(0.01 * torch.randn_like(v))
    with pytest.raises(RuntimeError):
        FS.GraphedShardedFitterStep({k: v.detach().cpu().requires_grad_(True) for k, v in fp.items()},
                                    {k: v.detach().cpu().requires_grad_(True) for k, v in sp.items()}, fn, n)


@pytest.mark.gpu
def test_graphed_full_step_with_priors_and_fused_adam_follows_torch_adam():
    """joint loss + parameter-space terms + Adam as ONE graph replay per optimisation step,
    against the eager step driven by torch.optim.Adam (optimize_to_joints.py:117-126, :153-175)."""
    import smilify_b200
    from smilify_b200.fitter_terms import FusedAdam
    from smilify_b200.smal_torch import SMAL

    dd = rigs.get_rig("stick")
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    dev = torch.device("cuda:0")
    model = SMAL(dev)
    n = 23
    frame, shared, R, T, target, vis = _problem(dd, n, dtype=torch.float32)
    J = frame["joint_rotations"].shape[1] + 1
    rs = np.random.RandomState(5)
    f = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32)).to(dev)
    consts = dict(max_limits=f(rs.uniform(0.1, 0.5, (J - 1, 3))), min_limits=f(-rs.uniform(0.1, 0.5, (J - 1, 3))),
                  betas_mean=f(np.zeros(5)), betas_prec=f(np.eye(5) + 0.1 * rs.standard_normal((5, 5))))
    weights = dict(w_limit=1.0, w_pose=0.5, w_splay=0.01, w_betas=0.3, w_temp=0.2)

    def make():
        fp = {k: v.to(dev).clone().requires_grad_(True) for k, v in frame.items()}
        sp = {k: v.to(dev).clone().requires_grad_(True) for k, v in shared.items()}
        fn = FS.with_prior_terms(FS.smal_joint_loss_fn(model, R.to(dev), T.to(dev), target.to(dev), vis.to(dev), get_skin=True),
                                 n, 0, weights, window=10, **consts)
        groups = [{"params": list(fp.values()) + [sp["betas"]], "lr": 5e-3}, {"params": [sp["fov"]], "lr": 1.0}]
        return fp, sp, fn, groups

    fp_g, sp_g, fn_g, groups_g = make()
    graphed = FS.GraphedShardedFitterStep(fp_g, sp_g, fn_g, n, optimizer=FusedAdam(groups_g, lr=5e-3))
    assert graphed._opt_in_graph
    fp_e, sp_e, fn_e, groups_e = make()
    eager = FS.ShardedFitterStep(fp_e, sp_e, fn_e, n)
    opt = torch.optim.Adam(groups_e, lr=5e-3)
    for it in range(4):
        lg = float(graphed())
        le = float(eager())
        opt.step()
        assert abs(lg - le) <= 2e-5 * abs(le), (it, lg, le)
        for k, v in list(fp_e.items()) + list(sp_e.items()):
            w = (fp_g if k in fp_g else sp_g)[k]
            assert float((w - v).abs().max()) <= 2e-5 * float(v.abs().max()), (it, k)


# ---- (f)2 host logic on the CPU: shards + halo exchange + prior terms, the oracle standing in for the kernel ----------
def _oracle_prior_fn(global_rot, joint_rot, trans, betas, weights, window, frame_offset, total_frames, halo_prev, halo_next,
                     **consts):
    """Same signature and sharding semantics as ``fitter_terms.fitter_prior_terms``, built from oracle/fitter_oracle.py."""
    import torch.nn.functional as F

    from oracle import fitter_oracle as FO

    n, N = global_rot.shape[0], int(total_frames)
    total = torch.zeros((), dtype=global_rot.dtype)
    ww = dict(weights)
    for w0 in range((frame_offset // window) * window, frame_offset + n, window):
        w1 = min(N, w0 + window)
        a, b = max(w0, frame_offset) - frame_offset, min(w1, frame_offset + n) - frame_offset
        part = dict(ww, w_betas=ww.get("w_betas", 0.0) if w0 >= frame_offset else 0.0)   # a window's betas term: once
        objs = FO.window_objective(global_rot[a:b], joint_rot[a:b], betas, part, **consts)
        frac = (b - a) / float(w1 - w0)                                                   # means run over the WHOLE window
        for k, v in objs.items():
            total = total + (v * frac if k in ("limit", "pose") else v)
    wt = ww.get("w_temp", 0.0)
    if wt > 0:
        J3 = joint_rot.shape[1] * 3
        seq = lambda h: (h[:3], h[3:3 + J3].reshape(-1, 3), h[3 + J3:])
        pair = lambda A, Bb: (F.mse_loss(A[0], Bb[0]) + F.mse_loss(A[1], Bb[1]) + F.mse_loss(A[2], Bb[2])) * wt
        fr = lambda i: (global_rot[i], joint_rot[i], trans[i])
        for i in range(n - 1):
            total = total + pair(fr(i), fr(i + 1))
        if halo_next is not None:
            total = total + pair(fr(n - 1), seq(halo_next.to(global_rot.dtype)))
        if halo_prev is not None:   # counted by the rank before: gradient only
            t = pair(seq(halo_prev.to(global_rot.dtype)), fr(0))
            total = total + (t - t.detach())
    return total


_PRIOR_W = dict(w_limit=0.7, w_pose=1.3, w_splay=0.05, w_betas=0.9, w_temp=0.4)


def _prior_consts(J, nB, dtype=torch.float64):
    rs = np.random.RandomState(9)
    t = lambda a: torch.from_numpy(np.asarray(a)).to(dtype)
    return dict(max_limits=t(rs.uniform(0.1, 0.4, (J - 1, 3))), min_limits=t(-rs.uniform(0.1, 0.4, (J - 1, 3))),
                betas_mean=t(rs.normal(0, 0.1, nB)), betas_prec=t(rs.normal(0, 1, (nB, nB))),
                pose_prior=None)


def _prior_worker(rank, world, port, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        from oracle import fitter_oracle as FO

        dd = rigs.get_rig("syn_k3")
        frame, shared, R, T, target, vis = _problem(dd, N_FRAMES)
        J, nB = frame["joint_rotations"].shape[1] + 1, shared["betas"].numel()
        consts = _prior_consts(J, nB)
        consts["pose_prior"] = FO.identity_pose_prior(J)
        lo, hi = FS.shard_range(N_FRAMES, rank, world)
        model = SO.build_model(dd, dtype=torch.float64)
        fp = {k: v[lo:hi].clone().requires_grad_(True) for k, v in frame.items()}
        sp = {k: v.clone().requires_grad_(True) for k, v in shared.items()}
        fn = FS.with_prior_terms(_oracle_loss_fn(model, R, T, target[lo:hi], vis[lo:hi]), N_FRAMES, lo, _PRIOR_W, window=4,
                                 prior_fn=_oracle_prior_fn, **consts)
        hp, hn = FS.exchange_halos(fp)
        loss = FS.ShardedFitterStep(fp, sp, fn, N_FRAMES)()
        out_q.put((rank, lo, hi, float(loss), {k: v.grad.numpy() for k, v in fp.items()}, {k: v.grad.numpy() for k, v in sp.items()},
                   None if hp is None else hp.numpy(), None if hn is None else hn.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world2_prior_terms_with_halos_match_single_process():
    from oracle import fitter_oracle as FO

    dd = rigs.get_rig("syn_k3")
    frame, shared, R, T, target, vis = _problem(dd, N_FRAMES)
    J, nB = frame["joint_rotations"].shape[1] + 1, shared["betas"].numel()
    consts = _prior_consts(J, nB)
    consts["pose_prior"] = FO.identity_pose_prior(J)
    model = SO.build_model(dd, dtype=torch.float64)
    fp = {k: v.clone().requires_grad_(True) for k, v in frame.items()}
    sp = {k: v.clone().requires_grad_(True) for k, v in shared.items()}
    joint = _oracle_loss_fn(model, R, T, target, vis)(fp, sp) / N_FRAMES
    ref = joint + FO.epoch_objective(fp["global_rotation"], fp["joint_rotations"], fp["trans"], sp["betas"], _PRIOR_W, 4, **consts)
    ref.backward()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_prior_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = lambda i: np.concatenate([frame["global_rotation"][i].numpy().ravel(), frame["joint_rotations"][i].numpy().ravel(),
                                     frame["trans"][i].numpy().ravel()]).astype(np.float32)
    for rank, lo, hi, loss, fg, sg, hp, hn in res:
        assert abs(loss - float(ref)) <= 1e-6 * abs(float(ref))
        assert (hp is None) == (lo == 0) and (hn is None) == (hi == N_FRAMES)
        if hp is not None:
            np.testing.assert_array_equal(hp, full(lo - 1))     # the LAST frame of the rank before
        if hn is not None:
            np.testing.assert_array_equal(hn, full(hi))         # the FIRST frame of the rank after
        for k in fg:
            np.testing.assert_allclose(fg[k], fp[k].grad[lo:hi].numpy(), rtol=1e-7, atol=1e-12)
        for k in sg:
            np.testing.assert_allclose(sg[k], sp[k].grad.numpy(), rtol=2e-6, atol=1e-9)
