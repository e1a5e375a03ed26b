"""Parity at BASELINE.json's full sizes through size-independent properties.

* configs[1]  STICK, B = 4096: every sample of a batch is independent (smal_torch.py:223-365 has no
  cross-sample term), so a random subset of the big batch must equal the fp64 oracle run on that
  subset alone (1e-5 relative), and the big batch must equal the CUDA path run on the subset alone
  bit for bit (forward) -- batch position must not leak into a sample's result.
* configs[2]  mouse rig, 18-camera reprojection loss: the loss is a ratio of sums over frames
  (multiview_smil_regressor.py:2319-2388), so a batch built from k copies of a smaller batch has the SAME
  loss and 1/k of its gradients; the small batch is checked against the oracle.
* configs[4]  STICK, 5 cameras, propagate_scaling=True, PCA scale / translation weights
  (smil_image_regressor.py:1166-1232) feeding betas_logscale / betas_trans, same two properties.
"""
import numpy as np
import pytest
import torch

from oracle import projection_oracle as PO
from oracle import smal_oracle as SO
from tests import rigs
from tests.common import TOL, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cuda_model(rig, static=False):
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    cfg = smilify_b200.configure(rigs.REAL_RIGS[rig])
    cfg.ALLOW_LIMB_SCALING = True
    cfg.ignore_hardcoded_body = True
    cfg.STATIC_JOINT_LOCATIONS = static
    return SMAL(torch.device(DEV))


def _run_cuda(model, inp, gv, gj):
    t = {k: torch.from_numpy(v).to(DEV).requires_grad_(True) for k, v in inp.items()}
    verts, joints, Rs, vs = model(t["beta"], t["theta"], trans=t["trans"])
    gr = torch.autograd.grad([verts, joints], [t["beta"], t["theta"], t["trans"]],
                             [torch.from_numpy(gv).to(DEV), torch.from_numpy(gj).to(DEV)])
    torch.cuda.synchronize()
    out = dict(verts=verts, joints=joints, Rs=Rs, v_shaped=vs, g_beta=gr[0], g_theta=gr[1], g_trans=gr[2])
    return {k: v.detach().cpu().numpy() for k, v in out.items()}


def test_stick_b4096_subset_matches_oracle_and_small_batch():
    B = 4096
    dd = rigs.get_rig("stick")
    model = _cuda_model("stick")
    V, J = model._tables["V"], model._tables["J"]
    inp = rigs.make_inputs(dd, B, seed=11)
    inp = {k: inp[k] for k in ("beta", "theta", "trans")}
    rs = np.random.RandomState(5)
    gv = rs.standard_normal((B, V, 3)).astype(np.float32)
    gj = rs.standard_normal((B, J, 3)).astype(np.float32)
    big = _run_cuda(model, inp, gv, gj)
    for k, v in big.items():
        assert np.isfinite(v).all(), k
    # first / last samples, both halves of sample pairs, group boundaries of the persistent passes, random ones
    sel = np.unique(np.concatenate([[0, 1, 2, 7, 8, 9, 31, 32, 33, B - 2, B - 1], rs.choice(B, 21, replace=False)]))
    sub_inp = {k: np.ascontiguousarray(v[sel]) for k, v in inp.items()}
    # (a) fp64 oracle on the subset
    m = SO.build_model(dd, dtype=torch.float64)
    t = {k: torch.from_numpy(v).double().requires_grad_(True) for k, v in sub_inp.items()}
    verts, joints, Rs, vs = SO.smal_forward(m, t["beta"], t["theta"], t["trans"])
    ((verts * torch.from_numpy(gv[sel]).double()).sum() + (joints * torch.from_numpy(gj[sel]).double()).sum()).backward()
    ref = dict(verts=verts, joints=joints, Rs=Rs, v_shaped=vs)
    for k, v in ref.items():
        assert rel(big[k][sel], v.detach().numpy()) <= TOL, k
    for k in ("beta", "theta", "trans"):
        assert rel(big["g_" + k][sel], t[k].grad.numpy()) <= TOL, k
    # (b) the same subset through the CUDA path alone: position in the batch must not matter
    small = _run_cuda(model, sub_inp, np.ascontiguousarray(gv[sel]), np.ascontiguousarray(gj[sel]))
    for k in ("verts", "joints", "Rs", "v_shaped"):
        assert np.array_equal(big[k][sel], small[k]), k
    for k in ("g_beta", "g_theta", "g_trans"):
        assert rel(big[k][sel], small[k]) <= 1e-6, k


def _world_joints(joints, trans, scale=10.0):
    """``_predict_canonical_joints_3d`` placement (multiview_smil_regressor.py:1522-1604): root-centred,
    scaled, translated -- plain tensor algebra on both sides."""
    root = joints[:, :1]
    return (joints - root) * scale * 0.1 + trans[:, None]


def _multiview_problem(rig, static, B, nviews, seed, with_pca):
    dd = rigs.get_rig(rig)
    rs = np.random.RandomState(seed)
    J = np.asarray(dd["weights"]).shape[1]
    nB = np.asarray(dd["shapedirs"]).shape[-1]
    p = dict(beta=rs.normal(0, 0.5, (B, nB)), theta=rs.normal(0, 0.3, (B, J, 3)), trans=rs.normal(0, 0.05, (B, 3)))
    if with_pca:
        p["scale_w"] = rs.normal(0, 0.5, (B, nB))
        p["trans_w"] = rs.normal(0, 0.5, (B, nB))
    R, T = PO.look_at_ring(nviews, 3.0, 15.0, dtype=torch.float64)
    cam = dict(R=R[None].expand(B, nviews, 3, 3).clone(), T=T[None].expand(B, nviews, 3).clone(),
               fov=torch.full((B, nviews), 60.0, dtype=torch.float64), asp=torch.ones(B, nviews, dtype=torch.float64))
    tgt = torch.from_numpy(rs.uniform(0.2, 0.8, (B, nviews, J, 2)))
    vis = torch.from_numpy((rs.uniform(size=(B, nviews, J)) < 0.8).astype(np.float64))
    return dd, p, cam, tgt, vis


def _oracle_multiview(dd, static, p, cam, tgt, vis, propagate, with_pca):
    m = SO.build_model(dd, dtype=torch.float64, static_joints=static)
    t = {k: torch.from_numpy(np.asarray(v)).double().requires_grad_(True) for k, v in p.items()}
    ls = bt = None
    if with_pca:
        sd = torch.from_numpy(np.asarray(dd["scaledirs"], dtype=np.float64)).permute(1, 2, 0)   # (C,J,3) -> (J,3,C)
        td = torch.from_numpy(np.asarray(dd["transdirs"], dtype=np.float64)).permute(1, 2, 0)
        ls, bt = SO.pca_weights_to_joint_values(sd, td, t["scale_w"], t["trans_w"])
    joints = SO.smal_forward(m, t["beta"], t["theta"], None, betas_logscale=ls, betas_trans=bt, get_skin=False,
                             propagate_scaling=propagate)
    pred = PO.multiview_project(_world_joints(joints, t["trans"]), cam["R"], cam["T"], cam["fov"], cam["asp"], 512.0)
    loss = PO.multiview_keypoint_loss(pred, tgt, vis)
    loss.backward()
    return float(loss.detach()), {k: v.grad.numpy() for k, v in t.items() if v.grad is not None}


def _cuda_multiview(model, dd, p, cam, tgt, vis, propagate, with_pca, reps=1):
    from smilify_b200 import losses

    f32 = lambda a: torch.from_numpy(np.asarray(a)).float().repeat((reps,) + (1,) * (np.asarray(a).ndim - 1)).to(DEV)
    t = {k: f32(v).requires_grad_(True) for k, v in p.items()}
    ls = bt = None
    if with_pca:
        sd = torch.from_numpy(np.asarray(dd["scaledirs"], dtype=np.float32)).permute(1, 2, 0).contiguous().to(DEV)
        td = torch.from_numpy(np.asarray(dd["transdirs"], dtype=np.float32)).permute(1, 2, 0).contiguous().to(DEV)
        ls, bt = losses.pca_joint_values(t["scale_w"], t["trans_w"], sd, td)
    joints = model(t["beta"], t["theta"], betas_logscale=ls, betas_trans=bt, get_skin=False, propagate_scaling=propagate)
    c = {k: f32(v.numpy()) for k, v in cam.items()}
    loss, _ = losses.multiview_keypoint_loss(_world_joints(joints, t["trans"]), c["R"], c["T"], c["fov"], c["asp"],
                                             f32(tgt.numpy()), f32(vis.numpy()), None, None, 512.0)
    loss.backward()
    torch.cuda.synchronize()
    return float(loss.detach()), {k: v.grad.cpu().numpy() for k, v in t.items() if v.grad is not None}


@pytest.mark.parametrize("rig,static,nviews,Bs,reps,propagate,with_pca", [
    ("mouse", True, 18, 128, 8, False, False),     # configs[2]: 1 024 frames per GPU x 18 cameras
    ("stick", False, 5, 128, 8, True, True),       # configs[4]: batch 1 024, 5 cameras, PCA scale / translation
], ids=["mouse18cam", "stick5cam_pca"])
def test_multiview_training_loss_fullsize(rig, static, nviews, Bs, reps, propagate, with_pca):
    dd, p, cam, tgt, vis = _multiview_problem(rig, static, Bs, nviews, 23, with_pca)
    model = _cuda_model(rig, static)
    ref_loss, ref_g = _oracle_multiview(dd, static, p, cam, tgt, vis, propagate, with_pca)
    loss, g = _cuda_multiview(model, dd, p, cam, tgt, vis, propagate, with_pca)
    assert abs(loss - ref_loss) <= TOL * abs(ref_loss)
    assert set(g) == set(ref_g)   # a static-joint rig has no beta gradient through its joints (smal_torch.py:257-260)
    for k, v in ref_g.items():
        assert rel(g[k], v) <= 2e-5, k      # per-frame parameters through SMAL + projection + loss (see test_gpu_losses)
    # k copies of the batch: same loss, 1/k of every per-frame gradient
    loss_k, g_k = _cuda_multiview(model, dd, p, cam, tgt, vis, propagate, with_pca, reps=reps)
    assert abs(loss_k - loss) <= 2e-6 * abs(loss)
    for k, v in g.items():
        for r in range(reps):
            assert rel(g_k[k][r * Bs:(r + 1) * Bs] * reps, v) <= 2e-5, (k, r)
