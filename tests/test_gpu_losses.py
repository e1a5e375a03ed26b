"""GPU parity of the reprojection / keypoint losses and of the stand-alone batch_lbs drop-ins,
called through the public Python surface (-> C ABI), against the fp64 oracles.

Tolerance: 1e-5 relative on the loss and on projected points.  Gradients of the projection + loss chain are held to
TWICE the error of the reference's OWN torch graph evaluated in fp32 against its fp64 evaluation on these very scenes
(oracle/pin_against_reference.py::pin_loss_chains -> tests/golden/PIN_REPORT.json "loss_chain_fp32_floor_worst":
grad_joints 6.6e-6, grad_T 1.0e-5, grad_R 1.7e-5, grad_fov 4.0e-5 -- the residual q - t is a ~2 px difference of ~256 px
fp32 pixel coordinates, and a shared camera sums thousands of such terms with cancellation).  north_star's 1e-5 is
therefore met where the reference itself meets it (joints) and documented where fp32 cannot (cameras); the loss
arithmetic alone (3-D loss, given points) is held to 1e-5."""
import numpy as np
import pytest
import torch

from oracle import projection_oracle as PO
from oracle import smal_oracle as SO
from tests.common import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GRAD_TOL = {"joints": 1.4e-5, "trans": 1.4e-5, "T": 2.1e-5, "R": 3.5e-5, "fov": 8.0e-5}   # 2 x measured fp32 floor


def _scene(B, V, J, seed, shared_cam=False):
    rs = np.random.RandomState(seed)
    joints = torch.from_numpy(rs.uniform(-0.4, 0.4, (B, J, 3)))
    R, T = PO.look_at_ring(V, 3.0, 15.0, dtype=torch.float64)
    if shared_cam:
        R, T = R[None], T[None]
        fov = torch.full((1, V), 60.0, dtype=torch.float64)
        asp = torch.ones(1, V, dtype=torch.float64)
    else:
        R = R[None].expand(B, V, 3, 3).clone()
        T = (T[None].expand(B, V, 3) + torch.from_numpy(rs.normal(0, 0.05, (B, V, 3)))).clone()
        fov = torch.from_numpy(rs.uniform(40, 70, (B, V)))
        asp = torch.from_numpy(rs.uniform(0.8, 1.3, (B, V)))
    return joints, R, T, fov, asp, rs


def _leaf(t, dev=None):
    t = t.detach().clone()
    if dev is not None:
        t = t.to(dev, torch.float32)
    return t.requires_grad_(True)


@pytest.mark.parametrize("shared_cam", [False, True])
@pytest.mark.parametrize("B,V,J", [(5, 3, 11), (3, 18, 40), (2, 1, 55), (64, 5, 55)])
def test_multiview_loss(B, V, J, shared_cam):
    from smilify_b200 import losses

    joints, R, T, fov, asp, rs = _scene(B, V, J, 7, shared_cam)
    S = 512.0
    with torch.no_grad():
        clean = PO.multiview_project(joints, R.expand(B, V, 3, 3), T.expand(B, V, 3), fov.expand(B, V), asp.expand(B, V), S)
    tgt = clean + torch.from_numpy(rs.normal(0, 2.0 / S, clean.shape))
    tgt[0, 0, 0] = 1.5  # out-of-bounds GT is masked out
    vis = torch.from_numpy((rs.uniform(size=(B, V, J)) < 0.8).astype(np.float64))
    vmask = torch.ones(B, V, dtype=torch.float64)
    vmask[-1, -1] = 0.0
    jw = torch.from_numpy(rs.uniform(0.5, 2.0, J))
    o = [_leaf(t) for t in (joints, R, T, fov)]
    pred = PO.multiview_project(o[0], o[1].expand(B, V, 3, 3), o[2].expand(B, V, 3), o[3].expand(B, V), asp.expand(B, V), S)
    ref = PO.multiview_keypoint_loss(pred, tgt, vis, vmask, jw)
    ref.backward()
    c = [_leaf(t, DEV) for t in (joints, R, T, fov)]
    f32 = lambda t: t.to(DEV, torch.float32)
    loss, proj = losses.multiview_keypoint_loss(c[0], c[1], c[2], c[3], f32(asp), f32(tgt), f32(vis), f32(vmask), f32(jw), S)
    loss.backward()
    assert rel(proj.cpu().numpy(), pred.detach().numpy()) < 1e-5
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    for name, a, b in zip(("joints", "R", "T", "fov"), c, o):
        assert a.grad.shape == a.shape, name
        assert rel(a.grad.cpu().numpy(), b.grad.numpy()) < GRAD_TOL[name], name


def test_multiview_loss_no_valid_joint():
    from smilify_b200 import losses

    joints, R, T, fov, asp, rs = _scene(2, 2, 5, 1)
    f32 = lambda t: t.to(DEV, torch.float32)
    j = _leaf(joints, DEV)
    loss, _ = losses.multiview_keypoint_loss(j, f32(R), f32(T), f32(fov), f32(asp), torch.full((2, 2, 5, 2), 0.5, device=DEV),
                                             torch.zeros(2, 2, 5, device=DEV), image_size=512)
    loss.backward()
    assert abs(loss.item() - 1e-8) < 1e-12 and float(j.grad.abs().max()) == 0.0


@pytest.mark.parametrize("N,J", [(10, 55), (33, 31)])
def test_fitter_joint_loss(N, J):
    from smilify_b200 import losses

    rs = np.random.RandomState(3)
    joints = torch.from_numpy(rs.uniform(-0.5, 0.5, (N, J, 3)))
    trans = torch.from_numpy(rs.normal(0, 0.1, (N, 3)))
    R, T = PO.look_at_default(dtype=torch.float64)
    S = 512.0
    fov = torch.tensor([60.0], dtype=torch.float64)
    tgt = torch.from_numpy(rs.uniform(0, S, (N, J, 2)))
    vis = torch.from_numpy((rs.uniform(size=(N, J)) < 0.7).astype(np.float64))
    oj, ot, of = _leaf(joints), _leaf(trans), _leaf(fov)
    ref, rproj = PO.fitter_joint_loss(oj, ot, R, T, of, tgt, vis, 25.0, S)
    ref.backward()
    cj, ct, cf = _leaf(joints, DEV), _leaf(trans, DEV), _leaf(fov, DEV)
    f32 = lambda t: t.to(DEV, torch.float32)
    loss, proj = losses.fitter_joint_loss(cj + ct[:, None, :], f32(R), f32(T), cf, f32(tgt), f32(vis), 25.0, S)
    loss.backward()
    assert rel(proj.cpu().numpy(), rproj.detach().numpy()) < 1e-5
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert rel(cj.grad.cpu().numpy(), oj.grad.numpy()) < GRAD_TOL["joints"]
    assert rel(ct.grad.cpu().numpy(), ot.grad.numpy()) < GRAD_TOL["trans"]
    assert cf.grad.shape == (1,) and rel(cf.grad.cpu().numpy(), of.grad.numpy()) < GRAD_TOL["fov"]


def test_singleview_and_3d_loss():
    from smilify_b200 import losses

    B, J = 9, 55
    rs = np.random.RandomState(5)
    joints = torch.from_numpy(rs.uniform(-0.5, 0.5, (B, J, 3)))
    R, T = PO.look_at_default(dtype=torch.float64)
    R = R.expand(B, 3, 3).clone()
    T = (T.expand(B, 3) + torch.from_numpy(rs.normal(0, 0.05, (B, 3)))).clone()
    fov = torch.from_numpy(rs.uniform(40, 70, (B,)))
    tgt = torch.from_numpy(rs.uniform(0.0, 1.0, (B, J, 2)))
    vis = torch.from_numpy((rs.uniform(size=(B, J)) < 0.7).astype(np.float64))
    vis[2] = 0.0  # a sample without any valid joint is skipped
    smask = torch.ones(B, dtype=torch.float64)
    smask[4] = 0.0
    jw = torch.from_numpy(rs.uniform(0.5, 2.0, J))
    oj = _leaf(joints)
    pred = PO.singleview_project(oj, R, T, fov, 512.0)
    ref = PO.singleview_keypoint_loss(pred, tgt, vis, jw, smask)
    ref.backward()
    f32 = lambda t: t.to(DEV, torch.float32)
    cj = _leaf(joints, DEV)
    loss, proj = losses.singleview_keypoint_loss(cj, f32(R), f32(T), f32(fov), f32(tgt), f32(vis), f32(jw), f32(smask), 512)
    loss.backward()
    assert rel(proj.cpu().numpy(), pred.detach().numpy()) < 1e-5
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert rel(cj.grad.cpu().numpy(), oj.grad.numpy()) < GRAD_TOL["joints"]
    # 3D keypoint loss
    t3 = torch.from_numpy(rs.uniform(-0.5, 0.5, (B, J, 3)))
    t3[1, 3] = 0.0  # all-zero GT joint is invalid
    oj = _leaf(joints)
    ref = PO.keypoint3d_loss(oj, t3, vis, jw, smask)
    ref.backward()
    cj = _leaf(joints, DEV)
    loss = losses.keypoint3d_loss(cj, f32(t3), f32(vis), f32(jw), f32(smask))
    loss.backward()
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert rel(cj.grad.cpu().numpy(), oj.grad.numpy()) < 1e-5


def test_batch_lbs_dropins():
    from smilify_b200.batch_lbs import batch_global_rigid_transformation, batch_rodrigues
    from tests import rigs

    rs = np.random.RandomState(2)
    th = torch.from_numpy(np.concatenate([rs.normal(0, 0.5, (500, 3)), np.zeros((3, 3))]))
    G = torch.from_numpy(rs.standard_normal((503, 3, 3)))
    o = _leaf(th)
    (SO.rodrigues(o) * G).sum().backward()
    c = _leaf(th, DEV)
    Rc = batch_rodrigues(c)
    (Rc * G.to(DEV, torch.float32)).sum().backward()
    assert rel(Rc.detach().cpu().numpy(), SO.rodrigues(th).numpy()) < 1e-5
    assert rel(c.grad.cpu().numpy(), o.grad.numpy()) < 1e-5
    # chain over the STICK tree
    dd = rigs.get_rig("stick")
    parents = np.asarray(dd["kintree_table"])[0].astype(np.int32)
    J, B = len(parents), 6
    Rs = SO.rodrigues(torch.from_numpy(rs.normal(0, 0.3, (B * J, 3)))).reshape(B, J, 3, 3)
    Js = torch.from_numpy(rs.uniform(-1, 1, (B, J, 3)))
    ls = torch.from_numpy(rs.normal(0, 0.05, (B, J, 3)))
    bt = torch.from_numpy(rs.normal(0, 0.01, (B, J, 3)))
    gJ = torch.from_numpy(rs.standard_normal((B, J, 3)))
    gA = torch.from_numpy(rs.standard_normal((B, J, 4, 4)))
    for prop in (False, True):
        o = [_leaf(t) for t in (Rs, Js, ls, bt)]
        nJ, A = SO.global_rigid_transformation(o[0], o[1], parents, o[2], o[3], prop)
        ((nJ * gJ).sum() + (A * gA).sum()).backward()
        c = [_leaf(t, DEV) for t in (Rs, Js, ls, bt)]
        nJc, Ac = batch_global_rigid_transformation(c[0], c[1], parents, betas_logscale=c[2], betas_trans=c[3],
                                                    propagate_scaling=prop, num_joints=J)
        ((nJc * gJ.to(DEV, torch.float32)).sum() + (Ac * gA.to(DEV, torch.float32)).sum()).backward()
        assert Ac.shape == (B, J, 4, 4)
        assert rel(nJc.detach().cpu().numpy(), nJ.detach().numpy()) < 1e-5
        assert rel(Ac.detach().cpu().numpy(), A.detach().numpy()) < 1e-5
        for name, a, b in zip(("Rs", "Js", "ls", "bt"), c, o):
            assert rel(a.grad.cpu().numpy(), b.grad.numpy()) < 1e-5, (prop, name)


# ---- against values the REFERENCE'S OWN loss code produced (frozen by oracle/pin_against_reference.py) ----------
def _t(a, grad=False):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV, torch.float32)
    return t.requires_grad_(True) if grad else t


@pytest.mark.parametrize("name", list(__import__("tests.rigs", fromlist=["LOSS_SCENES"]).LOSS_SCENES))
def test_multiview_loss_vs_frozen_reference_code(name):
    """CUDA projection + loss against `_compute_batched_keypoint_loss` of the reference executed on the oracle's
    projected points (tests/golden/loss_chain_pins.npz), gradients chained through the projection in fp64."""
    import os

    from smilify_b200 import losses
    from tests import rigs

    z = np.load(os.path.join(rigs.GOLDEN, "loss_chain_pins.npz"))
    B, V, J, seed, shared = rigs.LOSS_SCENES[name]
    sc = rigs.loss_scene(B, V, J, seed, shared)
    c = {k: _t(sc[k], grad=True) for k in ("joints", "R", "T", "fov")}
    loss, proj = losses.multiview_keypoint_loss(c["joints"], c["R"], c["T"], c["fov"], _t(sc["aspect"]), _t(sc["target"]),
                                                _t(sc["vis"]), _t(sc["view_mask"]), _t(sc["joint_weights"]),
                                                float(sc["image_size"]))
    loss.backward()
    assert rel(proj.detach().cpu().numpy(), z[name + "/pred"]) < 1e-5
    assert abs(loss.item() - float(z[name + "/loss"])) <= 1e-5 * abs(float(z[name + "/loss"]))
    for k, v in c.items():
        assert v.grad.shape == v.shape, k
        assert rel(v.grad.cpu().numpy(), z[name + "/grad_" + k]) < GRAD_TOL[k], (k, rel(v.grad.cpu().numpy(), z[name + "/grad_" + k]))


def _pins(prefix):
    import os

    from tests import rigs

    z = np.load(os.path.join(rigs.GOLDEN, "loss_pins.npz"))
    return z, sorted({k.split("/")[1] for k in z.files if k.startswith(prefix + "/")})


def _joints_for_points(pred_yx, S):
    """World joints that the camera R = I, T = 0, fov = 90 degrees, aspect 1 projects onto the given (y,x) pixels:
    screen = S/2 (1 - X/Z) at Z = 1."""
    p = np.asarray(pred_yx, dtype=np.float64)
    return np.stack([1.0 - 2.0 * p[..., 1] / S, 1.0 - 2.0 * p[..., 0] / S, np.ones(p.shape[:-1])], axis=-1)


def test_keypoint3d_loss_vs_frozen_reference_code():
    from smilify_b200 import losses

    z, names = _pins("singleview")
    for n in names:
        p = "singleview/%s/" % n
        j3 = _t(z[p + "j3"], grad=True)
        jw = _t(z[p + "joint_weights"]) if p + "joint_weights" in z.files else None
        sm = _t(z[p + "sample_mask"].astype(np.float32)) if p + "sample_mask" in z.files else None
        loss = losses.keypoint3d_loss(j3, _t(z[p + "t3"]), _t(z[p + "vis"]), jw, sm)
        loss.backward()
        ref, gref = float(z[p + "loss3"]), z[p + "grad_j3"]
        if np.isnan(ref):   # the reference's inf * 0 -> NaN for a non-finite predicted joint
            assert np.isnan(loss.item()), n
            g = j3.grad.cpu().numpy()
            assert np.array_equal(np.isnan(g), np.isnan(gref)), n
            continue
        assert abs(loss.item() - ref) <= 1e-5 * abs(ref), n
        assert rel(j3.grad.cpu().numpy(), gref) < 1e-5, n


def test_singleview_loss_vs_frozen_reference_code():
    """`_compute_batch_keypoint_loss` of the reference on given 2-D points; the CUDA path gets joints that a
    trivial camera maps onto exactly those points (so only the loss arithmetic is compared)."""
    from smilify_b200 import losses

    z, names = _pins("singleview")
    S = 512.0
    for n in names:
        if n.endswith("nonfinite"):
            continue   # non-finite 2-D points cannot be produced through a finite camera; covered on the CPU side
        p = "singleview/%s/" % n
        pred = z[p + "pred"].astype(np.float64)
        B = pred.shape[0]
        joints = _t(_joints_for_points(pred * (S + 1e-8), S), grad=True)
        R = torch.eye(3, device=DEV).expand(B, 3, 3)
        T = torch.zeros(B, 3, device=DEV)
        fov = torch.full((B,), 90.0, device=DEV)
        jw = _t(z[p + "joint_weights"]) if p + "joint_weights" in z.files else None
        sm = _t(z[p + "sample_mask"].astype(np.float32)) if p + "sample_mask" in z.files else None
        loss, proj = losses.singleview_keypoint_loss(joints, R, T, fov, _t(z[p + "target"]), _t(z[p + "vis"]), jw, sm, S)
        assert np.abs(proj.detach().cpu().numpy() - pred).max() < 2e-6, n
        ref = float(z[p + "loss"])
        assert abs(loss.item() - ref) <= 2e-5 * abs(ref), (n, loss.item(), ref)   # points reproduced to ~1e-6 only


def test_fitter_objective_vs_frozen_reference_code():
    """fitter.py:292-296 (-1 fill on both sides + F.mse_loss) executed by the reference on given pixels."""
    from smilify_b200 import losses

    z, names = _pins("fitter")
    S = 512.0
    for n in names:
        p = "fitter/%s/" % n
        pred = z[p + "pred"].astype(np.float64)
        N = pred.shape[0]
        joints = _t(_joints_for_points(pred, S), grad=True)
        loss, proj = losses.fitter_joint_loss(joints, torch.eye(3, device=DEV)[None], torch.zeros(1, 3, device=DEV),
                                              torch.tensor([90.0], device=DEV), _t(z[p + "target"]), _t(z[p + "vis"]),
                                              float(z[p + "w_j2d"]), S)
        ref = float(z[p + "loss"])
        vis = z[p + "vis"] > 0
        if vis.any():
            assert np.abs(proj.detach().cpu().numpy()[vis] - pred[vis]).max() < 5e-4, n   # pixels, fp32 camera round trip
        assert abs(loss.item() - ref) <= 2e-5 * max(abs(ref), 1e-30), (n, loss.item(), ref)


def test_pca_joint_values_vs_frozen_reference_code():
    from smilify_b200 import losses

    z, names = _pins("pca")
    for n in names:
        p = "pca/%s/" % n
        sw = _t(z[p + "scale_w"], grad=True)
        tw = _t(z[p + "trans_w"], grad=True) if z[p + "trans_w"].size else None
        ls, bt = losses.pca_joint_values(sw, tw, _t(z[p + "scaledirs"]), _t(z[p + "transdirs"]))
        ((ls * _t(z[p + "g_ls"])).sum() + ((bt * _t(z[p + "g_bt"])).sum() if tw is not None else 0.0)).backward()
        assert rel(ls.detach().cpu().numpy(), z[p + "log_scales"]) < 1e-5, n
        assert rel(bt.detach().cpu().numpy(), z[p + "betas_trans"]) < 1e-5, n
        assert rel(sw.grad.cpu().numpy(), z[p + "grad_scale_w"]) < 1e-5, n
        if tw is not None:
            assert rel(tw.grad.cpu().numpy(), z[p + "grad_trans_w"]) < 1e-5, n
