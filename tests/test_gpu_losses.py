"""GPU parity of the reprojection / keypoint losses and of the stand-alone batch_lbs drop-ins,
called through the public Python surface (-> C ABI), against the fp64 oracles.

Tolerance: 1e-5 relative on the loss and on projected points; 1e-4 relative (max-norm) on loss
gradients, whose fp32 value is a difference of O(1e2)-pixel terms."""
import numpy as np
import pytest
import torch

from oracle import projection_oracle as PO
from oracle import smal_oracle as SO
from tests.common import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


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
        assert rel(a.grad.cpu().numpy(), b.grad.numpy()) < 1e-4, name


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
    assert rel(cj.grad.cpu().numpy(), oj.grad.numpy()) < 1e-4
    assert rel(ct.grad.cpu().numpy(), ot.grad.numpy()) < 1e-4
    assert cf.grad.shape == (1,) and rel(cf.grad.cpu().numpy(), of.grad.numpy()) < 1e-4


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
    assert rel(cj.grad.cpu().numpy(), oj.grad.numpy()) < 1e-4
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
