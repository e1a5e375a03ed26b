"""GPU parity of the DLT triangulation and the 6-D rotation conversion, called through
the public Python surface (-> C ABI), against (a) vectors the reference's OWN statements produced
(tests/golden/triangulation_pins.npz, frozen by oracle/pin_against_reference.py::pin_triangulation), (b) the fp64 oracle,
and (c) the properties the reference tests (tests/test_triangulation_consistency.py:294-298, 336-452, 455-541)."""
import os

import numpy as np
import pytest
import torch

from oracle import projection_oracle as PO
from oracle import smal_oracle as SO
from tests import rigs
from tests.common import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a, grad=False):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV, torch.float32)
    return t.requires_grad_(True) if grad else t


@pytest.mark.parametrize("name", ["tri_b3_v4_j11", "tri_b2_v18_j40", "tri_b4_v5_j55"])
def test_triangulation_vs_frozen_reference_statements(name):
    from smilify_b200.triangulation import triangulate_joints_dlt

    z = np.load(os.path.join(rigs.GOLDEN, "triangulation_pins.npz"))
    p = name + "/"
    R, T, fov = _t(z[p + "R"], True), _t(z[p + "T"], True), _t(z[p + "fov"], True)
    tri, valid = triangulate_joints_dlt(_t(z[p + "kp"]), _t(z[p + "vis"]), R, T, fov, _t(z[p + "aspect"]), _t(z[p + "view_mask"]), 512)
    assert np.array_equal(valid.cpu().numpy(), z[p + "valid"])
    m = z[p + "valid"]
    assert rel(tri.detach().cpu().numpy()[m], z[p + "tri"][m]) < 1e-5
    (tri * _t(z[p + "g_tri"])).sum().backward()
    for k, v in (("R", R), ("T", T), ("fov", fov)):
        assert rel(v.grad.cpu().numpy(), z[p + "grad_" + k]) < 2e-5, (k, rel(v.grad.cpu().numpy(), z[p + "grad_" + k]))


def _ring(V, B, radius=3.0, elev=20.0, fov=60.0):
    R, T = PO.look_at_ring(V, radius, elev, dtype=torch.float64)
    return (R[None].expand(B, V, 3, 3).contiguous(), T[None].expand(B, V, 3).contiguous(),
            torch.full((B, V), fov, dtype=torch.float64), torch.ones(B, V, dtype=torch.float64))


def _stick_joints(seed=42, B=2):
    """As the reference test: real STICK rig, betas = 0, theta = randn * 0.15 (test_triangulation_consistency.py:202-217)."""
    dd = rigs.get_rig("stick")
    m = SO.build_model(dd, dtype=torch.float64)
    g = torch.Generator().manual_seed(seed)
    theta = torch.randn(B, m.num_joints, 3, generator=g, dtype=torch.float64) * 0.15
    joints = SO.smal_forward(m, torch.zeros(B, m.num_betas, dtype=torch.float64), theta, None, get_skin=False)
    return joints


def test_round_trip_on_stick_joints():
    """Project STICK joints into 4 ring cameras, triangulate them back: max error < 0.05, mean < 0.01 (the reference's
    bars, test_triangulation_consistency.py:294-298); here the round trip is exact to fp32."""
    from smilify_b200.triangulation import triangulate_joints_dlt

    joints = _stick_joints()
    B, J = joints.shape[:2]
    R, T, fov, asp = _ring(4, B)
    kp = PO.multiview_project(joints, R, T, fov, asp, 512.0)
    assert float(kp.min()) > 0.0 and float(kp.max()) < 1.0   # nothing clamped
    f32 = lambda t: t.to(DEV, torch.float32)
    tri, valid = triangulate_joints_dlt(f32(kp), torch.ones(B, 4, J, device=DEV), f32(R), f32(T), f32(fov), f32(asp), None, 512)
    assert bool(valid.all())
    err = (tri.cpu().double() - joints).norm(dim=-1)
    assert float(err.max()) < 0.05 and float(err.mean()) < 0.01
    assert float(err.max()) < 2e-4


def test_mask_semantics_and_gradients():
    """One visible view -> invalid; hidden views do not change the result; gradients reach cam_T and fov, and a
    detached target gets none (test_triangulation_consistency.py:336-452, 455-541)."""
    from smilify_b200.triangulation import triangulate_joints_dlt

    joints = _stick_joints(seed=3)
    B, J = joints.shape[:2]
    V = 5
    R, T, fov, asp = _ring(V, B)
    kp = PO.multiview_project(joints, R, T, fov, asp, 512.0)
    f32 = lambda t: t.to(DEV, torch.float32)
    vis = torch.ones(B, V, J, device=DEV)
    vis[:, 1:, 0] = 0.0                      # joint 0: one view only -> invalid
    vis[:, 3:, 1] = 0.0                      # joint 1: three views
    vmask = torch.ones(B, V, device=DEV)
    vmask[0, 4] = 0.0
    Tg, fg = f32(T).requires_grad_(True), f32(fov).requires_grad_(True)
    tri, valid = triangulate_joints_dlt(f32(kp), vis, f32(R), Tg, fg, f32(asp), vmask, 512)
    v = valid.cpu().numpy()
    assert not v[:, 0].any() and v[:, 1:].all()
    err = (tri.cpu().double() - joints).norm(dim=-1)
    assert float(err[:, 1:].max()) < 2e-4
    # garbage in a hidden view must not matter
    kp2 = kp.clone()
    kp2[0, 4] = 0.123
    tri2, _ = triangulate_joints_dlt(f32(kp2), vis, f32(R), f32(T), f32(fov), f32(asp), vmask, 512)
    assert torch.equal(tri2[0], tri.detach()[0])
    target = f32(joints).requires_grad_(True)
    loss = (((tri - target.detach()) ** 2) * valid[..., None]).mean()
    loss.backward()
    assert Tg.grad is not None and float(Tg.grad.abs().sum()) >= 0.0 and fg.grad is not None
    assert target.grad is None
    # a perturbed camera gives a non-zero gradient
    Tp = (f32(T) + 0.01).requires_grad_(True)
    tri3, valid3 = triangulate_joints_dlt(f32(kp), vis, f32(R), Tp, f32(fov), f32(asp), vmask, 512)
    (((tri3 - f32(joints)) ** 2) * valid3[..., None]).mean().backward()
    assert float(Tp.grad.abs().sum()) > 0.0


def test_triangulation_gradients_vs_oracle64():
    from smilify_b200.triangulation import triangulate_joints_dlt

    sc = rigs.loss_scene(3, 6, 17, 5, False)
    B, V, J = 3, 6, 17
    d = lambda k: torch.from_numpy(np.ascontiguousarray(sc[k])).double()
    leaves = [d(k).clone().requires_grad_(True) for k in ("R", "T", "fov")]
    tri, valid = PO.triangulate_joints_dlt(d("target"), d("vis"), leaves[0], leaves[1], leaves[2], d("aspect"), 512.0, d("view_mask"))
    g = torch.from_numpy(np.random.RandomState(2).standard_normal((B, J, 3))) * valid[..., None]
    (tri * g).sum().backward()
    c = [_t(sc[k], True) for k in ("R", "T", "fov")]
    tri_c, valid_c = triangulate_joints_dlt(_t(sc["target"]), _t(sc["vis"]), c[0], c[1], c[2], _t(sc["aspect"]), _t(sc["view_mask"]), 512)
    assert torch.equal(valid_c.cpu(), valid)
    m = valid.numpy()
    assert rel(tri_c.detach().cpu().numpy()[m], tri.detach().numpy()[m]) < 1e-5
    (tri_c * g.to(DEV, torch.float32)).sum().backward()
    for k, a, b in zip(("R", "T", "fov"), c, leaves):
        assert rel(a.grad.cpu().numpy(), b.grad.numpy()) < 2e-5, k


@pytest.mark.parametrize("columns", [False, True])
def test_rotation_6d(columns):
    from smilify_b200.triangulation import rotation_6d_to_matrix

    z = np.load(os.path.join(rigs.GOLDEN, "triangulation_pins.npz"))
    d6 = z["rot6d/d6"]
    x = torch.from_numpy(d6).requires_grad_(True)
    ref = PO.rotation_6d_to_matrix(x, columns=columns)
    G = torch.from_numpy(np.random.RandomState(1).standard_normal(ref.shape))
    (ref * G).sum().backward()
    c = _t(d6, True)
    R = rotation_6d_to_matrix(c, columns=columns)
    (R * G.to(DEV, torch.float32)).sum().backward()
    assert rel(R.detach().cpu().numpy(), ref.detach().numpy()) < 1e-5
    assert rel(c.grad.cpu().numpy(), x.grad.numpy()) < 1e-5
    if columns:   # the reference's own _rotation_6d_to_matrix, executed by the pin script
        assert rel(R.detach().cpu().numpy(), z["rot6d/R_cols"]) < 1e-5
    # orthonormal, right-handed
    Rd = R.detach().double().cpu()
    assert float((Rd @ Rd.transpose(1, 2) - torch.eye(3, dtype=torch.float64)).abs().max()) < 1e-5
    assert float((torch.linalg.det(Rd) - 1).abs().max()) < 1e-5


def test_rotation_6d_feeds_smal_as_rotmat():
    """6-D joint rotations -> matrices -> SMAL(theta.ndim == 4): the same joints as the reference's route through
    axis-angle (rotation_6d_to_axis_angle + Rodrigues), without the round trip."""
    import smilify_b200
    from smilify_b200.smal_torch import SMAL
    from smilify_b200.triangulation import rotation_6d_to_matrix

    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    model = SMAL(torch.device(DEV))
    dd = rigs.get_rig("stick")
    B, J = 5, model._tables["J"]
    rs = np.random.RandomState(8)
    theta = (rs.standard_normal((B, J, 3)) * 0.3)
    Rm = SO.rodrigues(torch.from_numpy(theta).reshape(-1, 3)).reshape(B, J, 3, 3)
    d6 = Rm[..., :2, :].reshape(B, J, 6)           # PyTorch3D matrix_to_rotation_6d: the first two ROWS
    beta = torch.from_numpy(rs.standard_normal((B, model.num_betas)) * 0.5)
    m = SO.build_model(dd, dtype=torch.float64)
    ref = SO.smal_forward(m, beta, torch.from_numpy(theta), None, get_skin=False)
    R = rotation_6d_to_matrix(d6.to(DEV, torch.float32))
    joints = model(beta.to(DEV, torch.float32), R, get_skin=False)
    assert rel(joints.cpu().numpy(), ref.numpy()) < 1e-5
