"""The drop-in boundary as the reference's callers use it:
module shadowing, attribute surface, ``no_grad`` inference, autocast, ``get_skin=False``, the
``J_transformed`` side effect, error behaviour."""
import sys

import numpy as np
import pytest
import torch

from tests import rigs
from tests.common import TOL, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(rig="stick", static=False):
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    cfg = smilify_b200.configure(rigs.REAL_RIGS[rig])
    cfg.ALLOW_LIMB_SCALING = True
    cfg.ignore_hardcoded_body = True
    cfg.STATIC_JOINT_LOCATIONS = static
    return SMAL(torch.device(DEV))


def _inputs(rig, B, seed=3):
    dd = rigs.get_rig(rig)
    inp = rigs.make_inputs(dd, B, seed=seed)
    return {k: torch.from_numpy(inp[k]).to(DEV) for k in ("beta", "theta", "trans")}


def test_install_shadows_the_reference_modules():
    import smilify_b200

    saved = {k: sys.modules.get(k) for k in ("smal_model", "smal_model.smal_torch", "smal_model.batch_lbs")}
    try:
        cls = smilify_b200.install(force=True)
        from smal_model.batch_lbs import batch_global_rigid_transformation, batch_rodrigues  # noqa: F401
        from smal_model.smal_torch import SMAL   # what fitter.py:15 / trainer.py:21 execute

        assert SMAL is cls and SMAL.__module__ == "smilify_b200.smal_torch"
        R = batch_rodrigues(torch.zeros(4, 3, device=DEV))
        assert torch.allclose(R, torch.eye(3, device=DEV).expand(4, 3, 3), atol=1e-7)   # fitter init pose
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_attribute_surface_and_plain_attributes():
    m = _model()
    V, J = 3020, 55
    assert m.faces.dtype == torch.int64 and m.faces.shape[1] == 3          # fitter.py:287, trainer.py:131
    assert tuple(m.v_template.shape) == (V, 3) and m.size == [V, 3]       # trainer.py:140
    assert m.num_betas == m.shapedirs.shape[0] and tuple(m.shapedirs.shape) == (m.num_betas, 3 * V)
    assert tuple(m.J_regressor.shape) == (V, J) and tuple(m.weights.shape) == (V, J)
    assert m.parents.dtype == np.int32 and m.parents.shape == (J,)
    assert tuple(m.posedirs.shape) == (9 * (J - 1), 3 * V)
    assert len(m.state_dict()) == 0 and len(list(m.parameters())) == 0    # nothing registered: checkpoints / DDP unaffected


def test_no_grad_inference_and_get_skin_false():
    m = _model()
    t = _inputs("stick", 9)
    with torch.no_grad():                                                  # fitter.py:422-429
        verts, joints, Rs, vs = m(t["beta"], t["theta"], trans=t["trans"])
        j_only = m(t["beta"], t["theta"], trans=t["trans"], get_skin=False)
    assert not verts.requires_grad and not joints.requires_grad
    assert tuple(verts.shape) == (9, 3020, 3) and tuple(Rs.shape) == (9, 55, 3, 3) and tuple(vs.shape) == (9, 3020, 3)
    assert torch.equal(j_only, joints)                                     # smal_torch.py:369-370
    g = {k: v.clone().requires_grad_(True) for k, v in t.items()}
    v2, j2, _, _ = m(g["beta"], g["theta"], trans=g["trans"])
    assert torch.equal(v2, verts) and torch.equal(j2, joints)
    assert m.J_transformed is not None and tuple(m.J_transformed.shape) == (9, 55, 3)   # smal_torch.py:307


def test_static_joints_skip_trans_on_joints():
    m = _model("mouse", static=True)
    t = _inputs("mouse", 5)
    verts, joints, _, _ = m(t["beta"], t["theta"], trans=t["trans"])
    assert joints is m.J_transformed                                       # smal_torch.py:343-346
    v0, j0, _, _ = m(t["beta"], t["theta"], trans=None)
    assert torch.equal(joints, j0)                                         # static joints carry no trans
    assert rel((verts - v0).cpu().numpy(), t["trans"][:, None].expand_as(verts).cpu().numpy()) < TOL


def test_autocast_and_half_inputs_compute_in_fp32():
    m = _model()
    t = _inputs("stick", 6)
    ref = m(t["beta"], t["theta"], trans=t["trans"])
    with torch.autocast("cuda", dtype=torch.bfloat16):                     # train_multiview_regressor.py:610
        out = m(t["beta"], t["theta"], trans=t["trans"])
    for a, b in zip(out, ref):
        assert a.dtype == torch.float32 and torch.equal(a, b)
    with torch.autocast("cuda", dtype=torch.float16):
        out16 = m(t["beta"].half(), t["theta"].half(), trans=t["trans"].half())
    ref16 = m(t["beta"].half().float(), t["theta"].half().float(), trans=t["trans"].half().float())
    assert out16[0].dtype == torch.float32 and torch.equal(out16[0], ref16[0])


def test_errors_are_python_exceptions_not_aborts():
    m = _model()
    t = _inputs("stick", 3)
    with pytest.raises((RuntimeError, ValueError)):
        m(t["beta"].cpu(), t["theta"].cpu(), trans=t["trans"].cpu())       # no CPU fallback
    with pytest.raises((RuntimeError, ValueError)):
        m(torch.zeros(3, m.num_betas + 2, device=DEV), t["theta"], trans=t["trans"])   # more betas than shapedirs
    verts, _, _, _ = m(t["beta"], t["theta"], trans=t["trans"])           # the model still works afterwards
    assert torch.isfinite(verts).all()
    wrong = m(t["beta"], torch.zeros(3, 7, 3, device=DEV), trans=t["trans"])   # wrong-width pose -> zeros (smal_torch.py:281-286)
    zero = m(t["beta"], torch.zeros(3, 55, 3, device=DEV), trans=t["trans"])
    assert torch.equal(wrong[0], zero[0])


def test_v_template_override_is_differentiable():
    """The reference graph differentiates through ``v_template`` (smal_torch.py:223-232: v_shaped = v_template + blend);
    the CUDA path returns the batch sum of dL/dv_shaped."""
    import smilify_b200
    from oracle import smal_oracle as SO
    from smilify_b200.smal_torch import SMAL
    from tests import rigs
    from tests.common import TOL, rel

    dd = rigs.get_rig("stick")
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    dev = torch.device("cuda:0")
    model = SMAL(dev)
    B = 3
    inp = rigs.make_inputs(dd, B, seed=4)
    rs = np.random.RandomState(2)
    V = np.asarray(dd["v_template"]).shape[0]
    vt = (np.asarray(dd["v_template"]) + 0.01 * rs.standard_normal((V, 3))).astype(np.float32)
    gv = rs.standard_normal((B, V, 3)).astype(np.float32)
    t = {k: torch.from_numpy(inp[k]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans")}
    vt_c = torch.from_numpy(vt).to(dev).requires_grad_(True)
    verts, joints, _, _ = model(t["beta"], t["theta"], trans=t["trans"], v_template=vt_c)
    (verts * torch.from_numpy(gv).to(dev)).sum().backward()
    m = SO.build_model(dd, dtype=torch.float64)
    o = {k: torch.from_numpy(inp[k]).double().requires_grad_(True) for k in ("beta", "theta", "trans")}
    vt_o = torch.from_numpy(vt).double().requires_grad_(True)
    out = SO.smal_forward_full(m, o["beta"], o["theta"], o["trans"], None, None, None, vt_o, False, True, True)
    (out["verts"] * torch.from_numpy(gv).double()).sum().backward()
    assert rel(vt_c.grad.cpu().numpy(), vt_o.grad.numpy()) <= TOL
    for k in t:
        assert rel(t[k].grad.cpu().numpy(), o[k].grad.numpy()) <= TOL, k


def test_smal_from_flat_rig_file_equals_smal_from_pickle(tmp_path):
    """``config.SMAL_FILE`` pointing at the flat pickle-free export gives the same model."""
    import smilify_b200
    from smilify_b200 import model_io
    from smilify_b200.smal_torch import SMAL
    from tests import rigs

    dev = torch.device("cuda:0")
    dd = rigs.get_rig("stick")
    inp = rigs.make_inputs(dd, 5, seed=9)
    t = {k: torch.from_numpy(inp[k]).to(dev) for k in ("beta", "theta", "trans")}
    outs = []
    for path in (rigs.REAL_RIGS["stick"], model_io.export_flat(rigs.REAL_RIGS["stick"], str(tmp_path / "stick_flat.npz"))):
        smilify_b200.configure(path)
        model = SMAL(dev)
        outs.append(model(t["beta"], t["theta"], trans=t["trans"]))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
