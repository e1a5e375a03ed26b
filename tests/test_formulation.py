"""CPU: the packed/sparse/closed-form formulation (tests/emulate.py, float64) equals the dense
oracle and its autograd.  This pins the packer tables and the hand-derived backward that the
CUDA kernels transcribe."""
import numpy as np
import pytest
import torch

from oracle import smal_oracle as SO
from smilify_b200 import packer
from tests import emulate as E
from tests import rigs


def _pack(dd, static):
    m = SO.build_model(dd, dtype=torch.float64, static_joints=static)
    t = packer.pack_model(m.v_template.numpy(), m.shapedirs.numpy(), m.posedirs.numpy(), m.J_regressor.numpy(),
                          m.weights.numpy(), m.parents, m.J.numpy() if m.J is not None else None)
    return m, t


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("rig,B,propagate,nb_used", [
    ("stick", 3, False, None), ("stick", 2, True, 3), ("mouse", 2, False, None), ("syn_k3", 4, False, None),
    ("syn_static_k4", 3, True, None), ("syn_wide", 3, False, None), ("syn_chain", 3, True, None),
])
def test_fast_path_matches_oracle(rig, B, propagate, nb_used):
    dd = rigs.get_rig(rig)
    static = bool(dd.get("static_joint_locs", False))
    m, t = _pack(dd, static)
    inp = rigs.make_inputs(dd, B, seed=3, with_scale=True, with_trans_off=True, nb_used=nb_used)
    up = rigs.make_upstream(B, m.num_verts, m.num_joints, seed=9)
    rs = np.random.RandomState(0)
    g_Rs = rs.standard_normal((B, m.num_joints, 3, 3))
    tt = {k: torch.from_numpy(v.astype(np.float64)).requires_grad_(True) for k, v in inp.items()}
    o = SO.smal_forward_full(m, tt["beta"], tt["theta"], tt["trans"], None, tt["betas_logscale"], tt["betas_trans"],
                             None, propagate)
    loss = (o["verts"] * torch.from_numpy(up["g_verts"]).double()).sum() + \
           (o["joints"] * torch.from_numpy(up["g_joints"]).double()).sum() + (o["Rs"] * torch.from_numpy(g_Rs)).sum()
    loss.backward()
    a = {k: v.astype(np.float64) for k, v in inp.items()}
    fw = E.fast_forward(t, a["beta"], a["theta"], a["trans"], a["betas_logscale"], a["betas_trans"], propagate, static)
    for k in ("verts", "joints", "Rs", "v_shaped", "J_transformed"):
        assert rel(fw[k], o[k].detach().numpy()) < 5e-7, k
    g = E.fast_backward(t, a["beta"], a["theta"], a["trans"], a["betas_logscale"], a["betas_trans"], propagate, static,
                        fw, up["g_verts"].astype(np.float64), up["g_joints"].astype(np.float64), g_Rs)
    for k in ("beta", "theta", "trans", "betas_logscale", "betas_trans"):
        ref = tt[k].grad.numpy() if tt[k].grad is not None else np.zeros_like(a[k])
        assert rel(g[k], ref) < 5e-6, (k, rel(g[k], ref))
    if t.get("vm_ok"):   # the vertex-major tables (two resident bones per warp) give the same gradients
        g = E.fast_backward(t, a["beta"], a["theta"], a["trans"], a["betas_logscale"], a["betas_trans"], propagate, static,
                            fw, up["g_verts"].astype(np.float64), up["g_joints"].astype(np.float64), g_Rs, vertex_major=True)
        for k in ("beta", "theta", "trans", "betas_logscale", "betas_trans"):
            ref = tt[k].grad.numpy() if tt[k].grad is not None else np.zeros_like(a[k])
            assert rel(g[k], ref) < 5e-6, ("vertex-major", k, rel(g[k], ref))


def test_rodrigues_backward_at_zero():
    th = np.zeros((2, 3))
    G = np.random.RandomState(1).standard_normal((2, 3, 3))
    tt = torch.zeros(2, 3, dtype=torch.float64, requires_grad=True)
    (SO.rodrigues(tt) * torch.from_numpy(G)).sum().backward()
    assert rel(E.rodrigues_bwd(th, G), tt.grad.numpy()) < 1e-6


def test_tree_tables_reject_bad_parent():
    with pytest.raises(IndexError):
        packer.tree_tables(np.array([-1, 0, 3, 1], dtype=np.int32))
