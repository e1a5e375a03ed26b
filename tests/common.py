"""Shared drivers for the parity tests: run one case of tests/rigs.CASES through the CUDA
path (smilify_b200) or through the oracle, with identical seeded inputs."""
from __future__ import annotations

import os
import tempfile

import numpy as np
import torch

from tests import rigs

TOL = 1e-5  # north_star: 1e-5 relative on fp32 vertices, joints and gradients
_RIG_FILES = {}


def rel(a, b):
    """max|a-b| / max|b| -- the relative error the 1e-5 bar is stated on."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.shape != b.shape:
        return float("inf")
    if b.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def rig_file(name):
    """Path of an on-disk rig the product loader can read (synthetic rigs are written to a temp .npz)."""
    if name in rigs.REAL_RIGS:
        return rigs.REAL_RIGS[name]
    if name not in _RIG_FILES:
        dd = rigs.get_rig(name)
        path = os.path.join(tempfile.gettempdir(), "smil_rig_%s_%d.npz" % (name, os.getpid()))
        save = {}
        for k, v in dd.items():
            if k == "J_names":
                save[k] = np.array([str(s) for s in v])
            elif k == "static_joint_locs":
                save[k] = np.array(bool(v))
            else:
                a = np.asarray(v)
                save[k] = a.astype(np.float32) if a.dtype.kind == "f" else a
        np.savez(path, **save)
        _RIG_FILES[name] = path
    return _RIG_FILES[name]


def run_oracle_case(case, dtype=torch.float64, want_grads=True):
    from oracle import smal_oracle as SO

    dd = rigs.get_rig(case["rig"])
    inp, call, cfg, up = rigs.case_inputs(case, dd)
    static = cfg.get("STATIC_JOINT_LOCATIONS", bool(dd.get("static_joint_locs", False)))
    m = SO.build_model(dd, dtype=dtype, static_joints=static)
    t = {k: torch.from_numpy(v.copy()).to(dtype).requires_grad_(want_grads) for k, v in inp.items()}
    kw = dict(call)
    get_skin = kw.pop("get_skin", True)
    vt = torch.from_numpy(kw["v_template"]).to(dtype) if "v_template" in kw else None
    o = SO.smal_forward_full(m, t["beta"], t["theta"], t.get("trans"), t.get("del_v"), t.get("betas_logscale"),
                             t.get("betas_trans"), vt, kw.get("propagate_scaling", False),
                             cfg.get("ALLOW_LIMB_SCALING", True), cfg.get("ignore_hardcoded_body", True))
    keys = ("verts", "joints", "Rs", "v_shaped") if get_skin else ("joints",)
    grads = {}
    if want_grads:
        loss = 0.0
        for k in keys:
            g = up.get("g_" + k)
            if g is not None and o[k].requires_grad:
                loss = loss + (o[k] * torch.from_numpy(g).to(dtype)).sum()
        if isinstance(loss, torch.Tensor):
            loss.backward()
            grads = {k: v.grad.numpy().copy() for k, v in t.items() if v.grad is not None}
    outs = {k: o[k].detach().numpy().copy() for k in keys + ("J_transformed",)}
    return outs, grads


def make_cuda_model(case, device="cuda:0", force_generic=False):
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    dd = rigs.get_rig(case["rig"])
    cfgmod = smilify_b200.configure(rig_file(case["rig"]))
    cfgmod.ALLOW_LIMB_SCALING = True
    cfgmod.ignore_hardcoded_body = True
    for k, v in case["cfg"].items():
        setattr(cfgmod, k, v)
    model = SMAL(torch.device(device))
    model._force_generic = force_generic
    return model, dd


def run_cuda_case(case, device="cuda:0", force_generic=False, want_grads=True):
    model, dd = make_cuda_model(case, device, force_generic)
    inp, call, cfg, up = rigs.case_inputs(case, dd)
    dev = torch.device(device)
    t = {k: torch.from_numpy(v.copy()).to(dev).requires_grad_(want_grads) for k, v in inp.items()}
    kw = dict(call)
    if "v_template" in kw:
        kw["v_template"] = torch.from_numpy(kw["v_template"]).to(dev)
    res = model(t["beta"], t["theta"], trans=t.get("trans"), del_v=t.get("del_v"),
                betas_logscale=t.get("betas_logscale"), betas_trans=t.get("betas_trans"), **kw)
    if kw.get("get_skin", True) is False:
        outs = {"joints": res}
    else:
        outs = dict(zip(("verts", "joints", "Rs", "v_shaped"), res))
    grads = {}
    if want_grads:
        loss = 0.0
        for k in ("verts", "joints", "Rs", "v_shaped"):
            g = up.get("g_" + k)
            if g is not None and k in outs and outs[k].requires_grad:
                loss = loss + (outs[k] * torch.from_numpy(g).to(dev)).sum()
        if isinstance(loss, torch.Tensor):
            loss.backward()
            grads = {k: v.grad.detach().cpu().numpy().copy() for k, v in t.items() if v.grad is not None}
    outs["J_transformed"] = model.J_transformed
    torch.cuda.synchronize()
    return {k: v.detach().cpu().numpy().copy() for k, v in outs.items()}, grads


def compare_case(case, device="cuda:0", force_generic=False):
    """-> list of (tensor name, rel err vs oracle64, rel err vs frozen reference fp32 or None)."""
    o_out, o_grad = run_oracle_case(case)
    c_out, c_grad = run_cuda_case(case, device, force_generic)
    gold = np.load(os.path.join(rigs.GOLDEN, "case_%s.npz" % case["name"]))
    s = case["stride"]
    rows = []
    for k, v in o_out.items():
        g = None
        if "out_" + k in gold.files:
            cv = c_out[k][:, ::s] if k in ("verts", "v_shaped") else c_out[k]
            g = rel(cv, gold["out_" + k])
        rows.append(("out." + k, rel(c_out.get(k, np.zeros(0)), v), g))
    assert set(c_grad) == set(o_grad), (case["name"], sorted(c_grad), sorted(o_grad))
    for k, v in o_grad.items():
        g = None
        if "grad_" + k in gold.files:
            cv = c_grad[k][:, ::s] if k == "del_v" else c_grad[k]
            g = rel(cv, gold["grad_" + k])
        rows.append(("grad." + k, rel(c_grad[k], v), g))
    return rows
