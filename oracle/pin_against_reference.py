#!/usr/bin/env python
"""Pin the oracle against the REAL reference and freeze golden vectors.

Runs only in the build container (needs /root/reference).  It
  1. converts the rig pickles that are present into compact fixture ``.npz`` files
     (``tests/golden/models``) -- data only, no reference source is copied;
  2. imports the unmodified reference ``smal_model.smal_torch.SMAL`` through a ``config``
     shim (the default ``config.SMAL_FILE`` is a missing blob, config.py:48) and runs every
     case of ``tests/rigs.CASES`` forward + backward in fp32;
  3. checks ``oracle/smal_oracle.py`` (fp32 and fp64) against those runs;
  4. freezes the reference's outputs/gradients to ``tests/golden/case_*.npz``;
  5. cross-checks ``oracle/projection_oracle.py`` against the reference's own numpy
     ``project_world_to_pixel`` o ``_sleap_to_pytorch3d_camera`` and freezes those vectors.

Usage:  python oracle/pin_against_reference.py [--ref /root/reference]
"""
from __future__ import annotations

import argparse
import ast
import importlib.util
import json
import os
import pickle
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import projection_oracle as PO  # noqa: E402
from oracle import smal_oracle as SO  # noqa: E402
from tests import rigs  # noqa: E402


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = np.abs(a - b).max() if a.size else 0.0
    return float(d / max(np.abs(b).max() if b.size else 0.0, 1e-30))


def install_config_shim(ref, smal_file):
    """exec the reference's config.py with SMAL_FILE pointed at ``smal_file``."""
    src = open(os.path.join(ref, "config.py")).read()
    marker = 'SMAL_FILE = join("3D_model_prep", "SMIL_OmniAnt.pkl")'
    assert marker in src, "reference config.py layout changed"
    src = src.replace(marker, "SMAL_FILE = %r" % smal_file)
    mod = types.ModuleType("config")
    mod.__file__ = os.path.join(ref, "config.py")
    cwd = os.getcwd()
    os.chdir(ref)
    try:
        exec(compile(src, mod.__file__, "exec"), mod.__dict__)
    finally:
        os.chdir(cwd)
    sys.modules["config"] = mod
    return mod


def export_fixture(dd, path):
    keep = {}
    for k in ("f", "v_template", "shapedirs", "posedirs", "J_regressor", "J", "weights", "kintree_table",
              "scaledirs", "transdirs", "sym_verts", "shape_mean_betas", "shape_cov"):
        if k in dd and isinstance(dd[k], np.ndarray):
            a = dd[k]
            keep[k] = a.astype(np.float32) if a.dtype.kind == "f" else a
    keep["J_names"] = np.array([str(s) for s in dd["J_names"]])
    keep["static_joint_locs"] = np.array(bool(dd.get("static_joint_locs", False)))
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **keep)


def run_reference(SMAL, cfgmod, inp, call, cfg, up):
    """One fp32 forward+backward through the unmodified reference."""
    defaults = dict(ALLOW_LIMB_SCALING=True, ignore_hardcoded_body=True)
    for k, v in {**defaults, **cfg}.items():
        setattr(cfgmod, k, v)
    model = SMAL(torch.device("cpu"))
    t = {k: torch.from_numpy(v.copy()).requires_grad_(True) for k, v in inp.items()}
    kw = dict(call)
    if "v_template" in kw:
        kw["v_template"] = torch.from_numpy(kw["v_template"])
    args = dict(trans=t.get("trans"), del_v=t.get("del_v"), betas_logscale=t.get("betas_logscale"),
                betas_trans=t.get("betas_trans"), **kw)
    res = model(t["beta"], t["theta"], **args)
    if kw.get("get_skin", True) is False:
        outs = {"joints": res}
    else:
        outs = dict(zip(("verts", "joints", "Rs", "v_shaped"), res))
    outs["J_transformed"] = model.J_transformed
    loss = 0.0
    for k in ("verts", "joints", "Rs", "v_shaped"):
        g = up.get("g_" + k)
        if g is not None and k in outs and outs[k].requires_grad:
            loss = loss + (outs[k] * torch.from_numpy(g)).sum()
    grads = {}
    if isinstance(loss, torch.Tensor):
        loss.backward()
        for k, v in t.items():
            if v.grad is not None:
                grads[k] = v.grad.numpy().copy()
    return {k: v.detach().numpy().copy() for k, v in outs.items()}, grads


def run_oracle(dd, dtype, inp, call, cfg, up):
    static = cfg.get("STATIC_JOINT_LOCATIONS", bool(dd.get("static_joint_locs", False)))
    m = SO.build_model(dd, dtype=dtype, static_joints=static)
    t = {k: torch.from_numpy(v.copy()).to(dtype).requires_grad_(True) for k, v in inp.items()}
    kw = dict(call)
    get_skin = kw.pop("get_skin", True)
    if "v_template" in kw:
        kw["v_template"] = torch.from_numpy(kw["v_template"]).to(dtype)
    o = SO.smal_forward_full(
        m, t["beta"], t["theta"], t.get("trans"), t.get("del_v"), t.get("betas_logscale"), t.get("betas_trans"),
        kw.get("v_template"), kw.get("propagate_scaling", False), cfg.get("ALLOW_LIMB_SCALING", True),
        cfg.get("ignore_hardcoded_body", True))
    keys = ("verts", "joints", "Rs", "v_shaped") if get_skin else ("joints",)
    loss = 0.0
    for k in keys:
        g = up.get("g_" + k)
        if g is not None and o[k].requires_grad:
            loss = loss + (o[k] * torch.from_numpy(g).to(dtype)).sum()
    grads = {}
    if isinstance(loss, torch.Tensor):
        loss.backward()
        grads = {k: v.grad.numpy().copy() for k, v in t.items() if v.grad is not None}
    outs = {k: o[k].detach().numpy().copy() for k in keys + ("J_transformed",)}
    return outs, grads


def pin_smal(ref, report):
    tmp = tempfile.mkdtemp(prefix="smil_pin_")
    rig_files = {
        "stick": os.path.join(ref, "3D_model_prep", "SMILy_STICK.pkl"),
        "mouse": os.path.join(ref, "3D_model_prep", "SMILy_Mouse_static_joints_REPOSE.pkl"),
    }
    for name, spec in rigs.SYNTH_SPECS.items():
        p = os.path.join(tmp, name + ".pkl")
        with open(p, "wb") as fh:
            pickle.dump(rigs.synthetic_rig(**spec), fh, protocol=2)
        rig_files[name] = p
    for name in ("stick", "mouse"):
        export_fixture(SO.load_model_dict(rig_files[name]), rigs.REAL_RIGS[name])

    worst = {"o32_fwd": 0.0, "o32_grad": 0.0, "o64_fwd": 0.0, "o64_grad": 0.0}
    for case in rigs.CASES:
        rig = case["rig"]
        # fixture / generator view of the rig (what the tests will see)
        dd = rigs.get_rig(rig)
        inp, call, cfg, up = rigs.case_inputs(case, dd)
        cfgmod = install_config_shim(ref, rig_files[rig])
        if "STATIC_JOINT_LOCATIONS" in cfg:
            cfgmod.STATIC_JOINT_LOCATIONS = cfg["STATIC_JOINT_LOCATIONS"]
        for mname in [k for k in sys.modules if k.startswith("smal_model")]:
            del sys.modules[mname]
        from smal_model.smal_torch import SMAL  # the unmodified reference

        r_out, r_grad = run_reference(SMAL, cfgmod, inp, call, cfg, up)
        entry = {}
        for tag, dt in (("o32", torch.float32), ("o64", torch.float64)):
            o_out, o_grad = run_oracle(dd, dt, inp, call, cfg, up)
            assert set(o_grad) == set(r_grad), (case["name"], set(o_grad), set(r_grad))
            ef = max(rel(o_out[k], r_out[k]) for k in r_out if k in o_out)
            eg = max([rel(o_grad[k], r_grad[k]) for k in r_grad] or [0.0])
            for k in r_out:
                assert o_out[k].shape == r_out[k].shape, (case["name"], k, o_out[k].shape, r_out[k].shape)
            entry[tag] = {"fwd": ef, "grad": eg}
            worst[tag + "_fwd"] = max(worst[tag + "_fwd"], ef)
            worst[tag + "_grad"] = max(worst[tag + "_grad"], eg)
        report["cases"][case["name"]] = entry
        print("%-20s o32 fwd %.2e grad %.2e | o64 fwd %.2e grad %.2e" % (
            case["name"], entry["o32"]["fwd"], entry["o32"]["grad"], entry["o64"]["fwd"], entry["o64"]["grad"]))
        # freeze
        s = case["stride"]
        save = {"in_" + k: v for k, v in inp.items() if k != "del_v"}
        for k, v in r_out.items():
            save["out_" + k] = v[:, ::s] if k in ("verts", "v_shaped") else v
        for k, v in r_grad.items():
            save["grad_" + k] = v[:, ::s] if k == "del_v" else v
        np.savez_compressed(os.path.join(rigs.GOLDEN, "case_%s.npz" % case["name"]), **save)
    report["worst"] = worst
    assert worst["o32_fwd"] < 2e-6 and worst["o64_fwd"] < 2e-6, worst
    assert worst["o32_grad"] < 2e-5 and worst["o64_grad"] < 2e-5, worst


def pin_projection(ref, report):
    """oracle FoV screen projection == reference numpy OpenCV projection (centred principal point)."""
    spec = importlib.util.spec_from_file_location(
        "_cf", os.path.join(ref, "smal_fitter", "multiview_common", "canonical_frame.py"))
    cf = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cf)
    src = open(os.path.join(ref, "smal_fitter", "sleap_data", "sleap_multiview_dataset.py")).read()
    fn = None
    for node in ast.walk(ast.parse(src)):
        if isinstance(node, ast.FunctionDef) and node.name == "_sleap_to_pytorch3d_camera":
            node.decorator_list = []
            fn = ast.get_source_segment(src, node)
    assert fn is not None
    import textwrap
    from typing import Tuple

    ns = {"np": np, "Tuple": Tuple}
    exec(textwrap.dedent(fn), ns)
    to_p3d = ns["_sleap_to_pytorch3d_camera"]

    rs = np.random.RandomState(11)
    S = 512.0
    rows = {k: [] for k in ("R_cv", "t_cv", "K", "wh", "X", "px", "R_p3d", "T_p3d", "fov", "aspect")}
    worst = 0.0
    for _ in range(200):
        A = rs.standard_normal((3, 3))
        Q, _r = np.linalg.qr(A)
        if np.linalg.det(Q) < 0:
            Q[:, 0] *= -1
        t = np.array([rs.uniform(-0.3, 0.3), rs.uniform(-0.3, 0.3), rs.uniform(2.0, 4.0)])
        W, H = rs.choice([640, 1280, 512]), rs.choice([480, 1024, 512])
        fx, fy = rs.uniform(500, 1500), rs.uniform(500, 1500)
        K = np.array([[fx, 0, W / 2.0], [0, fy, H / 2.0], [0, 0, 1.0]])
        X = rs.uniform(-0.5, 0.5, size=(16, 3))
        px = cf.project_world_to_pixel(X, Q, t, K)  # (16,2) [x,y] pixels in WxH
        Rp, Tp, fov, asp = to_p3d(Q.astype(np.float64), t.astype(np.float64), K, np.array([W, H]))
        scr = PO.fov_project_screen(
            torch.from_numpy(X)[None], torch.from_numpy(Rp.astype(np.float64))[None],
            torch.from_numpy(Tp.astype(np.float64))[None], torch.tensor([fov], dtype=torch.float64),
            torch.tensor([asp], dtype=torch.float64), S)[0].numpy()
        mine = np.stack([scr[:, 0] / S * W, scr[:, 1] / S * H], axis=1)
        ok = np.isfinite(px).all(axis=1)
        worst = max(worst, float(np.abs(mine[ok] - px[ok]).max()))
        for k, v in (("R_cv", Q), ("t_cv", t), ("K", K), ("wh", np.array([W, H], dtype=np.float64)), ("X", X),
                     ("px", px), ("R_p3d", Rp), ("T_p3d", Tp), ("fov", fov), ("aspect", asp)):
            rows[k].append(np.asarray(v, dtype=np.float64))
    # R_p3d/T_p3d are float32-rounded by the reference helper -> ~1e-4 px at 1e3 px scale
    print("projection cross-check: worst |oracle - reference numpy| = %.3e px over 200 cameras" % worst)
    report["projection_crosscheck_px"] = worst
    assert worst < 5e-3
    np.savez_compressed(os.path.join(rigs.GOLDEN, "projection_opencv.npz"), **{k: np.stack(v) for k, v in rows.items()})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    sys.path.insert(0, a.ref)
    torch.manual_seed(0)
    os.makedirs(rigs.GOLDEN, exist_ok=True)
    report = {"reference": a.ref, "torch": torch.__version__, "cases": {}}
    pin_smal(a.ref, report)
    pin_projection(a.ref, report)
    with open(os.path.join(rigs.GOLDEN, "PIN_REPORT.json"), "w") as fh:
        json.dump(report, fh, indent=1, sort_keys=True)
    print("pinned:", json.dumps(report["worst"]))


if __name__ == "__main__":
    main()
