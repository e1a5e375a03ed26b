"""Synthetic SMPL-style rigs and seeded input generators shared by the tests, the pin
script and bench.py.  Pure numpy; nothing here touches the GPU or the reference."""
from __future__ import annotations

import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
MODEL_DIR = os.path.join(GOLDEN, "models")

REAL_RIGS = {
    "stick": os.path.join(MODEL_DIR, "SMILy_STICK.npz"),
    "mouse": os.path.join(MODEL_DIR, "SMILy_Mouse_static_joints_REPOSE.npz"),
}


def synthetic_rig(seed=0, V=257, J=9, nB=4, K=3, posedirs=False, static=False, wide=False, chain=False):
    """A random rig in the on-disk SMPL schema (the keys ``SMAL.__init__`` reads,
    /root/reference/smal_model/smal_torch.py:92-196).

    ``K`` = max skinning weights per vertex; ``wide`` = star-shaped tree (many children on
    the root); ``chain`` = a single deep chain; otherwise a random tree with parent[i] < i.
    """
    rs = np.random.RandomState(seed)
    parents = np.zeros(J, dtype=np.int64)
    parents[0] = -1
    for i in range(1, J):
        if wide:
            parents[i] = 0 if i < J - 2 else i - 1
        elif chain:
            parents[i] = i - 1
        else:
            parents[i] = rs.randint(0, i)
    v_template = rs.uniform(-1, 1, size=(V, 3))
    shapedirs = rs.normal(0, 0.03, size=(V, 3, nB))
    weights = np.zeros((V, J), dtype=np.float32)
    for v in range(V):
        k = rs.randint(1, K + 1)
        idx = rs.choice(J, size=min(k, J), replace=False)
        w = rs.uniform(0.1, 1.0, size=len(idx))
        weights[v, idx] = (w / w.sum()).astype(np.float32)
    jreg = np.zeros((J, V), dtype=np.float32)
    if not static:
        for j in range(J):
            n = rs.randint(1, min(12, V))
            idx = rs.choice(V, size=n, replace=False)
            w = rs.uniform(0.1, 1.0, size=n)
            jreg[j, idx] = (w / w.sum()).astype(np.float32)
    Jloc = rs.uniform(-1, 1, size=(J, 3)).astype(np.float32)
    if posedirs:
        pd = rs.normal(0, 0.01, size=(V, 3, 9 * (J - 1)))
    else:
        pd = np.zeros((0,))
    F = max(4, V // 2)
    faces = rs.randint(0, V, size=(F, 3)).astype(np.int32)
    dd = {
        "f": faces,
        "v_template": v_template,
        "shapedirs": shapedirs,
        "posedirs": pd,
        "J_regressor": jreg,
        "J": Jloc,
        "weights": weights,
        "kintree_table": np.stack([parents, np.arange(J)]).astype(np.int32),
        "J_names": ["j%d" % i for i in range(J)],
        "scaledirs": rs.normal(0, 0.05, size=(nB, J, 3)),
        "transdirs": rs.normal(0, 0.05, size=(nB, J, 3)),
        "sym_verts": np.arange(3),
    }
    if static:
        dd["static_joint_locs"] = True
    return dd


SYNTH_SPECS = {
    # name: kwargs -- small rigs that exercise every branch the real rigs do not
    "syn_k3": dict(seed=1, V=257, J=9, nB=4, K=3),
    "syn_pose": dict(seed=2, V=130, J=7, nB=3, K=2, posedirs=True),
    "syn_static_k4": dict(seed=3, V=203, J=12, nB=2, K=4, static=True),
    "syn_wide": dict(seed=4, V=96, J=40, nB=6, K=2, wide=True),
    "syn_chain": dict(seed=5, V=64, J=21, nB=1, K=2, chain=True),
    "syn_j35": dict(seed=6, V=3100, J=35, nB=5, K=2),  # legacy 35-joint branch (6 vertex picks)
}


def padded_rig(base, nB=None, J=None, seed=13):
    """Stand-ins for the two BASELINE rigs whose pickles are missing:
    ``nB``: the base rig's geometry with its shape space padded to ``nB`` directions -- random directions made
    orthogonal to each other, scaled to the abs-max of the rig's own shapedirs (config 4: SMIL_OmniAnt has nB = 13);
    ``J``: extra LEAF joints hung on random existing joints, no vertex bound to them (config 3: the Falkner mouse has
    J = 40, 9 more than the rig present); scaledirs / transdirs are padded alike."""
    from oracle.smal_oracle import load_model_dict

    dd = dict(load_model_dict(REAL_RIGS[base]))
    rs = np.random.RandomState(seed)
    sd = np.asarray(dd["shapedirs"], dtype=np.float64)            # (V, 3, nB0)
    V, _, nB0 = sd.shape
    J0 = np.asarray(dd["kintree_table"]).shape[1]
    if nB is not None and nB > nB0:
        extra = rs.standard_normal((3 * V, nB - nB0))
        q, _r = np.linalg.qr(extra)
        q = q / np.abs(q).max(axis=0, keepdims=True) * np.abs(sd).max()
        dd["shapedirs"] = np.concatenate([sd, q.reshape(V, 3, nB - nB0)], axis=2)
        for k in ("scaledirs", "transdirs"):
            if k in dd:
                a = np.asarray(dd[k], dtype=np.float64)           # (nB0, J, 3)
                dd[k] = np.concatenate([a, rs.normal(0, np.abs(a).std() + 1e-3, (nB - nB0,) + a.shape[1:])], axis=0)
    if J is not None and J > J0:
        add = J - J0
        kt = np.asarray(dd["kintree_table"]).astype(np.int64)
        par = list(kt[0])
        for _ in range(add):
            par.append(int(rs.randint(0, J0)))
        dd["kintree_table"] = np.stack([np.asarray(par), np.arange(J)]).astype(np.int32)
        dd["weights"] = np.concatenate([np.asarray(dd["weights"]), np.zeros((V, add), np.float32)], axis=1)
        jr = np.asarray(dd["J_regressor"])
        jr = jr if jr.shape[0] == J0 else jr.T                      # (J0, V)
        dd["J_regressor"] = np.concatenate([jr, np.zeros((add, V), jr.dtype)], axis=0)
        Jl = np.asarray(dd["J"], dtype=np.float64)
        dd["J"] = np.concatenate([Jl, Jl[np.asarray(par[J0:])] + rs.normal(0, 0.05, (add, 3))], axis=0)
        dd["J_names"] = list(dd["J_names"]) + ["pad%d" % i for i in range(add)]
        for k in ("scaledirs", "transdirs"):
            if k in dd:
                a = np.asarray(dd[k], dtype=np.float64)
                dd[k] = np.concatenate([a, np.zeros((a.shape[0], add, 3))], axis=1)
    return dd


PADDED_SPECS = {
    "stick13": dict(base="stick", nB=13),    # config 4 stand-in for SMIL_OmniAnt (J = 55, nB = 13)
    "mouse40": dict(base="mouse", J=40),     # config 3 stand-in for the Falkner mouse (J = 40)
}


def get_rig(name):
    if name in REAL_RIGS:
        from oracle.smal_oracle import load_model_dict

        return load_model_dict(REAL_RIGS[name])
    if name in PADDED_SPECS:
        return padded_rig(**PADDED_SPECS[name])
    return synthetic_rig(**SYNTH_SPECS[name])


def make_inputs(dd, B, seed=0, theta_std=0.3, beta_std=0.5, with_scale=False, with_trans_off=False, with_del_v=False,
                zero_theta=False, rotmat=False, nb_used=None):
    """Seeded fp32 inputs with the distributions of SURVEY.md section 8(d), config 1."""
    rs = np.random.RandomState(seed)
    V = np.asarray(dd["v_template"]).shape[0]
    J = np.asarray(dd["kintree_table"]).shape[1]
    nB = np.asarray(dd["shapedirs"]).shape[-1]
    out = {
        "beta": (rs.standard_normal((B, nB)) * beta_std).astype(np.float32),
        "theta": (rs.standard_normal((B, J, 3)) * theta_std).astype(np.float32),
        "trans": rs.standard_normal((B, 3)).astype(np.float32),
    }
    ls = (rs.standard_normal((B, J, 3)) * 0.05).astype(np.float32)
    bt = (rs.standard_normal((B, J, 3)) * 0.01).astype(np.float32)
    dv = (rs.standard_normal((B, V, 3)) * 1e-3).astype(np.float32)
    if with_scale:
        out["betas_logscale"] = ls
    if with_trans_off:
        out["betas_trans"] = bt
    if with_del_v:
        out["del_v"] = dv
    if zero_theta:
        out["theta"] = np.zeros_like(out["theta"])
    if nb_used is not None:
        out["beta"] = out["beta"][:, :nb_used].copy()
    if rotmat:
        # rotation matrices via an independent closed form (unit axis, exact angle)
        th = out["theta"].astype(np.float64).reshape(-1, 3)
        ang = np.linalg.norm(th, axis=1, keepdims=True)
        ax = th / np.maximum(ang, 1e-12)
        Kx = np.zeros((th.shape[0], 3, 3))
        Kx[:, 0, 1], Kx[:, 0, 2] = -ax[:, 2], ax[:, 1]
        Kx[:, 1, 0], Kx[:, 1, 2] = ax[:, 2], -ax[:, 0]
        Kx[:, 2, 0], Kx[:, 2, 1] = -ax[:, 1], ax[:, 0]
        R = np.eye(3)[None] + np.sin(ang)[:, :, None] * Kx + (1 - np.cos(ang))[:, :, None] * (Kx @ Kx)
        out["theta"] = R.reshape(B, J, 3, 3).astype(np.float32)
    return out


def make_upstream(B, V, nJ_out, seed=1):
    rs = np.random.RandomState(seed + 1000)
    return {
        "g_verts": rs.standard_normal((B, V, 3)).astype(np.float32),
        "g_joints": rs.standard_normal((B, nJ_out, 3)).astype(np.float32),
        "g_Rs": rs.standard_normal((B, 0, 3, 3)).astype(np.float32),
    }


# --------------------------------------------------------------------------------------
# the golden case table (SURVEY.md section 8(d), config 1 and its variants)
# --------------------------------------------------------------------------------------
def _c(name, rig, B, inp=None, call=None, cfg=None, up_all=False, stride=1):
    return dict(name=name, rig=rig, B=B, inp=inp or {}, call=call or {}, cfg=cfg or {}, up_all=up_all, stride=stride)


CASES = [
    # STICK (V=3020, J=55, nB=5, regressed joints)
    _c("stick_base", "stick", 32, stride=8),
    _c("stick_full", "stick", 3, dict(with_scale=True, with_trans_off=True), up_all=True),
    _c("stick_theta0", "stick", 3, dict(zero_theta=True, with_scale=True, with_trans_off=True), stride=8),
    _c("stick_prop", "stick", 3, dict(with_scale=True, with_trans_off=True), dict(propagate_scaling=True), stride=8),
    _c("stick_rotmat", "stick", 3, dict(rotmat=True), stride=8, up_all=True),
    _c("stick_beta3", "stick", 3, dict(nb_used=3), stride=8),
    _c("stick_delv", "stick", 3, dict(with_del_v=True, with_scale=True), stride=8, up_all=True),
    _c("stick_noscale", "stick", 3, dict(with_scale=True, with_trans_off=True), cfg=dict(ALLOW_LIMB_SCALING=False),
       stride=8),
    _c("stick_notrans", "stick", 3, dict(), dict(trans=None), stride=8),
    _c("stick_jointsonly", "stick", 3, dict(with_scale=True), dict(get_skin=False), stride=8),
    _c("stick_vtemplate", "stick", 3, dict(), dict(v_template="perturbed"), stride=8),
    _c("stick_badtheta", "stick", 3, dict(), dict(theta_width=7), stride=8),
    _c("stick_b1_nobeta", "stick", 1, dict(nb_used=0), stride=8),
    # mouse (V=11263, J=31, nB=3, static joints)
    _c("mouse_base", "mouse", 4, stride=16),
    _c("mouse_full", "mouse", 2, dict(with_scale=True, with_trans_off=True), dict(propagate_scaling=True), up_all=True,
       stride=4),
    _c("mouse_regress", "mouse", 2, dict(with_scale=True), cfg=dict(STATIC_JOINT_LOCATIONS=False), stride=16),
    # synthetic rigs
    _c("syn_k3_full", "syn_k3", 5, dict(with_scale=True, with_trans_off=True), up_all=True),
    _c("syn_k3_delv_prop", "syn_k3", 5, dict(with_del_v=True, with_scale=True, with_trans_off=True),
       dict(propagate_scaling=True), up_all=True),
    _c("syn_pose_full", "syn_pose", 5, dict(with_scale=True, with_trans_off=True), up_all=True),
    _c("syn_pose_rotmat", "syn_pose", 4, dict(rotmat=True), up_all=True),
    _c("syn_static_k4", "syn_static_k4", 5, dict(with_scale=True, with_trans_off=True), up_all=True),
    _c("syn_wide", "syn_wide", 4, dict(with_scale=True, with_trans_off=True), up_all=True),
    _c("syn_chain", "syn_chain", 4, dict(with_scale=True, with_trans_off=True), dict(propagate_scaling=True),
       up_all=True),
    _c("syn_j35_legacy", "syn_j35", 2, dict(), cfg=dict(ignore_hardcoded_body=False), stride=8),
    # odd batches (sample-pair tails), betas_trans without log-scales on the fused backward
    _c("stick_odd5", "stick", 5, stride=8),
    _c("stick_btrans", "stick", 5, dict(with_trans_off=True), stride=8),
    _c("mouse_odd7", "mouse", 7, dict(with_trans_off=True), stride=16),
    # stand-ins for the missing BASELINE rigs: nB = 13 (OmniAnt), J = 40 (Falkner mouse)
    _c("stick13_base", "stick13", 6, stride=8),
    _c("stick13_full", "stick13", 3, dict(with_scale=True, with_trans_off=True), dict(propagate_scaling=True), stride=8),
    _c("mouse40_base", "mouse40", 3, dict(with_trans_off=True), stride=16),
]
CASE_BY_NAME = {c["name"]: c for c in CASES}


def case_inputs(case, dd):
    """Materialise a case: (inputs dict of np arrays, call kwargs, cfg flags, upstream grads)."""
    V = np.asarray(dd["v_template"]).shape[0]
    J = np.asarray(dd["kintree_table"]).shape[1]
    inp = make_inputs(dd, case["B"], seed=abs(hash_name(case["name"])) % 10000, **case["inp"])
    call = dict(case["call"])
    if "trans" in call and call["trans"] is None:
        inp.pop("trans")
        call.pop("trans")
    if call.get("v_template") == "perturbed":
        rs = np.random.RandomState(77)
        call["v_template"] = (np.asarray(dd["v_template"]) + rs.standard_normal((V, 3)) * 0.01).astype(np.float32)
    if "theta_width" in call:
        inp["theta"] = inp["theta"][:, : call.pop("theta_width")].copy()
    static = bool(dd.get("static_joint_locs", False))
    static = case["cfg"].get("STATIC_JOINT_LOCATIONS", static)
    legacy = J == 35 and not case["cfg"].get("ignore_hardcoded_body", True)
    nJ_out = J + (6 if legacy else 0)
    up = make_upstream(case["B"], V, nJ_out, seed=case["B"])
    rs = np.random.RandomState(5)
    if case["up_all"]:
        up["g_Rs"] = rs.standard_normal((case["B"], J, 3, 3)).astype(np.float32)
        up["g_v_shaped"] = rs.standard_normal((case["B"], V, 3)).astype(np.float32)
    else:
        up.pop("g_Rs")
    return inp, call, dict(case["cfg"]), up


def hash_name(s):
    h = 0
    for ch in s:
        h = (h * 131 + ord(ch)) % 1000003
    return h


# ---- seeded scenes of the reprojection-loss parity tests (tests/test_gpu_losses.py, oracle/pin_against_reference.py) ----
LOSS_SCENES = {
    # name: (B, V, J, seed, shared camera)
    "mv_b5_v3_j11": (5, 3, 11, 7, False),
    "mv_b3_v18_j40": (3, 18, 40, 7, False),
    "mv_b2_v1_j55": (2, 1, 55, 7, False),
    "mv_b64_v5_j55": (64, 5, 55, 7, False),
    "mv_b5_v3_j11_shared": (5, 3, 11, 7, True),
    "mv_b3_v18_j40_shared": (3, 18, 40, 7, True),
    "mv_b2_v1_j55_shared": (2, 1, 55, 7, True),
    "mv_b64_v5_j55_shared": (64, 5, 55, 7, True),
}


def loss_scene(B, V, J, seed, shared_cam=False, S=512.0):
    """fp32-representable joints / ring cameras / noisy targets / visibility / masks / joint weights of one
    multi-view scene (SURVEY.md section 8(d), config 3 set-up: look_at ring, radius 3, elevation 15 degrees)."""
    import torch

    from oracle import projection_oracle as PO

    rs = np.random.RandomState(seed)
    f = lambda a: np.asarray(a, dtype=np.float32)
    joints = f(rs.uniform(-0.4, 0.4, (B, J, 3)))
    R, T = PO.look_at_ring(V, 3.0, 15.0, dtype=torch.float64)
    R, T = R.numpy(), T.numpy()
    if shared_cam:
        R, T = f(R[None]), f(T[None])
        fov, asp = f(np.full((1, V), 60.0)), f(np.ones((1, V)))
    else:
        R = f(np.broadcast_to(R[None], (B, V, 3, 3)))
        T = f(np.broadcast_to(T[None], (B, V, 3)) + rs.normal(0, 0.05, (B, V, 3)))
        fov, asp = f(rs.uniform(40, 70, (B, V))), f(rs.uniform(0.8, 1.3, (B, V)))
    t64 = lambda a: torch.from_numpy(np.ascontiguousarray(a)).double()
    with torch.no_grad():
        clean = PO.multiview_project(t64(joints), t64(R).expand(B, V, 3, 3), t64(T).expand(B, V, 3), t64(fov).expand(B, V),
                                     t64(asp).expand(B, V), S).numpy()
    tgt = f(clean + rs.normal(0, 2.0 / S, clean.shape))
    tgt[0, 0, 0] = 1.5  # out-of-bounds GT is masked out
    vis = f(rs.uniform(size=(B, V, J)) < 0.8)
    vmask = np.ones((B, V), dtype=np.float32)
    vmask[-1, -1] = 0.0
    jw = f(rs.uniform(0.5, 2.0, J))
    return dict(joints=joints, R=R, T=T, fov=fov, aspect=asp, target=tgt, vis=vis, view_mask=vmask, joint_weights=jw,
                image_size=np.float32(S))
