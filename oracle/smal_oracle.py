"""CPU restatement of SMILify's SMAL/SMIL forward pass (torch-CPU, any float dtype).

TEST INFRASTRUCTURE.  This file is the oracle the CUDA path is checked against and
the ``cpu_baseline`` that ``bench.py`` times.  The product (``smilify_b200/``) never
imports it.

Parity status: PINNED.  ``oracle/pin_against_reference.py`` imports the real
reference (``/root/reference/smal_model/smal_torch.py`` through a ``config`` shim) in
the build container, checks this restatement against it on every rig present
(forward outputs and all input gradients) and freezes the reference's own outputs
into ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` re-checks the oracle
against those frozen vectors wherever the tests run.

The restatement is deliberately *dense and op-by-op* like the reference (dense
``W @ A`` skinning, dense joint regressor, per-joint Python loop) so that (a) it is
an independent check of the sparse/fused CUDA formulation and (b) its CPU timing
has the cost structure of the reference's own CPU path.

All ``file:line`` citations are relative to ``/root/reference``.
"""
from __future__ import annotations

import pickle
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch


# --------------------------------------------------------------------------------------
# model loading (smal_model/smal_torch.py:21-73, 88-196)
# --------------------------------------------------------------------------------------
class _Chumpy:
    """Stand-in for ``chumpy.ch.Ch`` in legacy pickles (smal_torch.py:21-58)."""

    def __init__(self, *a, **k):
        self.x = np.array(a[0]) if a else np.array([])

    def __setstate__(self, state):
        if isinstance(state, dict):
            self.x = np.array(state.get("x", []))
        else:
            self.x = np.array(state[0] if isinstance(state, (tuple, list)) else state)


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module == "chumpy.ch" and name == "Ch":
            return _Chumpy
        return super().find_class(module, name)


def _as_array(x):
    if isinstance(x, _Chumpy):
        return np.asarray(x.x)
    if hasattr(x, "todense"):
        return np.asarray(x.todense())
    return x


def load_model_dict(path: str) -> dict:
    """Read a SMPL-style ``.pkl`` (latin1, chumpy-safe) or a fixture ``.npz`` into a dict."""
    if str(path).endswith(".npz"):
        with np.load(path, allow_pickle=False) as z:
            d = {k: z[k] for k in z.files}
        if "static_joint_locs" in d:
            d["static_joint_locs"] = bool(d["static_joint_locs"])
        return d
    with open(path, "rb") as fh:
        raw = _Unpickler(fh, encoding="latin1").load()
    return {k: _as_array(v) for k, v in raw.items()}


@dataclass
class OracleModel:
    """Dense model tensors exactly as ``SMAL.__init__`` lays them out (smal_torch.py:88-196)."""

    v_template: torch.Tensor  # (V,3)
    shapedirs: torch.Tensor  # (nB, 3V)            smal_torch.py:124
    posedirs: torch.Tensor  # (9(J-1), 3V)        smal_torch.py:176-190 (zeros when empty)
    J_regressor: torch.Tensor  # (V, J)              smal_torch.py:169-172
    weights: torch.Tensor  # (V, J)              smal_torch.py:196
    parents: np.ndarray  # int32 (J,)          smal_torch.py:193
    faces: np.ndarray
    J: Optional[torch.Tensor] = None  # (J,3) static joints smal_torch.py:175-176
    static_joints: bool = False
    extras: dict = field(default_factory=dict)

    @property
    def num_joints(self) -> int:
        return self.J_regressor.shape[1]

    @property
    def num_verts(self) -> int:
        return self.v_template.shape[0]

    @property
    def num_betas(self) -> int:
        return self.shapedirs.shape[0]


def build_model(dd: dict, dtype=torch.float64, static_joints: Optional[bool] = None) -> OracleModel:
    """Dense tensors from a model dict.  Values are first rounded to fp32 like the
    reference's ``torch.Tensor(np_array)`` does, then widened to ``dtype``."""

    def t(a):
        return torch.from_numpy(np.asarray(a, dtype=np.float32).copy()).to(dtype)

    nb = np.asarray(dd["shapedirs"]).shape[-1]
    sd = np.reshape(np.asarray(dd["shapedirs"]), [-1, nb]).T  # (nB, 3V)
    jr = np.asarray(_as_array(dd["J_regressor"]))  # (J, V)
    V = np.asarray(dd["v_template"]).shape[0]
    nj = jr.shape[0]
    pd = np.asarray(dd["posedirs"])
    if pd.size != 0:
        pd = np.reshape(pd, [-1, pd.shape[-1]]).T
    else:
        pd = np.zeros(((nj - 1) * 9, V * 3))
    if static_joints is None:
        static_joints = bool(dd.get("static_joint_locs", False))
    return OracleModel(
        v_template=t(dd["v_template"]),
        shapedirs=t(sd),
        posedirs=t(pd),
        J_regressor=t(jr.T),
        weights=t(dd["weights"]),
        parents=np.asarray(dd["kintree_table"])[0].astype(np.int32),
        faces=np.asarray(dd["f"]).astype(np.int64),
        J=t(dd["J"]) if "J" in dd else None,
        static_joints=static_joints,
        extras={k: dd[k] for k in ("scaledirs", "transdirs") if k in dd},
    )


# --------------------------------------------------------------------------------------
# Rodrigues (smal_model/batch_lbs.py:10-50)
# --------------------------------------------------------------------------------------
def rodrigues(theta: torch.Tensor) -> torch.Tensor:
    """(N,3) axis-angle -> (N,3,3).  ``angle = ||theta + 1e-8||`` but the axis divides the
    *unshifted* theta (batch_lbs.py:37-38), so ``r`` is not exactly unit-length."""
    angle = torch.linalg.vector_norm(theta + 1e-8, dim=1, keepdim=True)  # (N,1)
    r = theta / angle
    c = torch.cos(angle)[:, :, None]
    s = torch.sin(angle)[:, :, None]
    rx, ry, rz = r[:, 0], r[:, 1], r[:, 2]
    zero = torch.zeros_like(rx)
    # batch_lbs.py:17-26: flat slots 1,2,3,5,6,7 <- -z, y, z, -x, -y, x
    K = torch.stack([zero, -rz, ry, rz, zero, -rx, -ry, rx, zero], dim=1).reshape(-1, 3, 3)
    outer = r[:, :, None] * r[:, None, :]
    eye = torch.eye(3, dtype=theta.dtype).expand(theta.shape[0], 3, 3)
    return c * eye + (1.0 - c) * outer + s * K


# --------------------------------------------------------------------------------------
# kinematic chain (smal_model/batch_lbs.py:75-197)
# --------------------------------------------------------------------------------------
def global_rigid_transformation(
    Rs: torch.Tensor,
    Js: torch.Tensor,
    parent: np.ndarray,
    betas_logscale: Optional[torch.Tensor] = None,
    betas_trans: Optional[torch.Tensor] = None,
    propagate_scaling: bool = False,
    allow_limb_scaling: bool = True,
    rotate_base: bool = False,
):
    """Returns ``(new_J (N,J,3), A (N,J,4,4))``."""
    N, nj = Rs.shape[0], parent.shape[0]
    dt = Rs.dtype
    R0 = Rs[:, 0]
    if rotate_base:  # batch_lbs.py:103-112
        R0 = R0 @ torch.diag(torch.tensor([1.0, -1.0, -1.0], dtype=dt))
    if not allow_limb_scaling:  # batch_lbs.py:123-124
        betas_logscale = None
    if betas_logscale is not None:
        scale = torch.exp(betas_logscale)  # batch_lbs.py:129-131
    else:
        scale = torch.ones(N, nj, 3, dtype=dt)
    S = torch.diag_embed(scale)  # batch_lbs.py:134
    offs = None
    if betas_trans is not None:  # batch_lbs.py:146-148 (y flipped for Unreal data)
        offs = betas_trans * torch.tensor([1.0, -1.0, 1.0], dtype=dt)

    bottom = torch.zeros(N, 1, 4, dtype=dt)
    bottom[:, 0, 3] = 1.0

    def rigid(R, t):  # batch_lbs.py:136-141
        return torch.cat([torch.cat([R, t[:, :, None]], dim=2), bottom], dim=1)

    G = [rigid(R0, Js[:, 0])]
    for i in range(1, nj):  # batch_lbs.py:155-180
        p = int(parent[i])
        t = Js[:, i] - Js[:, p]
        if offs is not None:
            t = t + offs[:, i]
        if propagate_scaling:
            Sp_inv = torch.eye(3, dtype=dt).expand(N, 3, 3)
        else:
            Sp_inv = torch.linalg.inv(S[:, p])
        G.append(G[p] @ rigid(Sp_inv @ Rs[:, i] @ S[:, i], t))
    G = torch.stack(G, dim=1)  # (N,J,4,4)
    new_J = G[:, :, :3, 3]
    # batch_lbs.py:190-195: subtract the rest-pose bone position from the translation column
    Jh = torch.cat([Js, torch.zeros(N, nj, 1, dtype=dt)], dim=2)[..., None]  # (N,J,4,1)
    moved = G @ Jh
    A = G - torch.nn.functional.pad(moved, (3, 0))
    return new_J, A


# --------------------------------------------------------------------------------------
# SMAL.__call__ (smal_model/smal_torch.py:198-370)
# --------------------------------------------------------------------------------------
def smal_forward(
    m: OracleModel,
    beta: torch.Tensor,
    theta: torch.Tensor,
    trans: Optional[torch.Tensor] = None,
    del_v: Optional[torch.Tensor] = None,
    betas_logscale: Optional[torch.Tensor] = None,
    betas_trans: Optional[torch.Tensor] = None,
    get_skin: bool = True,
    v_template: Optional[torch.Tensor] = None,
    propagate_scaling: bool = False,
    allow_limb_scaling: bool = True,
    ignore_hardcoded_body: bool = True,
):
    """Returns ``(verts, joints, Rs, v_shaped)`` (or ``joints``) plus nothing else; the
    module-level side effect ``self.J_transformed`` is returned by :func:`smal_forward_full`."""
    out = smal_forward_full(
        m, beta, theta, trans, del_v, betas_logscale, betas_trans, v_template, propagate_scaling,
        allow_limb_scaling, ignore_hardcoded_body,
    )
    if get_skin:
        return out["verts"], out["joints"], out["Rs"], out["v_shaped"]
    return out["joints"]


def smal_forward_full(
    m, beta, theta, trans=None, del_v=None, betas_logscale=None, betas_trans=None, v_template=None,
    propagate_scaling=False, allow_limb_scaling=True, ignore_hardcoded_body=True,
):
    dt = m.v_template.dtype
    V, nj = m.num_verts, m.num_joints
    nb = beta.shape[1]  # smal_torch.py:211
    vt = m.v_template if v_template is None else v_template
    # 1. shape blend (smal_torch.py:223-253)
    if nb > 0:
        blend = (beta @ m.shapedirs[:nb]).reshape(-1, V, 3)
        v_shaped = vt + blend if del_v is None else vt + del_v + blend
    else:
        v_shaped = vt.unsqueeze(0) if del_v is None else vt + del_v
    # 2. rest joints (smal_torch.py:257-264)
    if m.static_joints:
        J = m.J.unsqueeze(0).expand(v_shaped.shape[0], -1, -1)
    else:
        J = torch.stack([v_shaped[:, :, c] @ m.J_regressor for c in range(3)], dim=2)
    # 3. rotations + pose blend (smal_torch.py:281-301)
    if theta.shape[1] != nj:  # wrong-width pose is silently zeroed
        theta = torch.zeros(beta.shape[0] if nb > 0 else 1, nj, 3, dtype=dt)
    if theta.dim() == 4:
        Rs = theta
    else:
        Rs = rodrigues(theta.reshape(-1, 3)).reshape(-1, nj, 3, 3)
    feat = (Rs[:, 1:] - torch.eye(3, dtype=dt)).reshape(-1, (nj - 1) * 9)
    v_posed = (feat @ m.posedirs).reshape(-1, V, 3) + v_shaped
    # 4. chain (smal_torch.py:307-315)
    J_transformed, A = global_rigid_transformation(
        Rs, J, m.parents, betas_logscale, betas_trans, propagate_scaling, allow_limb_scaling
    )
    # 5. skinning (smal_torch.py:318-340)
    B = theta.shape[0]
    W = m.weights.repeat(B, 1).reshape(B, V, nj)
    T = (W @ A.reshape(B, nj, 16)).reshape(B, V, 4, 4)
    vh = torch.cat([v_posed, torch.ones(B, V, 1, dtype=dt)], dim=2)
    verts = (T @ vh[..., None])[:, :, :3, 0]
    if trans is None:
        trans = torch.zeros(B, 3, dtype=dt)
    verts = verts + trans[:, None, :]
    # 6. posed joints (smal_torch.py:343-365)
    if m.static_joints:
        joints = J_transformed
    else:
        joints = torch.stack([verts[:, :, c] @ m.J_regressor for c in range(3)], dim=2)
    if nj == 35 and not ignore_hardcoded_body:
        picks = [1863, 26, 2124, 150, 3055, 1097]
        joints = torch.cat([joints] + [verts[:, None, p] for p in picks], dim=1)
    return dict(verts=verts, joints=joints, Rs=Rs, v_shaped=v_shaped, J_transformed=J_transformed, A=A, J=J)


# --------------------------------------------------------------------------------------
# PCA scale/translation weights -> per-joint values
# (smal_fitter/neuralSMIL/smil_image_regressor.py:1166-1232)
# --------------------------------------------------------------------------------------
def pca_weights_to_joint_values(scaledirs, transdirs, scale_w, trans_w=None, translation_factor=0.01):
    """``scaledirs/transdirs`` are (J,3,C) as cached by the regressor (the pickle stores
    (C,J,3); the caller permutes).  Returns (log_beta_scales, betas_trans), both (B,J,3)."""
    scale = 1.0 + torch.einsum("jck,bk->bjc", scaledirs, scale_w)
    log_scales = torch.log(torch.clamp(scale, min=1e-4))
    if trans_w is not None:
        bt = torch.einsum("jck,bk->bjc", transdirs, trans_w) * translation_factor
    else:
        bt = torch.zeros_like(log_scales)
    return log_scales, bt
