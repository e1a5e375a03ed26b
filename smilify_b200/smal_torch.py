"""Drop-in for ``smal_model.smal_torch`` (/root/reference/smal_model/smal_torch.py).

``SMAL(device, shape_family_id=-1, dtype=torch.float)`` and
``SMAL.__call__(beta, theta, trans=None, del_v=None, betas_logscale=None, betas_trans=None,
get_skin=True, v_template=None, propagate_scaling=False)`` keep the reference's signature,
return values, side effect (``self.J_transformed``), plain (non-buffer) attributes and the
points in time at which the global ``config`` is read; the arithmetic runs in the sm_100a
kernels behind the C ABI (include/smil_b200.h).  CPU tensors raise: there is no CPU fallback.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from . import config as _config
from . import packer
from .functions import DeviceModel, SmalFunction
from .model_io import _ChArray, _RigUnpickler, load_rig, reference_layout

# smal_torch.py:353-365 -- legacy 35-joint dog rig: six extra "joints" picked from vertices
_LEGACY_35_PICKS = (1863, 26, 2124, 150, 3055, 1097)


def load_smal_model(file_path):
    """Same name and contract as the reference helper (smal_torch.py:61-73)."""
    return load_rig(file_path)


class CustomUnpickler(_RigUnpickler):
    """Same name and call pattern as the reference's unpickler for legacy chumpy pickles (smal_torch.py:21-58):
    ``CustomUnpickler(f).load()`` -- imported by smal_fitter/priors/pose_prior_35.py:8.  Chumpy arrays become inert
    wrappers with ``.r`` / ``.data`` / ``__array__`` (no chumpy import)."""

    ChumpyWrapper = _ChArray

    def __init__(self, file, encoding="latin1"):
        super().__init__(file, encoding=encoding)


def undo_chumpy(x):
    """smal_torch.py:78-84: the numeric payload of a chumpy variable, anything else unchanged."""
    if isinstance(x, np.ndarray):
        return x
    try:
        return x.r
    except AttributeError:
        return x


class SMAL(nn.Module):
    def __init__(self, device, shape_family_id=-1, dtype=torch.float):
        super().__init__()
        cfg = _config.active()
        device = torch.device(device)
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        dd = load_rig(cfg.SMAL_FILE)  # smal_torch.py:92
        if getattr(cfg, "DEBUG", False):
            print(cfg.SMAL_FILE)
            for key, value in dd.items():
                print(key, getattr(value, "shape", None))

        family = None
        if shape_family_id != -1:  # smal_torch.py:133-140
            family = np.asarray(load_rig_any(cfg.SMAL_DATA_FILE)["cluster_means"][shape_family_id])
        static = bool(getattr(cfg, "STATIC_JOINT_LOCATIONS", False))
        lay = reference_layout(dd, static_joints=static, shape_family_betas=family)

        # symmetry alignment (smal_torch.py:142-162): load-time, CPU; skipped by default (config.py:52)
        v_template = lay["v_template"]
        if getattr(cfg, "ignore_sym", True):
            self.left_inds = np.array([])
            self.right_inds = np.array([])
            self.center_inds = np.array([])
        else:
            try:
                from smal_model.smal_basics import align_smal_template_to_symmetry_axis
            except Exception as exc:  # pragma: no cover - needs the reference tree on sys.path
                raise NotImplementedError(
                    "config.ignore_sym=False needs the reference's smal_model.smal_basics on sys.path") from exc
            sym = dd.get("sym_verts")
            sym_file = None if getattr(cfg, "ignore_hardcoded_body", True) else cfg.SMAL_SYM_FILE
            v_sym, self.left_inds, self.right_inds, self.center_inds = align_smal_template_to_symmetry_axis(
                v_template.astype(np.float64), sym_file=sym_file, I=sym)
            v_template = np.asarray(v_sym, dtype=np.float32)

        # plain attributes, like the reference (they are absent from state_dict on purpose)
        self.f = dd["f"]
        self.faces = torch.from_numpy(self.f.astype(int)).to(device)
        self.size = [v_template.shape[0], 3]
        self.num_betas = lay["shapedirs"].shape[0]
        self.shapedirs = torch.from_numpy(lay["shapedirs"]).to(device)
        self.v_template = torch.from_numpy(np.ascontiguousarray(v_template)).to(device)
        self.J_regressor = torch.from_numpy(lay["J_regressor"]).to(device)
        if static:
            self.J = torch.from_numpy(lay["J"]).to(device)
        self.posedirs = torch.from_numpy(lay["posedirs"]).to(device)
        self.parents = lay["parents"]
        self.weights = torch.from_numpy(lay["weights"]).to(device)
        self.J_transformed = None

        self._device = device
        self._tables = packer.pack_model_cached(v_template, lay["shapedirs"], lay["posedirs"], lay["J_regressor"],
                                         lay["weights"], lay["parents"], lay["J"], source=cfg.SMAL_FILE)
        self._dm = DeviceModel(self._tables, device)  # raises on non-CUDA devices / missing library

    # the reference overrides __call__, not forward (smal_torch.py:198), so nn.Module hooks are bypassed
    def __call__(self, beta, theta, trans=None, del_v=None, betas_logscale=None, betas_trans=None, get_skin=True,
                 v_template=None, propagate_scaling=False):
        cfg = _config.active()
        nBetas = beta.shape[1]  # smal_torch.py:211
        NUM_JOINTS = self._tables["J"]

        # wrong-width pose is silently replaced by zeros (smal_torch.py:281-286)
        if theta.shape[1] != NUM_JOINTS:
            dim_x = beta.shape[0] if nBetas > 0 else 1
            theta = torch.zeros(dim_x, NUM_JOINTS, 3, device=beta.device)
        flags = 0
        if theta.dim() == 4:  # rotation matrices (smal_torch.py:288-289)
            flags |= _lib.F_THETA_IS_ROTMAT
        if getattr(cfg, "STATIC_JOINT_LOCATIONS", False):  # read per call (smal_torch.py:257,343)
            flags |= _lib.F_STATIC_JOINTS
        if propagate_scaling:
            flags |= _lib.F_PROPAGATE_SCALING
        if not get_skin:
            flags |= _lib.F_JOINTS_ONLY
        if not getattr(cfg, "ALLOW_LIMB_SCALING", True):  # batch_lbs.py:123-124
            betas_logscale = None
        if getattr(self, "_force_generic", False):
            flags |= _lib.F_FORCE_GENERIC
        legacy35 = NUM_JOINTS == 35 and not getattr(cfg, "ignore_hardcoded_body", True)  # smal_torch.py:353
        if legacy35 and not get_skin:
            flags &= ~_lib.F_JOINTS_ONLY  # the six extra joints are vertex picks

        if betas_trans is not None:
            betas_trans = betas_trans.to(beta.device)  # batch_lbs.py:148
        out_dtype = torch.float32
        tensors = [beta, theta, trans, del_v, betas_logscale, betas_trans, v_template]
        if any(t is not None and t.dtype == torch.float64 for t in tensors):
            out_dtype = torch.float64  # the reference would promote; compute stays fp32
        cast = [None if t is None else (t if t.dtype == torch.float32 or torch.is_autocast_enabled() else t.float())
                for t in tensors]
        beta_, theta_, trans_, del_v_, ls_, bt_, vt_ = cast

        verts, joints, Rs, v_shaped, Jt = SmalFunction.apply(self._dm, flags, vt_, beta_, theta_, trans_, del_v_, ls_, bt_)
        if flags & _lib.F_THETA_IS_ROTMAT:
            Rs = theta
        self.J_transformed = joints if (flags & _lib.F_STATIC_JOINTS) else Jt  # smal_torch.py:307,343-346

        if legacy35:
            joints = torch.cat([joints] + [verts[:, None, p] for p in _LEGACY_35_PICKS], dim=1)
        if out_dtype != torch.float32:
            verts, joints, v_shaped = verts.to(out_dtype), joints.to(out_dtype), v_shaped.to(out_dtype)
            if not (flags & _lib.F_THETA_IS_ROTMAT):
                Rs = Rs.to(out_dtype)
        if get_skin:
            return verts, joints, Rs, v_shaped
        return joints


def load_rig_any(path):
    """Unpickle an auxiliary SMAL data file (e.g. ``SMAL_DATA_FILE`` with ``cluster_means``)."""
    import pickle

    from .model_io import _RigUnpickler, _plain

    with open(path, "rb") as fh:
        raw = _RigUnpickler(fh, encoding="latin1").load()
    return {k: _plain(v) for k, v in raw.items()} if isinstance(raw, dict) else raw
