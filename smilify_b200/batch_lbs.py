"""CUDA-backed drop-ins for ``smal_model.batch_lbs`` (/root/reference/smal_model/batch_lbs.py).

Same names, argument meaning and return shapes as the reference module; the arithmetic runs in
the kernels behind ``smil_rodrigues_*`` / ``smil_chain_*`` (include/smil_b200.h)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from . import config as _config
from .functions import ChainFunction, RodriguesFunction, tree_only_model

_TREE_CACHE = {}


def batch_skew(vec, batch_size=None, opts=None):
    """(N,3) -> (N,3,3) skew-symmetric matrices (batch_lbs.py:10-28)."""
    if batch_size is None:
        batch_size = vec.shape[0]
    z = torch.zeros_like(vec[:, 0])
    return torch.stack([z, -vec[:, 2], vec[:, 1], vec[:, 2], z, -vec[:, 0], -vec[:, 1], vec[:, 0], z], dim=1).reshape(
        batch_size, 3, 3)


def batch_rodrigues(theta, opts=None):
    """Theta is (N,3); returns (N,3,3) (batch_lbs.py:31-50)."""
    return RodriguesFunction.apply(theta)


def batch_global_rigid_transformation(Rs, Js, parent, rotate_base=False, betas_logscale=None, betas_trans=None,
                                      propagate_scaling=False, opts=None, num_joints=35):
    """Returns ``(new_J (N,J,3), A (N,J,4,4))`` exactly like batch_lbs.py:75-197.

    ``config.ALLOW_LIMB_SCALING`` is read at call time like the reference does (batch_lbs.py:123)."""
    cfg = _config.active()
    if not getattr(cfg, "ALLOW_LIMB_SCALING", True):
        betas_logscale = None
    parent = np.asarray(parent)
    key = (parent.astype(np.int64).tobytes(), str(Rs.device))
    dm = _TREE_CACHE.get(key)
    if dm is None:
        dm = _TREE_CACHE[key] = tree_only_model(parent.astype(np.int32), Rs.device)
    if rotate_base:  # batch_lbs.py:103-112
        print("Flipping the SMPL coordinate frame!!!!")
        rot_x = torch.tensor([[1, 0, 0], [0, -1, 0], [0, 0, -1]], dtype=Rs.dtype, device=Rs.device)
        Rs = torch.cat([torch.matmul(Rs[:, :1], rot_x), Rs[:, 1:]], dim=1)
    flags = _lib.F_PROPAGATE_SCALING if propagate_scaling else 0
    if betas_trans is not None:
        betas_trans = betas_trans.to(Rs.device)
    new_J, A34 = ChainFunction.apply(dm, flags, Rs, Js, betas_logscale, betas_trans)
    N, J = A34.shape[:2]
    bottom = torch.zeros((N, J, 1, 4), dtype=A34.dtype, device=A34.device)
    bottom[..., 3] = 1.0
    return new_J, torch.cat([A34, bottom], dim=2)
