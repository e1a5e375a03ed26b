"""Stand-in for the reference's module-level ``config`` (/root/reference/config.py).

The reference's ``SMAL`` reads ``config.SMAL_FILE`` (+ symmetry flags) at construction and
``config.STATIC_JOINT_LOCATIONS / ALLOW_LIMB_SCALING / ignore_hardcoded_body / DEBUG`` at every
call (smal_torch.py:92,257,343,353; batch_lbs.py:123).  When this package is used inside the
reference tree, ``import config`` resolves to the reference's own module and that one is used
(see :func:`active`).  Stand-alone (tests, bench) this module provides the same attribute names."""
from __future__ import annotations

import sys
import types

SMAL_FILE = None
SMAL_DATA_FILE = None
SMAL_SYM_FILE = None
ignore_sym = True
ignore_hardcoded_body = True
STATIC_JOINT_LOCATIONS = False
ALLOW_LIMB_SCALING = True
DEBUG = False
CANONICAL_MODEL_JOINTS = None
N_POSE = None
N_BETAS = None
dd = None


def active() -> types.ModuleType:
    """The config module in force: the reference's ``config`` if one has been imported,
    otherwise this module."""
    mod = sys.modules.get("config")
    if mod is not None and hasattr(mod, "SMAL_FILE"):
        return mod
    return sys.modules[__name__]


def configure(smal_file, **overrides):
    """Stand-alone equivalent of ``apply_smal_file_override``
    (/root/reference/smal_fitter/neuralSMIL/configs/config_utils.py:238-282): point the fallback
    config at a rig file and re-derive ``dd, N_POSE, N_BETAS, STATIC_JOINT_LOCATIONS,
    CANONICAL_MODEL_JOINTS`` from it."""
    from .model_io import load_rig

    me = sys.modules[__name__]
    me.SMAL_FILE = str(smal_file)
    rig = load_rig(smal_file)
    me.dd = rig
    me.STATIC_JOINT_LOCATIONS = bool(rig.get("static_joint_locs", False))
    names = rig.get("J_names")
    nj = len(names) if names is not None else rig["kintree_table"].shape[1]
    me.N_POSE = nj - 1
    me.N_BETAS = rig["shapedirs"].shape[-1]
    me.CANONICAL_MODEL_JOINTS = list(range(nj))
    for k, v in overrides.items():
        setattr(me, k, v)
    return me
