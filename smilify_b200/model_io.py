"""Reading SMPL-style rig files: the ``.pkl`` schema written by SMILify's Blender exporter
(/root/reference/3D_model_prep/SMPL_exporter.py, read at smal_model/smal_torch.py:21-73,92-196
and config.py:76-140) and the compact ``.npz`` fixtures of this repository.

Keys: ``f, v_template, shapedirs, posedirs, J_regressor, J, weights, kintree_table, J_names`` and
optionally ``static_joint_locs, scaledirs, transdirs, shape_cov, shape_mean_betas, sym_verts``.
Legacy SMAL pickles hold ``chumpy`` arrays and a scipy-sparse ``J_regressor``; both are converted
to plain numpy without importing chumpy."""
from __future__ import annotations

import pickle

import numpy as np

REQUIRED_KEYS = ("f", "v_template", "shapedirs", "posedirs", "J_regressor", "weights", "kintree_table")


class _ChArray:
    """Inert stand-in for ``chumpy.ch.Ch``: keeps only the numeric payload."""

    def __init__(self, *args, **kwargs):
        self.payload = np.array(args[0]) if args else np.array([])

    def __setstate__(self, state):
        if isinstance(state, dict):
            self.payload = np.array(state.get("x", []))
        elif isinstance(state, (tuple, list)):
            self.payload = np.array(state[0])
        else:
            self.payload = np.array(state)

    def __array__(self, dtype=None, copy=None):
        return np.asarray(self.payload, dtype=dtype)


class _RigUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module.startswith("chumpy"):
            return _ChArray
        return super().find_class(module, name)


def _plain(value):
    if isinstance(value, _ChArray):
        return np.asarray(value.payload)
    if hasattr(value, "todense"):  # scipy sparse J_regressor (smal_torch.py:169-172)
        return np.asarray(value.todense())
    return value


def load_rig(path) -> dict:
    """Return the rig as a dict of numpy arrays / python values; raises KeyError on a missing key."""
    path = str(path)
    if path.endswith(".npz"):
        with np.load(path, allow_pickle=False) as z:
            dd = {k: z[k] for k in z.files}
        if "static_joint_locs" in dd:
            dd["static_joint_locs"] = bool(dd["static_joint_locs"])
        if "J_names" in dd:
            dd["J_names"] = [str(s) for s in dd["J_names"]]
    else:
        with open(path, "rb") as fh:
            raw = _RigUnpickler(fh, encoding="latin1").load()
        dd = {k: _plain(v) for k, v in raw.items()}
    for k in REQUIRED_KEYS:
        if k not in dd:
            raise KeyError("rig file %s lacks key %r" % (path, k))
    return dd


def reference_layout(dd: dict, static_joints: bool, shape_family_betas=None) -> dict:
    """The arrays ``SMAL.__init__`` derives (smal_torch.py:120-196), as float32/int32 numpy."""

    def f32(a):
        return np.ascontiguousarray(np.asarray(a), dtype=np.float32)

    v_template = np.asarray(dd["v_template"], dtype=np.float64)
    V = v_template.shape[0]
    nB = np.asarray(dd["shapedirs"]).shape[-1]
    shapedirs = np.reshape(np.asarray(dd["shapedirs"]), [-1, nB]).T.copy()          # (nB, 3V) :124
    if shape_family_betas is not None:                                               # :133-140
        v_template = v_template + np.matmul(np.asarray(shape_family_betas)[None, :], shapedirs).reshape(V, 3)
    jreg = np.asarray(_plain(dd["J_regressor"]))                                     # (J, V)
    J = jreg.shape[0]
    pd = np.asarray(dd["posedirs"])
    if pd.size != 0:                                                                 # :179-190
        posedirs = np.reshape(pd, [-1, pd.shape[-1]]).T
    else:
        posedirs = np.zeros(((J - 1) * 9, V * 3))
    return dict(
        faces=np.asarray(dd["f"]),
        v_template=f32(v_template), shapedirs=f32(shapedirs), posedirs=f32(posedirs), J_regressor=f32(jreg.T),
        weights=f32(dd["weights"]), parents=np.asarray(dd["kintree_table"])[0].astype(np.int32),   # :193
        J=f32(dd["J"]) if (static_joints and "J" in dd) else None,
    )
