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

    @property
    def r(self):      # chumpy's accessor for the underlying data (smal_torch.py:55-58)
        return self.payload

    @property
    def data(self):   # the attribute name the reference's ChumpyWrapper uses
        return self.payload


class _RigUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module.startswith("chumpy"):
            return getattr(type(self), "ChumpyWrapper", _ChArray)
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
            dd = {k: z[k] for k in z.files if not k.startswith("pack/") and k not in ("flat_format", "pack_key")}
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
    validate_rig(dd, path)
    return dd


def validate_rig(dd: dict, path="<rig>") -> None:
    """Schema check of an SMPL-style rig (the keys ``SMAL.__init__`` reads, smal_torch.py:92-196, written by
    3D_model_prep/SMPL_exporter.py): shapes that must agree, a kinematic table whose parents precede their children
    (the chain indexes a growing list with it, batch_lbs.py:157,179), finite numbers, skinning weights that sum to one.
    Raises ValueError naming the file and the offending key."""

    def bad(msg):
        raise ValueError("rig file %s: %s" % (path, msg))

    vt = np.asarray(dd["v_template"])
    if vt.ndim != 2 or vt.shape[1] != 3 or vt.shape[0] == 0:
        bad("v_template has shape %s, expected (V, 3)" % (vt.shape,))
    V = vt.shape[0]
    sd = np.asarray(dd["shapedirs"])
    if sd.ndim != 3 or sd.shape[:2] != (V, 3):
        bad("shapedirs has shape %s, expected (%d, 3, nB)" % (sd.shape, V))
    kt = np.asarray(dd["kintree_table"])
    if kt.ndim != 2 or kt.shape[0] < 1:
        bad("kintree_table has shape %s, expected (2, J)" % (kt.shape,))
    J = kt.shape[1]
    par = kt[0].astype(np.int64)
    for i in range(1, J):
        if not (0 <= par[i] < i):
            bad("kintree_table[0][%d] = %d: a parent must precede its child (0 <= parent < index)" % (i, par[i]))
    W = np.asarray(dd["weights"])
    if W.shape != (V, J):
        bad("weights has shape %s, expected (%d, %d)" % (W.shape, V, J))
    jr = np.asarray(_plain(dd["J_regressor"]))
    if jr.shape != (J, V):
        bad("J_regressor has shape %s, expected (%d, %d)" % (jr.shape, J, V))
    pd = np.asarray(dd["posedirs"])
    if pd.size and (pd.ndim != 3 or pd.shape[:2] != (V, 3) or pd.shape[2] != 9 * (J - 1)):
        bad("posedirs has shape %s, expected (%d, 3, %d) or empty" % (pd.shape, V, 9 * (J - 1)))
    f = np.asarray(dd["f"])
    if f.ndim != 2 or f.shape[1] != 3 or (f.size and (f.min() < 0 or f.max() >= V)):
        bad("f (faces) has shape %s / indices outside [0, %d)" % (f.shape, V))
    if bool(dd.get("static_joint_locs", False)):
        if "J" not in dd or np.asarray(dd["J"]).shape != (J, 3):
            bad("static_joint_locs is set but J is missing or not (%d, 3)" % J)
    for k in ("scaledirs", "transdirs"):
        if k in dd and np.asarray(dd[k]).shape[1:] != (J, 3):
            bad("%s has shape %s, expected (C, %d, 3)" % (k, np.asarray(dd[k]).shape, J))
    for k, a in (("v_template", vt), ("shapedirs", sd), ("weights", W), ("J_regressor", jr)):
        if not np.isfinite(a).all():
            bad("%s holds non-finite values" % k)
    rows = W.sum(axis=1)
    if W.min() < 0 or np.abs(rows - 1.0).max() > 1e-3:
        bad("weights: rows must be non-negative and sum to 1 (row sums in [%g, %g])" % (rows.min(), rows.max()))


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


FLAT_FORMAT = 1
_FLAT_SKIP = ("J_names",)


def export_flat(src_path, dst_path, static_joints=None) -> str:
    """write a rig as ONE flat, pickle-free ``.npz`` -- the validated arrays of the SMPL-style
    schema (``f, v_template, shapedirs, posedirs, J_regressor, J, weights, kintree_table, J_names, static_joint_locs`` and
    whatever else the source holds as plain arrays: ``scaledirs, transdirs, shape_cov, shape_mean_betas, sym_verts``) plus
    the PACKED device tables of the kernels (``pack/<name>``), so that a later ``SMAL`` construction neither unpickles
    chumpy objects nor spends seconds in the packer.  ``config.SMAL_FILE`` may point at the result: :func:`load_rig` reads
    ``.npz`` files with ``allow_pickle=False`` and ``packer.pack_model_cached`` is pre-seeded from the embedded tables
    (:func:`embedded_pack`).  ``static_joints`` defaults to the file's own ``static_joint_locs`` flag."""
    from . import packer

    dd = load_rig(src_path)
    static = bool(dd.get("static_joint_locs", False)) if static_joints is None else bool(static_joints)
    lay = reference_layout(dd, static_joints=static)
    tables = packer.pack_model(lay["v_template"], lay["shapedirs"], lay["posedirs"], lay["J_regressor"], lay["weights"],
                               lay["parents"], lay["J"])
    out = {"flat_format": np.asarray(FLAT_FORMAT), "pack_key": np.asarray(packer.pack_key(
        lay["v_template"], lay["shapedirs"], lay["posedirs"], lay["J_regressor"], lay["weights"], lay["parents"], lay["J"]))}
    for k, v in dd.items():
        if k == "J_names":
            out[k] = np.asarray([str(x) for x in v])
        elif k == "static_joint_locs":
            out[k] = np.asarray(bool(v))
        else:
            a = np.asarray(v)
            if a.dtype == object:      # nothing that needs pickle goes into the flat file
                continue
            out[k] = a
    for k, v in tables.items():
        out["pack/" + k] = np.asarray(v)
    np.savez(dst_path, **out)
    return str(dst_path) if str(dst_path).endswith(".npz") else str(dst_path) + ".npz"


def embedded_pack(path, key):
    """The packed tables embedded in a flat rig file, if they were made for exactly this rig layout and packer version
    (``key`` = ``packer.pack_key`` of the arrays about to be packed); else None."""
    path = str(path)
    if not path.endswith(".npz"):
        return None
    try:
        with np.load(path, allow_pickle=False) as z:
            if "pack_key" not in z.files or str(z["pack_key"]) != key:
                return None
            return {k[5:]: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("pack/")}
    except Exception:
        return None
