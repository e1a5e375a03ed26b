"""ctypes binding of ``libsmil_b200.so`` (the C ABI in ``include/smil_b200.h``).

There is no CPU fallback: if the shared library is missing or a call fails, a RuntimeError is
raised with the library's own message."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsmil_b200.so")
ABI_VERSION = 7

c_f = C.POINTER(C.c_float)
c_i = C.POINTER(C.c_int32)


class SmilTuning(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("fwd_gp", "fwd_topsync", "fwd_cps", "bwd_stages", "bwd_cps", "bwd_vm", "pose_v2",
                                          "pose_lean", "bwd_atomic", "bwd_slot_reduce", "bwd_stagger_ns")]


class SmilModelDesc(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("num_verts", C.c_int32), ("num_joints", C.c_int32), ("num_betas", C.c_int32),
        ("num_pose_basis", C.c_int32), ("skin_k", C.c_int32), ("jreg_nnz", C.c_int32), ("num_levels", C.c_int32),
        ("max_width", C.c_int32), ("num_pairs", C.c_int32), ("has_static", C.c_int32),
        ("fwd_chunk", C.c_int32), ("fwd_nchunk", C.c_int32), ("v_pad", C.c_int32),
        ("bwd_chunk", C.c_int32), ("bwd_ntile", C.c_int32), ("bwd_warps", C.c_int32), ("bwd_rec", C.c_int32),
        ("num_units", C.c_int32), ("num_item_pairs", C.c_int32), ("bwd_max_tile_pairs", C.c_int32), ("fused_ok", C.c_int32),
        ("v_template", c_f), ("shapedirs", c_f), ("posedirs", c_f), ("J_static", c_f),
        ("skin_idx", c_i), ("skin_w", c_f),
        ("jc_ptr", c_i), ("jc_vid", c_i), ("jc_val", c_f),
        ("jr_ptr", c_i), ("jr_jid", c_i), ("jr_val", c_f),
        ("parents", c_i), ("order", c_i), ("level_ptr", c_i), ("child_ptr", c_i), ("child_idx", c_i),
        ("JT", c_f), ("JS", c_f), ("jreg_rowsum", c_f),
        ("pj_ptr", c_i), ("pj_bone", c_i), ("pj_joint", c_i), ("pj_c0", c_f), ("pj_cm", c_f),
        ("pb_ptr", c_i), ("pb_pair", c_i), ("pbrec", c_f),
        ("vconst", c_f), ("ell_idx", c_i), ("ell_w", c_f), ("fwd_pos", c_i),
        ("rec2", c_f), ("unit_bone", c_i), ("unit_ptr", c_i), ("unit_tile", c_i), ("tile_unit_ptr", c_i),
        ("warp_unit_ptr", c_i), ("warp_unit", c_i), ("tjv_ptr", c_i), ("tjv_v", c_i),
        ("sb_ptr", c_i), ("unit_slot", c_i),
        ("vm_ok", C.c_int32), ("vm_chunk", C.c_int32), ("vm_ntile", C.c_int32), ("vm_recf", C.c_int32),
        ("vm_nslots", C.c_int32), ("vm_nseg", C.c_int32), ("vm_max_tile_recs", C.c_int32), ("vm_max_tile_segs", C.c_int32),
        ("vm_warps", C.c_int32), ("vm_nphase", C.c_int32),
        ("vm_rec", c_f), ("vm_seg", c_i), ("vm_tile_rec_ptr", c_i), ("vm_warp_seg_ptr", c_i), ("vm_sb_ptr", c_i),
        ("vm_slot_bone", c_i),
        ("tuning", SmilTuning),
    ]


class SmilModelInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "num_verts", "num_joints", "num_betas", "num_pose_basis", "skin_k", "jreg_nnz", "tree_depth",
        "max_level_width", "posedirs_zero", "joint_bone_pairs", "num_chunks", "num_ga_slots")]


class SmilProfileEntry(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("launches", C.c_int32), ("total_ms", C.c_float)]


class SmilFwdArgs(C.Structure):
    _fields_ = [
        ("batch", C.c_int32), ("num_betas_used", C.c_int32), ("flags", C.c_uint32), ("_pad", C.c_int32),
        ("beta", C.c_void_p), ("beta_stride", C.c_int64),
        ("theta", C.c_void_p), ("theta_stride", C.c_int64),
        ("trans", C.c_void_p), ("trans_stride", C.c_int64),
        ("betas_logscale", C.c_void_p), ("logscale_stride", C.c_int64),
        ("betas_trans", C.c_void_p), ("btrans_stride", C.c_int64),
        ("del_v", C.c_void_p), ("del_v_stride", C.c_int64),
        ("v_template", C.c_void_p),
        ("verts", C.c_void_p), ("joints", C.c_void_p), ("Rs", C.c_void_p), ("v_shaped", C.c_void_p),
        ("J_transformed", C.c_void_p),
        ("saved", C.c_void_p), ("saved_bytes", C.c_size_t),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


class SmilBwdArgs(C.Structure):
    _fields_ = [
        ("fwd", SmilFwdArgs),
        ("g_verts", C.c_void_p), ("g_joints", C.c_void_p), ("g_Rs", C.c_void_p), ("g_v_shaped", C.c_void_p),
        ("g_beta", C.c_void_p), ("g_theta", C.c_void_p), ("g_trans", C.c_void_p), ("g_logscale", C.c_void_p),
        ("g_btrans", C.c_void_p), ("g_del_v", C.c_void_p),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


class SmilChainArgs(C.Structure):
    _fields_ = [
        ("batch", C.c_int32), ("flags", C.c_uint32),
        ("Rs", C.c_void_p), ("Js", C.c_void_p), ("betas_logscale", C.c_void_p), ("betas_trans", C.c_void_p),
        ("new_J", C.c_void_p), ("A", C.c_void_p),
        ("g_new_J", C.c_void_p), ("g_A", C.c_void_p),
        ("g_Rs", C.c_void_p), ("g_Js", C.c_void_p), ("g_logscale", C.c_void_p), ("g_btrans", C.c_void_p),
    ]


class SmilProjArgs(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("batch", C.c_int32), ("num_views", C.c_int32), ("num_joints", C.c_int32),
        ("image_size", C.c_float), ("norm_size", C.c_float), ("clamp_lo", C.c_float), ("clamp_hi", C.c_float),
        ("weight", C.c_float), ("_pad", C.c_int32),
        ("joints", C.c_void_p), ("joints_stride", C.c_int64),
        ("cam_R", C.c_void_p), ("R_bstride", C.c_int64), ("R_vstride", C.c_int64),
        ("cam_T", C.c_void_p), ("T_bstride", C.c_int64), ("T_vstride", C.c_int64),
        ("fov_deg", C.c_void_p), ("fov_bstride", C.c_int64), ("fov_vstride", C.c_int64),
        ("aspect", C.c_void_p), ("asp_bstride", C.c_int64), ("asp_vstride", C.c_int64),
        ("target", C.c_void_p), ("visibility", C.c_void_p), ("view_mask", C.c_void_p),
        ("joint_weights", C.c_void_p), ("sample_mask", C.c_void_p),
        ("proj", C.c_void_p), ("loss", C.c_void_p), ("partials", C.c_void_p),
        ("g_loss", C.c_void_p), ("g_joints", C.c_void_p), ("g_cam_R", C.c_void_p), ("g_cam_T", C.c_void_p),
        ("g_fov", C.c_void_p),
    ]


class SmilKp3dArgs(C.Structure):
    _fields_ = [
        ("batch", C.c_int32), ("num_joints", C.c_int32),
        ("joints", C.c_void_p), ("target", C.c_void_p), ("visibility", C.c_void_p), ("joint_weights", C.c_void_p),
        ("sample_mask", C.c_void_p), ("loss", C.c_void_p), ("partials", C.c_void_p),
        ("g_loss", C.c_void_p), ("g_joints", C.c_void_p),
    ]


class SmilFitterPriorArgs(C.Structure):
    _fields_ = [
        ("num_frames", C.c_int32), ("num_joints", C.c_int32), ("num_betas", C.c_int32), ("window", C.c_int32),
        ("frame_offset", C.c_int64), ("total_frames", C.c_int64),
        ("global_rot", C.c_void_p), ("joint_rot", C.c_void_p), ("trans", C.c_void_p), ("betas", C.c_void_p),
        ("halo_prev", C.c_void_p), ("halo_next", C.c_void_p), ("max_limits", C.c_void_p), ("min_limits", C.c_void_p),
        ("rotation_mask", C.c_void_p), ("global_mask", C.c_void_p), ("pose_mean", C.c_void_p), ("pose_precs", C.c_void_p),
        ("pose_use", C.c_void_p), ("betas_mean", C.c_void_p), ("betas_prec", C.c_void_p),
        ("w_limit", C.c_float), ("w_pose", C.c_float), ("w_splay", C.c_float), ("w_betas", C.c_float),
        ("w_temp", C.c_float), ("g_scale", C.c_float), ("accumulate", C.c_int32),
        ("loss", C.c_void_p), ("partials", C.c_void_p), ("g_global_rot", C.c_void_p), ("g_joint_rot", C.c_void_p),
        ("g_trans", C.c_void_p), ("g_betas", C.c_void_p),
    ]


ADAM_MAX_TENSORS = 8


class SmilAdamArgs(C.Structure):
    _fields_ = [
        ("num_tensors", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("param", C.c_void_p * ADAM_MAX_TENSORS), ("grad", C.c_void_p * ADAM_MAX_TENSORS),
        ("exp_avg", C.c_void_p * ADAM_MAX_TENSORS), ("exp_avg_sq", C.c_void_p * ADAM_MAX_TENSORS),
        ("numel", C.c_int64 * ADAM_MAX_TENSORS), ("lr", C.c_float * ADAM_MAX_TENSORS), ("step", C.c_void_p),
    ]


class SmilTriArgs(C.Structure):
    _fields_ = [
        ("batch", C.c_int32), ("num_views", C.c_int32), ("num_joints", C.c_int32), ("image_size", C.c_float),
        ("keypoints", C.c_void_p), ("visibility", C.c_void_p), ("view_mask", C.c_void_p),
        ("cam_R", C.c_void_p), ("R_bstride", C.c_int64), ("R_vstride", C.c_int64),
        ("cam_T", C.c_void_p), ("T_bstride", C.c_int64), ("T_vstride", C.c_int64),
        ("fov_deg", C.c_void_p), ("fov_bstride", C.c_int64), ("fov_vstride", C.c_int64),
        ("aspect", C.c_void_p), ("asp_bstride", C.c_int64), ("asp_vstride", C.c_int64),
        ("tri", C.c_void_p), ("valid", C.c_void_p),
        ("g_tri", C.c_void_p), ("g_cam_R", C.c_void_p), ("g_cam_T", C.c_void_p), ("g_fov", C.c_void_p),
    ]


F_THETA_IS_ROTMAT = 1 << 0
F_STATIC_JOINTS = 1 << 1
F_PROPAGATE_SCALING = 1 << 2
F_JOINTS_ONLY = 1 << 3
F_FORCE_GENERIC = 1 << 4
LOSS_FITTER, LOSS_MULTIVIEW, LOSS_SINGLEVIEW = 0, 1, 2

# every symbol include/smil_b200.h declares
EXPORTS = (
    "smil_abi_version", "smil_last_error", "smil_launch_count", "smil_profile_enable", "smil_profile_collect",
    "smil_model_create", "smil_model_destroy", "smil_model_info",
    "smil_saved_bytes", "smil_fwd_scratch_bytes", "smil_bwd_scratch_bytes", "smil_forward", "smil_backward",
    "smil_rodrigues_forward", "smil_rodrigues_backward", "smil_chain_forward", "smil_chain_backward",
    "smil_project_loss_forward", "smil_project_loss_backward", "smil_keypoint3d_loss_forward",
    "smil_keypoint3d_loss_backward", "smil_triangulate_dlt_forward", "smil_triangulate_dlt_backward",
    "smil_rot6d_to_rotmat_forward", "smil_rot6d_to_rotmat_backward", "smil_keypoint3d_loss_forward_static",
    "smil_fitter_prior_terms", "smil_adam_step",
)
KP3D_STATIC_MAX_BATCH = 16384   # SMIL_KP3D_STATIC_MAX_BATCH

_lib = None


def lib():
    """The loaded library; raises loudly when it is missing (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "smilify_b200: %s is missing -- build it with `python -m smilify_b200.build` "
            "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.smil_abi_version.restype = C.c_int
    L.smil_last_error.restype = C.c_char_p
    L.smil_launch_count.restype = C.c_longlong
    L.smil_profile_enable.argtypes = [C.c_int]
    L.smil_profile_collect.argtypes = [C.POINTER(SmilProfileEntry), C.c_int, C.POINTER(C.c_int)]
    L.smil_model_create.argtypes = [C.POINTER(SmilModelDesc), C.POINTER(C.c_void_p)]
    L.smil_model_destroy.argtypes = [C.c_void_p]
    L.smil_model_destroy.restype = None
    L.smil_model_info.argtypes = [C.c_void_p, C.POINTER(SmilModelInfo)]
    L.smil_saved_bytes.argtypes = [C.c_void_p, C.c_int32, C.c_uint32]
    L.smil_saved_bytes.restype = C.c_size_t
    L.smil_fwd_scratch_bytes.argtypes = [C.c_void_p, C.POINTER(SmilFwdArgs)]
    L.smil_fwd_scratch_bytes.restype = C.c_size_t
    L.smil_bwd_scratch_bytes.argtypes = [C.c_void_p, C.POINTER(SmilBwdArgs)]
    L.smil_bwd_scratch_bytes.restype = C.c_size_t
    L.smil_forward.argtypes = [C.c_void_p, C.POINTER(SmilFwdArgs), C.c_void_p]
    L.smil_backward.argtypes = [C.c_void_p, C.POINTER(SmilBwdArgs), C.c_void_p]
    L.smil_rodrigues_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.smil_rodrigues_backward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    L.smil_chain_forward.argtypes = [C.c_void_p, C.POINTER(SmilChainArgs), C.c_void_p]
    L.smil_chain_backward.argtypes = [C.c_void_p, C.POINTER(SmilChainArgs), C.c_void_p]
    L.smil_project_loss_forward.argtypes = [C.POINTER(SmilProjArgs), C.c_void_p]
    L.smil_project_loss_backward.argtypes = [C.POINTER(SmilProjArgs), C.c_void_p]
    L.smil_keypoint3d_loss_forward.argtypes = [C.POINTER(SmilKp3dArgs), C.c_void_p]
    L.smil_keypoint3d_loss_backward.argtypes = [C.POINTER(SmilKp3dArgs), C.c_void_p]
    L.smil_keypoint3d_loss_forward_static.argtypes = [C.POINTER(SmilKp3dArgs), C.c_void_p]
    L.smil_fitter_prior_terms.argtypes = [C.POINTER(SmilFitterPriorArgs), C.c_void_p]
    L.smil_adam_step.argtypes = [C.POINTER(SmilAdamArgs), C.c_void_p]
    L.smil_triangulate_dlt_forward.argtypes = [C.POINTER(SmilTriArgs), C.c_void_p]
    L.smil_triangulate_dlt_backward.argtypes = [C.POINTER(SmilTriArgs), C.c_void_p]
    L.smil_rot6d_to_rotmat_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
    L.smil_rot6d_to_rotmat_backward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
    for name in EXPORTS:
        getattr(L, name)  # AttributeError here = the .so is older than include/smil_b200.h
    if L.smil_abi_version() != ABI_VERSION:
        raise RuntimeError("smilify_b200: ABI mismatch (library %d, binding %d)" % (L.smil_abi_version(), ABI_VERSION))
    _lib = L
    return L


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().smil_last_error()
        raise RuntimeError("%s failed (%d): %s" % (what, rc, msg.decode("utf-8", "replace") if msg else "?"))


def launch_count() -> int:
    return int(lib().smil_launch_count())


def profile_enable(max_launches: int):
    check(lib().smil_profile_enable(int(max_launches)), "smil_profile_enable")


def profile_collect() -> dict:
    """-> {kernel name: (launches, total device ms)} since the previous collect."""
    buf = (SmilProfileEntry * 64)()
    n = C.c_int(0)
    check(lib().smil_profile_collect(buf, 64, C.byref(n)), "smil_profile_collect")
    return {buf[i].name.decode(): (int(buf[i].launches), float(buf[i].total_ms)) for i in range(n.value)}


def _np_ptr(a: np.ndarray, ctype):
    if a is None or a.size == 0:
        return C.cast(None, C.POINTER(ctype))
    return a.ctypes.data_as(C.POINTER(ctype))


_DESC_F32 = ("v_template", "shapedirs", "posedirs", "J_static", "skin_w", "jc_val", "jr_val", "JT", "JS", "jreg_rowsum",
             "pj_c0", "pj_cm", "pbrec", "vconst", "ell_w", "rec2", "vm_rec")
_DESC_I32 = ("skin_idx", "jc_ptr", "jc_vid", "jr_ptr", "jr_jid", "parents", "order", "level_ptr", "child_ptr",
             "child_idx", "pj_ptr", "pj_bone", "pj_joint", "pb_ptr", "pb_pair", "ell_idx", "fwd_pos", "unit_bone", "unit_ptr",
             "unit_tile", "tile_unit_ptr", "warp_unit_ptr", "warp_unit", "tjv_ptr", "tjv_v", "sb_ptr", "unit_slot",
             "vm_seg", "vm_tile_rec_ptr", "vm_warp_seg_ptr", "vm_sb_ptr", "vm_slot_bone")


def tuning_from_env(overrides=None) -> dict:
    """Per-model tuning switches (SmilTuning): ``SMIL_*`` environment variables read at MODEL CREATION (development
    sweeps), overridden by ``overrides``.  Integers pass through; for the on/off switches "0" means off."""
    out = {}
    for name in ("fwd_gp", "fwd_cps", "bwd_stages", "bwd_cps", "bwd_stagger_ns"):
        v = os.environ.get("SMIL_" + name.upper())
        out[name] = int(v) if v else 0
    v = os.environ.get("SMIL_FWD_TOPSYNC")
    out["fwd_topsync"] = 0 if v in (None, "") else (1 if int(v) else 2)
    for name in ("bwd_vm", "pose_lean", "bwd_slot_reduce", "bwd_atomic"):
        v = os.environ.get("SMIL_" + name.upper())
        out[name] = 0 if v in (None, "") else (0 if int(v) else 2)
    v = os.environ.get("SMIL_POSE_V2")
    out["pose_v2"] = 1 if v not in (None, "") and int(v) else 0   # 1: several samples per warp (pose_kernels2.cuh)
    if os.environ.get("SMIL_POSE_LS") == "0":
        out["pose_v2"] = 2                                         # 2: warp per sample (pose_kernels.cuh) everywhere
    out.update(overrides or {})
    return out


def desc_from_tables(t: dict, device: int):
    """Fill a SmilModelDesc from packer tables.  Returns (desc, keepalive list)."""
    d = SmilModelDesc()
    for k, v in tuning_from_env(t.get("tuning")).items():
        setattr(d.tuning, k, int(v))
    keep = []
    d.device = device
    d.num_verts, d.num_joints, d.num_betas, d.num_pose_basis = t.get("V", 0), t["J"], t.get("nB", 0), t.get("P", 0)
    d.skin_k, d.jreg_nnz = t.get("K", 0), t.get("jreg_nnz", 0)
    d.num_levels, d.max_width, d.num_pairs = t["num_levels"], t["max_width"], t.get("npairs", 0)
    js = t.get("J_static")
    d.has_static = int(js is not None and js.size > 0)
    d.fwd_chunk, d.fwd_nchunk, d.v_pad = t.get("fwd_chunk", 0), t.get("fwd_nchunk", 0), t.get("Vpad", 0)
    d.bwd_chunk, d.bwd_ntile, d.bwd_warps, d.bwd_rec = (t.get("bwd_chunk", 0), t.get("bwd_ntile", 0),
                                                        t.get("bwd_warps", 0), t.get("bwd_rec2", 0))
    d.num_units, d.num_item_pairs = t.get("nunits", 0), t.get("npair_items", 0)
    d.bwd_max_tile_pairs = t.get("bwd_max_tile_pairs", 0)
    d.fused_ok = int(t.get("fused_ok", False))
    d.vm_ok = int(t.get("vm_ok", False))
    for name in ("vm_chunk", "vm_ntile", "vm_recf", "vm_nslots", "vm_nseg", "vm_max_tile_recs", "vm_max_tile_segs", "vm_nphase", "vm_warps"):
        setattr(d, name, int(t.get(name, 0)))
    for name in _DESC_F32:
        a = t.get(name)
        a = np.ascontiguousarray(a, dtype=np.float32) if a is not None else None
        keep.append(a)
        setattr(d, name, _np_ptr(a, C.c_float))
    for name in _DESC_I32:
        a = t.get("unit_pptr" if name == "unit_ptr" else name)  # the kernels index item PAIRS
        a = np.ascontiguousarray(a, dtype=np.int32) if a is not None else None
        keep.append(a)
        setattr(d, name, _np_ptr(a, C.c_int32))
    return d, keep
