"""DLT triangulation of joints from multi-view 2-D keypoints and 6-D rotations on the CUDA path
(``smil_triangulate_dlt_*`` / ``smil_rot6d_to_rotmat_*`` in include/smil_b200.h).

Replaces (paths relative to /root/reference):

* ``triangulate_joints_dlt``   -- ``MultiViewSMILImageRegressor._triangulate_joints_dlt``
  (smal_fitter/neuralSMIL/multiview_smil_regressor.py:1707-1838): GT 2-D keypoints + PREDICTED cameras ->
  world-space joints; differentiable with respect to the cameras (``cam_R``, ``cam_T``, ``fov``), which is what the
  reference uses it for (tests/test_triangulation_consistency.py:455-541);
* ``rotation_6d_to_matrix``    -- ``columns=True``: the reference's own ``_rotation_6d_to_matrix`` (cameras,
  multiview_smil_regressor.py:250-272); ``columns=False``: ``pytorch3d.transforms.rotation_6d_to_matrix``, the first
  half of ``rotation_6d_to_axis_angle`` (smil_image_regressor.py:87-96).  Its output can be passed to ``SMAL`` as a
  4-D ``theta`` (rotation matrices, smal_torch.py:288-289), which skips the reference's matrix -> axis-angle ->
  Rodrigues round trip.

CPU tensors raise: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from . import functions as _fn
from .functions import _check_cuda_f32, _stream
from .losses import _cams, _f, _inner_contig


class TriangulateFunction(torch.autograd.Function):
    """(cam_R, cam_T, fov) -> (tri, valid); keypoints / visibility / aspect / view mask are data."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, image_size, keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask):
        dev = keypoints.device
        for n, t in (("keypoints_2d", keypoints), ("visibility", visibility), ("cam_R", cam_R), ("cam_T", cam_T),
                     ("fov", fov), ("aspect", aspect), ("view_mask", view_mask)):
            _check_cuda_f32(n, t, dev)
        B, V, J = keypoints.shape[0], keypoints.shape[1], keypoints.shape[2]
        if tuple(keypoints.shape) != (B, V, J, 2) or tuple(visibility.shape) != (B, V, J):
            raise RuntimeError("keypoints_2d must be (B,V,J,2) and visibility (B,V,J); got %s %s" % (
                tuple(keypoints.shape), tuple(visibility.shape)))
        if tuple(cam_R.shape) != (B, V, 3, 3) or tuple(cam_T.shape) != (B, V, 3) or tuple(fov.shape) != (B, V):
            raise RuntimeError("cameras must be expanded to (B,V,3,3) / (B,V,3) / (B,V); got %s %s %s" % (
                tuple(cam_R.shape), tuple(cam_T.shape), tuple(fov.shape)))
        keypoints, visibility = keypoints.contiguous(), visibility.contiguous()
        cam_R, cam_T = _inner_contig(cam_R, 2), _inner_contig(cam_T, 1)
        view_mask = view_mask.contiguous() if view_mask is not None else None
        tri = _fn._empty((B, J, 3), device=dev, dtype=torch.float32)
        valid = _fn._empty((B, J), device=dev, dtype=torch.float32)
        a = TriangulateFunction._args(image_size, keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask)
        a.tri, a.valid = tri.data_ptr(), valid.data_ptr()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().smil_triangulate_dlt_forward(C.byref(a), _stream()), "smil_triangulate_dlt_forward")
        ctx.image_size = image_size
        ctx.save_for_backward(keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask)
        ctx.mark_non_differentiable(valid)
        ctx.set_materialize_grads(False)
        return tri, valid

    @staticmethod
    def _args(image_size, keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask):
        a = _lib.SmilTriArgs()
        a.batch, a.num_views, a.num_joints = keypoints.shape[0], keypoints.shape[1], keypoints.shape[2]
        a.image_size = float(image_size)
        a.keypoints, a.visibility = keypoints.data_ptr(), visibility.data_ptr()
        a.view_mask = view_mask.data_ptr() if view_mask is not None else None
        a.cam_R, a.R_bstride, a.R_vstride = cam_R.data_ptr(), cam_R.stride(0), cam_R.stride(1)
        a.cam_T, a.T_bstride, a.T_vstride = cam_T.data_ptr(), cam_T.stride(0), cam_T.stride(1)
        a.fov_deg, a.fov_bstride, a.fov_vstride = fov.data_ptr(), fov.stride(0), fov.stride(1)
        if aspect is not None:
            a.aspect, a.asp_bstride, a.asp_vstride = aspect.data_ptr(), aspect.stride(0), aspect.stride(1)
        return a

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, g_tri, _g_valid):
        keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask = ctx.saved_tensors
        if g_tri is None or not any(ctx.needs_input_grad[3:6]):
            return (None,) * 8
        B, V = keypoints.shape[0], keypoints.shape[1]
        dev = keypoints.device
        g_tri = g_tri.to(torch.float32).contiguous()
        a = TriangulateFunction._args(ctx.image_size, keypoints, visibility, cam_R, cam_T, fov, aspect, view_mask)
        a.g_tri = g_tri.data_ptr()
        gR = _fn._empty((B, V, 3, 3), device=dev, dtype=torch.float32) if ctx.needs_input_grad[3] else None
        gT = _fn._empty((B, V, 3), device=dev, dtype=torch.float32) if ctx.needs_input_grad[4] else None
        gf = _fn._empty((B, V), device=dev, dtype=torch.float32) if ctx.needs_input_grad[5] else None
        a.g_cam_R = gR.data_ptr() if gR is not None else None
        a.g_cam_T = gT.data_ptr() if gT is not None else None
        a.g_fov = gf.data_ptr() if gf is not None else None
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().smil_triangulate_dlt_backward(C.byref(a), _stream()), "smil_triangulate_dlt_backward")
        return None, None, None, gR, gT, gf, None, None


def triangulate_joints_dlt(keypoints_2d, visibility, cam_R, cam_T, fov, aspect=None, view_mask=None, image_size=512):
    """``_triangulate_joints_dlt`` (multiview_smil_regressor.py:1707-1838).  ``keypoints_2d`` (B,V,J,2) normalised (y,x);
    ``visibility`` (B,V,J); cameras broadcastable to (B,V,3,3) / (B,V,3) / (B,V) / (B,V); ``view_mask`` (B,V).
    Returns ``(triangulated (B,J,3), valid (B,J) bool)`` -- valid where at least two views see the joint."""
    B, V = keypoints_2d.shape[0], keypoints_2d.shape[1]
    R, T, f, asp = _cams(B, V, cam_R, cam_T, fov, aspect)
    tri, valid = TriangulateFunction.apply(float(image_size), _f(keypoints_2d), _f(visibility), R, T, f, asp, _f(view_mask))
    return tri, valid > 0.5


class Rot6dFunction(torch.autograd.Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, d6, columns):
        _check_cuda_f32("rotation_6d", d6, d6.device)
        flat = d6.reshape(-1, 6).contiguous()
        R = _fn._empty((flat.shape[0], 3, 3), device=d6.device, dtype=torch.float32)
        with torch.cuda.device(d6.device):
            _lib.check(_lib.lib().smil_rot6d_to_rotmat_forward(flat.data_ptr(), R.data_ptr(), flat.shape[0], int(columns),
                                                               _stream()), "smil_rot6d_to_rotmat_forward")
        ctx.columns = int(columns)
        ctx.save_for_backward(flat)
        ctx.shape = d6.shape
        return R.reshape(tuple(d6.shape[:-1]) + (3, 3))

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, gR):
        (flat,) = ctx.saved_tensors
        g = gR.to(torch.float32).reshape(-1, 9).contiguous()
        out = torch.empty_like(flat)
        with torch.cuda.device(flat.device):
            _lib.check(_lib.lib().smil_rot6d_to_rotmat_backward(flat.data_ptr(), g.data_ptr(), out.data_ptr(), flat.shape[0],
                                                                ctx.columns, _stream()), "smil_rot6d_to_rotmat_backward")
        return out.reshape(ctx.shape), None


def rotation_6d_to_matrix(rotation_6d, columns=False):
    """6-D rotation representation ``(..., 6)`` -> rotation matrices ``(..., 3, 3)`` by Gram-Schmidt.  ``columns=False``:
    b1, b2, b3 are the rows (``pytorch3d.transforms.rotation_6d_to_matrix``); ``columns=True``: the columns (the
    reference's ``_rotation_6d_to_matrix`` for camera rotations)."""
    if rotation_6d.shape[-1] != 6:
        raise RuntimeError("rotation_6d must have a trailing dimension of 6 (got %s)" % (tuple(rotation_6d.shape),))
    if not rotation_6d.is_cuda:
        raise RuntimeError("rotation_6d is on %s: smilify_b200 has no CPU fallback" % rotation_6d.device)
    return Rot6dFunction.apply(_f(rotation_6d), bool(columns))
