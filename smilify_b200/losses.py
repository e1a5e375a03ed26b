"""2D joint reprojection + keypoint losses on the CUDA path (``smil_project_loss_*`` /
``smil_keypoint3d_loss_*`` in include/smil_b200.h).

One fused kernel pair replaces, per flavour (paths relative to /root/reference):

* ``fitter_joint_loss``        -- ``Renderer.forward`` projection slice (smal_fitter/p3d_renderer.py:130-140)
  + the SMALFitter joint loss (smal_fitter/fitter.py:280-296);
* ``multiview_keypoint_loss``  -- ``_batch_project_joints_to_views`` + ``_compute_batched_keypoint_loss``
  (smal_fitter/neuralSMIL/multiview_smil_regressor.py:1623-1705, 2319-2388);
* ``singleview_keypoint_loss`` -- normalise/clamp + ``_compute_batch_keypoint_loss``
  (smal_fitter/neuralSMIL/smil_image_regressor.py:2727-2732, 3032-3115);
* ``keypoint3d_loss``          -- ``_compute_batch_keypoint_3d_loss`` (smil_image_regressor.py:3334-3407).

Cameras follow PyTorch3D's ``FoVPerspectiveCameras`` (row vectors, ``X @ R + T``, fov in degrees);
projected points are in the reference's ``(y, x)`` order.  Gradients flow to the joints and to the
camera ``R``, ``T`` and ``fov`` (``fov`` is an ``nn.Parameter`` in SMALFitter, fitter.py:231);
``aspect_ratio`` is data in every reference call site and gets no gradient.  CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from . import functions as _fn
from .functions import _check_cuda_f32, _stream


def _inner_contig(t, inner):
    """Make the trailing ``inner`` dims dense without materialising broadcast leading dims."""
    want, acc = [], 1
    for s in reversed(t.shape[t.dim() - inner:]):
        want.append(acc)
        acc *= s
    if list(t.stride()[t.dim() - inner:]) == list(reversed(want)):
        return t
    return t.contiguous()


class ProjectLossFunction(torch.autograd.Function):
    """(joints, cam_R, cam_T, fov) -> (loss, proj); everything else is data."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, kind, image_size, norm_size, clamp_lo, clamp_hi, weight, joints, cam_R, cam_T, fov, aspect, target,
                visibility, view_mask, joint_weights, sample_mask):
        dev = joints.device
        for n, t in (("joints", joints), ("cam_R", cam_R), ("cam_T", cam_T), ("fov", fov), ("aspect", aspect),
                     ("target", target), ("visibility", visibility), ("view_mask", view_mask),
                     ("joint_weights", joint_weights), ("sample_mask", sample_mask)):
            _check_cuda_f32(n, t, dev)
        B, Jc = joints.shape[0], joints.shape[1]
        V = cam_R.shape[1]
        if tuple(cam_R.shape) != (B, V, 3, 3) or tuple(cam_T.shape) != (B, V, 3) or tuple(fov.shape) != (B, V):
            raise RuntimeError("cameras must be expanded to (B,V,3,3) / (B,V,3) / (B,V); got %s %s %s" % (
                tuple(cam_R.shape), tuple(cam_T.shape), tuple(fov.shape)))
        if tuple(target.shape) != (B, V, Jc, 2):
            raise RuntimeError("target has shape %s, expected %s" % (tuple(target.shape), (B, V, Jc, 2)))
        joints = _inner_contig(joints, 2)
        cam_R, cam_T = _inner_contig(cam_R, 2), _inner_contig(cam_T, 1)
        target = target.contiguous()
        visibility = visibility.contiguous() if visibility is not None else None
        view_mask = view_mask.contiguous() if view_mask is not None else None
        joint_weights = joint_weights.contiguous() if joint_weights is not None else None
        sample_mask = sample_mask.contiguous() if sample_mask is not None else None
        opts = dict(device=dev, dtype=torch.float32)
        proj = _fn._empty((B, V, Jc, 2), **opts)
        loss = _fn._empty((), **opts)
        partials = _fn._empty(B * V * 3 + 4, **opts)
        a = ProjectLossFunction._args(kind, image_size, norm_size, clamp_lo, clamp_hi, weight, joints, cam_R, cam_T, fov,
                                      aspect, target, visibility, view_mask, joint_weights, sample_mask, partials)
        a.proj, a.loss = proj.data_ptr(), loss.data_ptr()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().smil_project_loss_forward(C.byref(a), _stream()), "smil_project_loss_forward")
        ctx.consts = (kind, image_size, norm_size, clamp_lo, clamp_hi, weight)
        ctx.save_for_backward(joints, cam_R, cam_T, fov, aspect, target, visibility, view_mask, joint_weights,
                              sample_mask, partials)
        ctx.mark_non_differentiable(proj)
        ctx.set_materialize_grads(False)
        return loss, proj

    @staticmethod
    def _args(kind, image_size, norm_size, clamp_lo, clamp_hi, weight, joints, cam_R, cam_T, fov, aspect, target,
              visibility, view_mask, joint_weights, sample_mask, partials):
        B, Jc = joints.shape[0], joints.shape[1]
        V = cam_R.shape[1]
        a = _lib.SmilProjArgs()
        a.kind, a.batch, a.num_views, a.num_joints = kind, B, V, Jc
        a.image_size, a.norm_size, a.clamp_lo, a.clamp_hi, a.weight = image_size, norm_size, clamp_lo, clamp_hi, weight
        a.joints, a.joints_stride = joints.data_ptr(), joints.stride(0)
        a.cam_R, a.R_bstride, a.R_vstride = cam_R.data_ptr(), cam_R.stride(0), cam_R.stride(1)
        a.cam_T, a.T_bstride, a.T_vstride = cam_T.data_ptr(), cam_T.stride(0), cam_T.stride(1)
        a.fov_deg, a.fov_bstride, a.fov_vstride = fov.data_ptr(), fov.stride(0), fov.stride(1)
        if aspect is not None:
            a.aspect, a.asp_bstride, a.asp_vstride = aspect.data_ptr(), aspect.stride(0), aspect.stride(1)
        a.target = target.data_ptr()
        a.visibility = visibility.data_ptr() if visibility is not None else None
        a.view_mask = view_mask.data_ptr() if view_mask is not None else None
        a.joint_weights = joint_weights.data_ptr() if joint_weights is not None else None
        a.sample_mask = sample_mask.data_ptr() if sample_mask is not None else None
        a.partials = partials.data_ptr()
        return a

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, g_loss, _g_proj):
        joints, cam_R, cam_T, fov, aspect, target, visibility, view_mask, joint_weights, sample_mask, partials = \
            ctx.saved_tensors
        need = ctx.needs_input_grad  # 6 consts, then joints, cam_R, cam_T, fov, ...
        if g_loss is None:
            return (None,) * 16
        B, Jc = joints.shape[0], joints.shape[1]
        V = cam_R.shape[1]
        dev = joints.device
        opts = dict(device=dev, dtype=torch.float32)
        gl = g_loss.to(torch.float32).reshape(1).contiguous()
        a = ProjectLossFunction._args(*ctx.consts, joints, cam_R, cam_T, fov, aspect, target, visibility, view_mask,
                                      joint_weights, sample_mask, partials)
        a.g_loss = gl.data_ptr()
        g_j = _fn._empty((B, Jc, 3), **opts) if need[6] else None
        g_R = _fn._empty((B, V, 3, 3), **opts) if need[7] else None
        g_T = _fn._empty((B, V, 3), **opts) if need[8] else None
        g_f = _fn._empty((B, V), **opts) if need[9] else None
        a.g_joints = g_j.data_ptr() if g_j is not None else None
        a.g_cam_R = g_R.data_ptr() if g_R is not None else None
        a.g_cam_T = g_T.data_ptr() if g_T is not None else None
        a.g_fov = g_f.data_ptr() if g_f is not None else None
        if g_j is not None or g_R is not None or g_T is not None or g_f is not None:
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().smil_project_loss_backward(C.byref(a), _stream()), "smil_project_loss_backward")
        return (None,) * 6 + (g_j, g_R, g_T, g_f) + (None,) * 6


class Keypoint3dLossFunction(torch.autograd.Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, joints, target, visibility, joint_weights, sample_mask):
        dev = joints.device
        for n, t in (("joints", joints), ("target", target), ("visibility", visibility),
                     ("joint_weights", joint_weights), ("sample_mask", sample_mask)):
            _check_cuda_f32(n, t, dev)
        if joints.dim() != 3 or joints.shape[2] != 3 or target.shape != joints.shape:
            raise RuntimeError("joints/target must both be (B,J,3); got %s %s" % (tuple(joints.shape), tuple(target.shape)))
        joints, target = joints.contiguous(), target.contiguous()
        tens = [None if t is None else t.contiguous() for t in (visibility, joint_weights, sample_mask)]
        B, J = joints.shape[:2]
        loss = _fn._empty((), device=dev, dtype=torch.float32)
        partials = _fn._empty(B * 3 + 4, device=dev, dtype=torch.float32)
        a = Keypoint3dLossFunction._args(joints, target, *tens, partials)
        a.loss = loss.data_ptr()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().smil_keypoint3d_loss_forward(C.byref(a), _stream()), "smil_keypoint3d_loss_forward")
        ctx.save_for_backward(joints, target, *tens, partials)
        ctx.set_materialize_grads(False)
        return loss

    @staticmethod
    def _args(joints, target, visibility, joint_weights, sample_mask, partials):
        a = _lib.SmilKp3dArgs()
        a.batch, a.num_joints = joints.shape[0], joints.shape[1]
        a.joints, a.target = joints.data_ptr(), target.data_ptr()
        a.visibility = visibility.data_ptr() if visibility is not None else None
        a.joint_weights = joint_weights.data_ptr() if joint_weights is not None else None
        a.sample_mask = sample_mask.data_ptr() if sample_mask is not None else None
        a.partials = partials.data_ptr()
        return a

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, g_loss):
        joints, target, visibility, joint_weights, sample_mask, partials = ctx.saved_tensors
        if g_loss is None or not ctx.needs_input_grad[0]:
            return (None,) * 5
        gl = g_loss.to(torch.float32).reshape(1).contiguous()
        a = Keypoint3dLossFunction._args(joints, target, visibility, joint_weights, sample_mask, partials)
        g_j = torch.empty_like(joints)
        a.g_loss, a.g_joints = gl.data_ptr(), g_j.data_ptr()
        with torch.cuda.device(joints.device):
            _lib.check(_lib.lib().smil_keypoint3d_loss_backward(C.byref(a), _stream()), "smil_keypoint3d_loss_backward")
        return g_j, None, None, None, None


# ---- public wrappers ------------------------------------------------------------------------------
def _f(t):
    return None if t is None else (t if t.dtype == torch.float32 else t.float())


def _cams(B, V, R, T, fov, aspect):
    """Broadcast camera tensors to (B,V,...) views (stride 0 where shared: nothing is materialised,
    and autograd sums the per-(frame,view) gradients back onto shared parameters)."""
    R = _f(R)
    R = R.reshape((1,) * (4 - R.dim()) + tuple(R.shape)) if R.dim() < 4 else R
    T = _f(T)
    T = T.reshape((1,) * (3 - T.dim()) + tuple(T.shape)) if T.dim() < 3 else T
    fov = _f(torch.as_tensor(fov, device=R.device))
    fov = fov.reshape((1,) * (2 - fov.dim()) + tuple(fov.shape)) if fov.dim() < 2 else fov
    if aspect is not None:
        aspect = _f(torch.as_tensor(aspect, device=R.device))
        aspect = aspect.reshape((1,) * (2 - aspect.dim()) + tuple(aspect.shape)) if aspect.dim() < 2 else aspect
        aspect = aspect.expand(B, V)
    return R.expand(B, V, 3, 3), T.expand(B, V, 3), fov.expand(B, V), aspect


def fitter_joint_loss(joints, cam_R, cam_T, fov, target_yx, visibility, w_j2d=1.0, image_size=512):
    """``w_j2d * F.mse_loss(proj, target)`` with invisible joints forced to -1 on both sides
    (fitter.py:292-296).  ``joints`` (N,Jc,3) already carry ``+trans`` and the canonical-joint
    selection (fitter.py:280-283); one camera per frame: ``cam_R`` (1|N,3,3), ``cam_T`` (1|N,3),
    ``fov`` (1|N,) degrees.  Returns ``(loss, proj (N,Jc,2) in (y,x) pixels after the -1 fill)``."""
    N, Jc = joints.shape[:2]
    R = _f(cam_R).reshape(-1, 1, 3, 3)
    T = _f(cam_T).reshape(-1, 1, 3)
    f = _f(torch.as_tensor(fov, device=joints.device)).reshape(-1, 1)
    R, T, f, _ = _cams(N, 1, R, T, f, None)
    loss, proj = ProjectLossFunction.apply(
        _lib.LOSS_FITTER, float(image_size), 1.0, 0.0, 0.0, float(w_j2d), _f(joints), R, T, f, None,
        _f(target_yx).reshape(N, 1, Jc, 2), _f(visibility).reshape(N, 1, Jc), None, None, None)
    return loss, proj.reshape(N, Jc, 2)


def multiview_keypoint_loss(joints, cam_R, cam_T, fov, aspect, target, visibility, view_mask=None, joint_weights=None,
                            image_size=512):
    """``_compute_batched_keypoint_loss(_batch_project_joints_to_views(...), ...)``
    (multiview_smil_regressor.py:1623-1705, 2319-2388).  ``joints`` (B,Jc,3) world space; cameras
    broadcastable to (B,V,3,3)/(B,V,3)/(B,V)/(B,V); ``target`` (B,V,Jc,2) normalised (y,x);
    ``visibility`` (B,V,Jc); ``view_mask`` (B,V).  Returns ``(loss, proj (B,V,Jc,2) in [0,1])``."""
    B, V = target.shape[0], target.shape[1]
    R, T, f, asp = _cams(B, V, cam_R, cam_T, fov, aspect)
    return ProjectLossFunction.apply(
        _lib.LOSS_MULTIVIEW, float(image_size), float(image_size) + 1e-8, 0.0, 1.0, 1.0, _f(joints), R, T, f, asp,
        _f(target), _f(visibility), _f(view_mask), _f(joint_weights), None)


def singleview_keypoint_loss(joints, cam_R, cam_T, fov, target, visibility, joint_weights=None, sample_mask=None,
                             image_size=512, norm_size=512.0):
    """Single-view flavour (smil_image_regressor.py:2727-2732, 3032-3115): one camera per sample,
    ``proj / 512`` clamped to [-10,10], per-sample normalisation, mean over samples with at least
    one valid joint; ``sample_mask`` marks samples whose keypoint data is present."""
    B, Jc = joints.shape[:2]
    R, T, f, _ = _cams(B, 1, _f(cam_R).reshape(-1, 1, 3, 3), _f(cam_T).reshape(-1, 1, 3),
                       _f(torch.as_tensor(fov, device=joints.device)).reshape(-1, 1), None)
    loss, proj = ProjectLossFunction.apply(
        _lib.LOSS_SINGLEVIEW, float(image_size), float(norm_size) + 1e-8, -10.0, 10.0, 1.0, _f(joints), R, T, f, None,
        _f(target).reshape(B, 1, Jc, 2), _f(visibility).reshape(B, 1, Jc), None, _f(joint_weights), _f(sample_mask))
    return loss, proj.reshape(B, Jc, 2)


def keypoint3d_loss(joints, target, visibility=None, joint_weights=None, sample_mask=None):
    """``_compute_batch_keypoint_3d_loss`` (smil_image_regressor.py:3334-3407)."""
    return Keypoint3dLossFunction.apply(_f(joints), _f(target), _f(visibility), _f(joint_weights), _f(sample_mask))


def pca_joint_values(scale_weights, trans_weights, scaledirs, transdirs, translation_factor=0.01, min_scale=1e-4):
    """``_transform_separate_pca_weights_to_joint_values`` (smil_image_regressor.py:1166-1232): PCA weights
    (B,C) -> ``(log_beta_scales, betas_trans)`` (B,J,3) that feed ``SMAL(..., betas_logscale=, betas_trans=)``.
    ``scaledirs`` / ``transdirs`` are (J,3,C) as the reference caches them (the pickle stores (C,J,3)).
    Stays in torch: two tiny einsums whose outputs are kernel inputs."""
    B = scale_weights.shape[0]
    if scaledirs is None or transdirs is None:
        J = 0 if scaledirs is None and transdirs is None else (scaledirs if scaledirs is not None else transdirs).shape[0]
        z = torch.zeros(B, J, 3, device=scale_weights.device)
        return z, z.clone()
    scale = torch.clamp(1.0 + torch.einsum("jck,bk->bjc", scaledirs, scale_weights), min=min_scale)
    log_beta_scales = torch.log(scale)
    if trans_weights is not None:
        betas_trans = torch.einsum("jck,bk->bjc", transdirs, trans_weights) * translation_factor
    else:
        betas_trans = torch.zeros(B, scaledirs.shape[0], 3, device=scale_weights.device)
    # The reference tests ``isfinite(...).all()`` on the host and only then sanitises; nan_to_num is the identity (and
    # passes gradients through unchanged) on finite values, so it is applied unconditionally here: no device->host
    # sync per call, and the function can be captured in a CUDA graph.
    log_beta_scales = torch.nan_to_num(log_beta_scales, nan=0.0, posinf=0.0, neginf=-10.0)
    if trans_weights is not None:
        betas_trans = torch.nan_to_num(betas_trans, nan=0.0, posinf=0.0, neginf=0.0)
    return log_beta_scales, betas_trans
