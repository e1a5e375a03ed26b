"""Parameter-space terms of the fitter objective and a fused Adam, on the device.

``fitter_prior_terms`` mirrors the argument meaning of the reference's own code:
  * joint limits, pose prior, splay, betas prior   /root/reference/smal_fitter/fitter.py:298-331
  * ``get_temporal``                               /root/reference/smal_fitter/fitter.py:337-350
as ``optimize_to_joints.py:154-172`` drives them: the window terms are evaluated per window of ``window`` consecutive
frames (``config.WINDOW_SIZE``) and summed, the temporal terms run over all consecutive frame pairs.  Value and
gradients come out of ONE kernel launch (+ a 1-block finalize); the autograd ``backward`` only scales the saved
gradients.  ``FusedAdam`` is ``torch.optim.Adam`` (defaults, optimize_to_joints.py:117-126) as one captured-graph-safe
launch over all parameter tensors.  CUDA only: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import torch

from . import _lib
from . import functions as _fn
from .functions import _check_cuda_f32, _stream


def identity_pose_prior():
    """The prior the reference builds for rigs without hard-coded body data (fitter.py:25-52): zero mean, identity
    precision, the root's three angles ignored -- the kernel's defaults when nothing is passed."""
    return dict(mean=None, precs=None, use=None)


class _FitterPriorFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, global_rot, joint_rot, trans, betas, consts, weights, window, frame_offset, total_frames, halo_prev, halo_next):
        dev = global_rot.device
        for n, t in (("global_rot", global_rot), ("joint_rot", joint_rot), ("trans", trans), ("betas", betas),
                     ("halo_prev", halo_prev), ("halo_next", halo_next)):
            _check_cuda_f32(n, t, dev)
        n = global_rot.shape[0]
        J = joint_rot.shape[1] + 1
        if tuple(global_rot.shape) != (n, 3) or tuple(joint_rot.shape) != (n, J - 1, 3):
            raise RuntimeError("global_rot (n, 3) / joint_rot (n, J-1, 3) expected, got %s / %s" % (tuple(global_rot.shape), tuple(joint_rot.shape)))
        if trans is not None and tuple(trans.shape) != (n, 3):
            raise RuntimeError("trans (n, 3) expected, got %s" % (tuple(trans.shape),))
        keep = []

        def ptr(t, shape=None, name=""):
            if t is None:
                return None
            t = t.detach()
            if t.dtype != torch.float32 or t.device != dev:
                t = t.to(device=dev, dtype=torch.float32)
            t = t.contiguous()
            if shape is not None and t.numel() != shape:
                raise RuntimeError("%s has %d elements, expected %d" % (name, t.numel(), shape))
            keep.append(t)
            return t.data_ptr()

        a = _lib.SmilFitterPriorArgs()
        a.num_frames, a.num_joints, a.window = n, J, int(window)
        a.num_betas = 0 if betas is None else betas.numel()
        a.frame_offset, a.total_frames = int(frame_offset), int(total_frames if total_frames is not None else n)
        a.global_rot, a.joint_rot, a.trans, a.betas = ptr(global_rot), ptr(joint_rot), ptr(trans), ptr(betas)
        P = 3 * J
        a.halo_prev, a.halo_next = ptr(halo_prev, P + 3, "halo_prev"), ptr(halo_next, P + 3, "halo_next")
        a.max_limits = ptr(consts.get("max_limits"), P - 3, "max_limits")
        a.min_limits = ptr(consts.get("min_limits"), P - 3, "min_limits")
        a.rotation_mask = ptr(consts.get("rotation_mask"), P - 3, "rotation_mask")
        a.global_mask = ptr(consts.get("global_mask"), 3, "global_mask")
        pp = consts.get("pose_prior") or {}
        a.pose_mean, a.pose_precs, a.pose_use = ptr(pp.get("mean"), P, "pose mean"), ptr(pp.get("precs"), P * P, "pose precs"), ptr(pp.get("use"), P, "pose use")
        nB = a.num_betas
        a.betas_mean, a.betas_prec = ptr(consts.get("betas_mean"), nB, "betas_mean"), ptr(consts.get("betas_prec"), nB * nB, "betas_prec")
        a.w_limit, a.w_pose, a.w_splay = (float(weights.get(k, 0.0)) for k in ("w_limit", "w_pose", "w_splay"))
        a.w_betas, a.w_temp = float(weights.get("w_betas", 0.0)), float(weights.get("w_temp", 0.0))
        a.g_scale, a.accumulate = 1.0, 0
        opts = dict(device=dev, dtype=torch.float32)
        loss = _fn._empty((), **opts)
        partials = _fn._empty(n, **opts)
        g_g, g_j = _fn._empty((n, 3), **opts), _fn._empty((n, J - 1, 3), **opts)
        g_t = _fn._empty((n, 3), **opts) if trans is not None else None
        g_b = _fn._empty(nB, **opts) if betas is not None else None
        a.loss, a.partials = loss.data_ptr(), partials.data_ptr()
        a.g_global_rot, a.g_joint_rot = g_g.data_ptr(), g_j.data_ptr()
        a.g_trans = g_t.data_ptr() if g_t is not None else None
        a.g_betas = g_b.data_ptr() if g_b is not None else None
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().smil_fitter_prior_terms(C.byref(a), _stream()), "smil_fitter_prior_terms")
        ctx.save_for_backward(g_g, g_j, *( [g_t] if g_t is not None else []), *([g_b] if g_b is not None else []))
        ctx.has = (g_t is not None, g_b is not None)
        ctx.betas_shape = None if betas is None else tuple(betas.shape)
        return loss

    @staticmethod
    def backward(ctx, g_loss):
        saved = list(ctx.saved_tensors)
        g_g, g_j = saved[0], saved[1]
        g_t = saved[2] if ctx.has[0] else None
        g_b = saved[2 + int(ctx.has[0])] if ctx.has[1] else None
        need = ctx.needs_input_grad
        return (g_g * g_loss if need[0] else None, g_j * g_loss if need[1] else None,
                (g_t * g_loss) if (g_t is not None and need[2]) else None,
                (g_b * g_loss).reshape(ctx.betas_shape) if (g_b is not None and need[3]) else None,
                None, None, None, None, None, None, None)


def fitter_prior_terms(global_rot, joint_rot, trans=None, betas=None, weights: Optional[Dict[str, float]] = None, window=10,
                       max_limits=None, min_limits=None, pose_prior=None, betas_mean=None, betas_prec=None,
                       rotation_mask=None, global_mask=None, frame_offset=0, total_frames=None, halo_prev=None,
                       halo_next=None):
    """Sum of the enabled terms (``weights``: w_limit, w_pose, w_splay, w_betas, w_temp; a term with weight 0 or without
    its constants is skipped, as in the reference).  ``pose_prior``: dict(mean, precs, use) or None for the identity
    prior of fitter.py:25-52.  Sharded sequences pass ``frame_offset`` / ``total_frames`` and, with ``w_temp > 0``, the
    neighbouring frames' parameters as ``halo_prev`` / ``halo_next`` (``[global(3) | joints(3(J-1)) | trans(3)]``)."""
    consts = dict(max_limits=max_limits, min_limits=min_limits, pose_prior=pose_prior if pose_prior is not None else identity_pose_prior(),
                  betas_mean=betas_mean, betas_prec=betas_prec, rotation_mask=rotation_mask, global_mask=global_mask)
    return _FitterPriorFunction.apply(global_rot, joint_rot, trans, betas, consts, dict(weights or {}), window, frame_offset,
                                      total_frames, halo_prev, halo_next)


class FusedAdam:
    """``torch.optim.Adam`` with default ``betas`` / ``eps``, no weight decay, no amsgrad (what
    optimize_to_joints.py:117-126 constructs), as ONE launch over all parameters.  ``params``: tensors or
    ``{"params": [...], "lr": ...}`` groups like the reference passes (its ``fov`` group has its own rate).  The step
    counter lives on the device, so ``step()`` can be captured in a CUDA graph and replayed."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.params, self.lrs = [], []
        for g in params:
            if isinstance(g, dict):
                for p in g["params"]:
                    self.params.append(p)
                    self.lrs.append(float(g.get("lr", lr)))
            else:
                self.params.append(g)
                self.lrs.append(float(lr))
        if not self.params:
            raise ValueError("FusedAdam: no parameters")
        if len(self.params) > _lib.ADAM_MAX_TENSORS:   # one launch, one tick of the device step counter
            raise ValueError("FusedAdam: at most %d parameter tensors (got %d)" % (_lib.ADAM_MAX_TENSORS, len(self.params)))
        dev = self.params[0].device
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous() or p.device != dev:
                raise RuntimeError("FusedAdam needs contiguous float32 CUDA parameters on one device (there is no CPU fallback)")
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.device = dev

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    def step(self, grads: Optional[Sequence[torch.Tensor]] = None):
        """``grads``: explicit gradient tensors (same order as the parameters); default ``p.grad``.  Parameters whose
        gradient is ``None`` are skipped (torch.optim does the same)."""
        L = _lib.lib()
        idx = [i for i, p in enumerate(self.params) if (grads[i] if grads is not None else p.grad) is not None]
        with torch.cuda.device(self.device):
            for c0 in range(0, len(idx), _lib.ADAM_MAX_TENSORS):
                chunk = idx[c0:c0 + _lib.ADAM_MAX_TENSORS]
                a = _lib.SmilAdamArgs()
                a.num_tensors = len(chunk)
                a.beta1, a.beta2, a.eps = self.betas[0], self.betas[1], self.eps
                keep = []
                for t, i in enumerate(chunk):
                    g = grads[i] if grads is not None else self.params[i].grad
                    if g.dtype != torch.float32 or not g.is_contiguous():
                        g = g.float().contiguous()
                    keep.append(g)
                    a.param[t], a.grad[t] = self.params[i].data_ptr(), g.data_ptr()
                    a.exp_avg[t], a.exp_avg_sq[t] = self.exp_avg[i].data_ptr(), self.exp_avg_sq[i].data_ptr()
                    a.numel[t], a.lr[t] = self.params[i].numel(), self.lrs[i]
                a.step = self.step_count.data_ptr()
                _lib.check(L.smil_adam_step(C.byref(a), _stream()), "smil_adam_step")
