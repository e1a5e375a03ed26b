"""torch.autograd surface of the CUDA path.

The reference has no custom ``autograd.Function`` -- it differentiates ``SMAL.__call__`` op by
op (/root/reference/smal_model/smal_torch.py:198-370).  Here one Function covers the whole
call: forward = ``smil_forward``, backward = ``smil_backward`` (include/smil_b200.h), both
enqueued on the caller's current CUDA stream with caller-owned (torch-allocated) buffers.
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np
import torch

from . import _lib
from . import packer


class DeviceModel:
    """Owner of one ``SmilModel*`` (packed constants resident on one CUDA device)."""

    def __init__(self, tables: dict, device: torch.device):
        if device.type != "cuda":
            raise RuntimeError("smilify_b200 runs on CUDA devices only (got %s); there is no CPU fallback" % device)
        self.device = device
        self.index = device.index if device.index is not None else torch.cuda.current_device()
        self.tables = tables
        L = _lib.lib()
        desc, keep = _lib.desc_from_tables(tables, self.index)
        handle = C.c_void_p()
        _lib.check(L.smil_model_create(C.byref(desc), C.byref(handle)), "smil_model_create")
        del keep
        self.handle = handle
        self._finalizer = weakref.finalize(self, L.smil_model_destroy, handle)
        info = _lib.SmilModelInfo()
        _lib.check(L.smil_model_info(handle, C.byref(info)), "smil_model_info")
        self.info = info

    @property
    def num_joints(self):
        return self.tables["J"]

    @property
    def num_verts(self):
        return self.tables.get("V", 0)


def tree_only_model(parents: np.ndarray, device: torch.device) -> DeviceModel:
    """A model that carries just the kinematic-tree tables (for the stand-alone chain op)."""
    t = dict(J=int(parents.shape[0]))
    t.update(packer.tree_tables(np.asarray(parents)))
    return DeviceModel(t, device)


def _empty(*size, **kw):
    """Every device buffer handed to the library (outputs, saved state, scratch) is allocated here, so that
    tests/test_gpu_guards.py can swap in an allocator with poisoned contents and guard bands."""
    return torch.empty(*size, **kw)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _prep(t, lead_shape_ok=True):
    """-> (tensor kept alive, data_ptr, per-sample stride in elements).  A batch dimension that
    was produced by ``expand`` (stride 0) is passed through as stride 0 without materialising."""
    if t is None:
        return None, None, 0
    if t.dim() >= 2 and t.stride(0) == 0 and t[0].is_contiguous():
        return t, t.data_ptr(), 0
    if not t.is_contiguous():
        t = t.contiguous()
    return t, t.data_ptr(), (t.stride(0) if t.dim() >= 1 else 0)


def _check_cuda_f32(name, t, device):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError("%s is on %s: smilify_b200 has no CPU fallback, move it to %s" % (name, t.device, device))
    if t.device != device:
        raise RuntimeError("%s is on %s but the model lives on %s" % (name, t.device, device))
    if t.dtype != torch.float32:
        raise RuntimeError("%s must be float32 (got %s)" % (name, t.dtype))


class SmalFunction(torch.autograd.Function):
    """(beta, theta, trans, del_v, betas_logscale, betas_trans) -> (verts, joints, Rs, v_shaped, J_transformed)."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, dm: DeviceModel, flags: int, v_template, beta, theta, trans, del_v, betas_logscale, betas_trans):
        L = _lib.lib()
        dev = dm.device
        for n, t in (("beta", beta), ("theta", theta), ("trans", trans), ("del_v", del_v),
                     ("betas_logscale", betas_logscale), ("betas_trans", betas_trans), ("v_template", v_template)):
            _check_cuda_f32(n, t, dev)
        B = theta.shape[0]
        J, V = dm.num_joints, dm.num_verts
        nb = beta.shape[1]
        rotmat = bool(flags & _lib.F_THETA_IS_ROTMAT)
        joints_only = bool(flags & _lib.F_JOINTS_ONLY)
        if beta.shape[0] != B:
            raise RuntimeError("beta has batch %d but theta has batch %d" % (beta.shape[0], B))
        if theta.shape[1:] != ((J, 3, 3) if rotmat else (J, 3)):
            raise RuntimeError("theta has shape %s, expected (B, %d, %s)" % (tuple(theta.shape), J, "3, 3" if rotmat else "3"))
        for n, t in (("betas_logscale", betas_logscale), ("betas_trans", betas_trans)):
            if t is not None and tuple(t.shape) != (B, J, 3):
                raise RuntimeError("%s has shape %s, expected (%d, %d, 3)" % (n, tuple(t.shape), B, J))
        if trans is not None and tuple(trans.shape) != (B, 3):
            raise RuntimeError("trans has shape %s, expected (%d, 3)" % (tuple(trans.shape), B))
        if del_v is not None and tuple(del_v.shape) not in ((B, V, 3), (V, 3), (1, V, 3)):
            raise RuntimeError("del_v has shape %s, expected (%d, %d, 3) or (%d, 3)" % (tuple(del_v.shape), B, V, V))
        if v_template is not None and tuple(v_template.shape) != (V, 3):
            raise RuntimeError("v_template has shape %s, expected (%d, 3)" % (tuple(v_template.shape), V))

        with torch.cuda.device(dev):
            a = _lib.SmilFwdArgs()
            a.batch, a.num_betas_used, a.flags = B, nb, flags
            keep = []
            t, a.beta, a.beta_stride = _prep(beta); keep.append(t)
            t, a.theta, a.theta_stride = _prep(theta); keep.append(t)
            t, a.trans, a.trans_stride = _prep(trans); keep.append(t)
            t, a.betas_logscale, a.logscale_stride = _prep(betas_logscale); keep.append(t)
            t, a.betas_trans, a.btrans_stride = _prep(betas_trans); keep.append(t)
            if del_v is not None:
                dv = del_v.contiguous()
                keep.append(dv)
                a.del_v = dv.data_ptr()
                a.del_v_stride = V * 3 if (dv.dim() == 3 and dv.shape[0] == B) else 0
            if v_template is not None:
                vt = v_template.contiguous()
                keep.append(vt)
                a.v_template = vt.data_ptr()
            opts = dict(device=dev, dtype=torch.float32)
            joints = _empty((B, J, 3), **opts)
            Jt = _empty((B, J, 3), **opts)
            verts = v_shaped = None
            if not joints_only:
                verts = _empty((B, V, 3), **opts)
                v_shaped = _empty((B, V, 3), **opts)
                a.verts, a.v_shaped = verts.data_ptr(), v_shaped.data_ptr()
            Rs = None
            if not rotmat:
                Rs = _empty((B, J, 3, 3), **opts)
                a.Rs = Rs.data_ptr()
            a.joints, a.J_transformed = joints.data_ptr(), Jt.data_ptr()
            saved = _empty(L.smil_saved_bytes(dm.handle, B, flags), device=dev, dtype=torch.uint8)
            a.saved, a.saved_bytes = saved.data_ptr(), saved.numel()
            scratch = None
            nscratch = L.smil_fwd_scratch_bytes(dm.handle, C.byref(a))  # 0 on the fused path
            if nscratch:
                scratch = _empty(nscratch, device=dev, dtype=torch.uint8)
                a.scratch, a.scratch_bytes = scratch.data_ptr(), scratch.numel()
            _lib.check(L.smil_forward(dm.handle, C.byref(a), _stream()), "smil_forward")
            del scratch

        ctx.dm, ctx.flags, ctx.nb = dm, flags, nb
        # an output nobody differentiates arrives as None in backward (not a zero tensor), so the
        # kernels skip it; a zero g_v_shaped would also push the call off the fused path
        ctx.set_materialize_grads(False)
        ctx.has = (trans is not None, del_v is not None, betas_logscale is not None, betas_trans is not None,
                   v_template is not None)
        ctx.del_v_shape = tuple(del_v.shape) if del_v is not None else None
        ctx.save_for_backward(beta, theta, trans, del_v, betas_logscale, betas_trans, v_template, saved, v_shaped, Rs)
        ctx.mark_non_differentiable(Jt)
        return (verts if verts is not None else joints.new_empty(0), joints,
                Rs if Rs is not None else joints.new_empty(0),
                v_shaped if v_shaped is not None else joints.new_empty(0), Jt)

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, g_verts, g_joints, g_Rs, g_v_shaped, _g_Jt):
        L = _lib.lib()
        dm, flags, nb = ctx.dm, ctx.flags, ctx.nb
        beta, theta, trans, del_v, ls, bt, v_template, saved, v_shaped, Rs = ctx.saved_tensors
        dev = dm.device
        B = theta.shape[0]
        J, V = dm.num_joints, dm.num_verts
        rotmat = bool(flags & _lib.F_THETA_IS_ROTMAT)
        joints_only = bool(flags & _lib.F_JOINTS_ONLY)
        need = ctx.needs_input_grad  # (dm, flags, v_template, beta, theta, trans, del_v, ls, bt)

        def up(g, shape):
            if g is None or g.numel() == 0:
                return None
            g = g.to(torch.float32)
            if tuple(g.shape) != shape:
                g = g.expand(shape)
            g = g.contiguous()
            if g.data_ptr() % 16:   # a contiguous VIEW with a storage offset (narrow() of a cat / stack gradient): the TMA
                g = g.clone()       # bulk copies of the fused backward need 16-byte aligned rows
            return g

        with torch.cuda.device(dev):
            g = _lib.SmilBwdArgs()
            a = g.fwd
            a.batch, a.num_betas_used, a.flags = B, nb, flags
            keep = []
            t, a.beta, a.beta_stride = _prep(beta); keep.append(t)
            t, a.theta, a.theta_stride = _prep(theta); keep.append(t)
            t, a.trans, a.trans_stride = _prep(trans); keep.append(t)
            t, a.betas_logscale, a.logscale_stride = _prep(ls); keep.append(t)
            t, a.betas_trans, a.btrans_stride = _prep(bt); keep.append(t)
            if del_v is not None:
                dv = del_v.contiguous(); keep.append(dv)
                a.del_v = dv.data_ptr()
                a.del_v_stride = V * 3 if (dv.dim() == 3 and dv.shape[0] == B) else 0
            if v_template is not None:
                vt = v_template.contiguous(); keep.append(vt)
                a.v_template = vt.data_ptr()
            a.saved, a.saved_bytes = saved.data_ptr(), saved.numel()
            if v_shaped is not None:
                a.v_shaped = v_shaped.data_ptr()
            if Rs is not None:
                a.Rs = Rs.data_ptr()
            gv = None if joints_only else up(g_verts, (B, V, 3))
            gj = up(g_joints, (B, J, 3))
            gR = None if rotmat else up(g_Rs, (B, J, 3, 3))
            gvs = None if joints_only else up(g_v_shaped, (B, V, 3))
            keep += [gv, gj, gR, gvs]
            g.g_verts = gv.data_ptr() if gv is not None else None
            g.g_joints = gj.data_ptr() if gj is not None else None
            g.g_Rs = gR.data_ptr() if gR is not None else None
            g.g_v_shaped = gvs.data_ptr() if gvs is not None else None
            opts = dict(device=dev, dtype=torch.float32)
            # static-joint rigs: joints / Rs depend on neither the shape nor trans (smal_torch.py:257-260, 343-346),
            # so with no upstream gradient on the vertices those inputs get None, like the reference graph gives
            shape_path = not (bool(flags & _lib.F_STATIC_JOINTS) and gv is None and gvs is None)
            o_beta = _empty((B, nb), **opts) if (need[3] and nb > 0 and shape_path) else None
            o_theta = _empty(tuple(theta.shape), **opts) if need[4] else None
            o_trans = _empty((B, 3), **opts) if (need[5] and trans is not None and shape_path) else None
            want_vt = need[2] and v_template is not None and shape_path   # smal_torch.py:223-232: v_template is differentiable too
            o_delv = _empty((B, V, 3), **opts) if (((need[6] and del_v is not None) or want_vt) and shape_path) else None
            o_ls = _empty((B, J, 3), **opts) if (need[7] and ls is not None) else None
            o_bt = _empty((B, J, 3), **opts) if (need[8] and bt is not None) else None
            g.g_beta = o_beta.data_ptr() if o_beta is not None else None
            g.g_theta = o_theta.data_ptr() if o_theta is not None else None
            g.g_trans = o_trans.data_ptr() if o_trans is not None else None
            g.g_del_v = o_delv.data_ptr() if o_delv is not None else None
            g.g_logscale = o_ls.data_ptr() if o_ls is not None else None
            g.g_btrans = o_bt.data_ptr() if o_bt is not None else None
            scratch = _empty(L.smil_bwd_scratch_bytes(dm.handle, C.byref(g)), device=dev, dtype=torch.uint8)
            g.scratch, g.scratch_bytes = scratch.data_ptr(), scratch.numel()
            _lib.check(L.smil_backward(dm.handle, C.byref(g), _stream()), "smil_backward")
        o_vt = o_delv.sum(0) if (o_delv is not None and want_vt) else None
        if not (need[6] and del_v is not None):
            o_delv = None
        if o_delv is not None and ctx.del_v_shape != (B, V, 3):
            o_delv = o_delv.sum(0).reshape(ctx.del_v_shape)
        return None, None, o_vt, o_beta, o_theta, o_trans, o_delv, o_ls, o_bt


class RodriguesFunction(torch.autograd.Function):
    """Drop-in for ``batch_rodrigues`` (/root/reference/smal_model/batch_lbs.py:31-50)."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, theta):
        _check_cuda_f32("theta", theta, theta.device)
        th = theta.contiguous()
        R = _empty((th.shape[0], 3, 3), device=th.device, dtype=torch.float32)
        with torch.cuda.device(th.device):
            _lib.check(_lib.lib().smil_rodrigues_forward(th.data_ptr(), R.data_ptr(), th.shape[0], _stream()),
                       "smil_rodrigues_forward")
        ctx.save_for_backward(th)
        return R

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, gR):
        (th,) = ctx.saved_tensors
        gR = gR.to(torch.float32).contiguous()
        out = torch.empty_like(th)
        with torch.cuda.device(th.device):
            _lib.check(_lib.lib().smil_rodrigues_backward(th.data_ptr(), gR.data_ptr(), out.data_ptr(), th.shape[0],
                                                          _stream()), "smil_rodrigues_backward")
        return out


class ChainFunction(torch.autograd.Function):
    """Drop-in for ``batch_global_rigid_transformation`` (/root/reference/smal_model/batch_lbs.py:75-197)."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, dm: DeviceModel, flags: int, Rs, Js, ls, bt):
        B, J = Rs.shape[0], dm.num_joints
        for n, t in (("Rs", Rs), ("Js", Js), ("betas_logscale", ls), ("betas_trans", bt)):
            _check_cuda_f32(n, t, dm.device)
        Rs, Js = Rs.contiguous(), Js.contiguous()
        ls = ls.contiguous() if ls is not None else None
        bt = bt.contiguous() if bt is not None else None
        newJ = _empty((B, J, 3), device=dm.device, dtype=torch.float32)
        A = _empty((B, J, 3, 4), device=dm.device, dtype=torch.float32)
        a = _lib.SmilChainArgs()
        a.batch, a.flags = B, flags
        a.Rs, a.Js = Rs.data_ptr(), Js.data_ptr()
        a.betas_logscale = ls.data_ptr() if ls is not None else None
        a.betas_trans = bt.data_ptr() if bt is not None else None
        a.new_J, a.A = newJ.data_ptr(), A.data_ptr()
        with torch.cuda.device(dm.device):
            _lib.check(_lib.lib().smil_chain_forward(dm.handle, C.byref(a), _stream()), "smil_chain_forward")
        ctx.dm, ctx.flags = dm, flags
        ctx.save_for_backward(Rs, Js, ls, bt)
        return newJ, A

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, g_newJ, g_A):
        dm = ctx.dm
        Rs, Js, ls, bt = ctx.saved_tensors
        B, J = Rs.shape[0], dm.num_joints
        opts = dict(device=dm.device, dtype=torch.float32)
        a = _lib.SmilChainArgs()
        a.batch, a.flags = B, ctx.flags
        a.Rs, a.Js = Rs.data_ptr(), Js.data_ptr()
        a.betas_logscale = ls.data_ptr() if ls is not None else None
        a.betas_trans = bt.data_ptr() if bt is not None else None
        gJ = g_newJ.to(torch.float32).contiguous() if g_newJ is not None else None
        gA = g_A.to(torch.float32).contiguous() if g_A is not None else None
        a.g_new_J = gJ.data_ptr() if gJ is not None else None
        a.g_A = gA.data_ptr() if gA is not None else None
        o_Rs, o_Js = _empty((B, J, 3, 3), **opts), _empty((B, J, 3), **opts)
        o_ls = _empty((B, J, 3), **opts) if ls is not None else None
        o_bt = _empty((B, J, 3), **opts) if bt is not None else None
        a.g_Rs, a.g_Js = o_Rs.data_ptr(), o_Js.data_ptr()
        a.g_logscale = o_ls.data_ptr() if o_ls is not None else None
        a.g_btrans = o_bt.data_ptr() if o_bt is not None else None
        with torch.cuda.device(dm.device):
            _lib.check(_lib.lib().smil_chain_backward(dm.handle, C.byref(a), _stream()), "smil_chain_backward")
        return None, None, o_Rs, o_Js, o_ls, o_bt
