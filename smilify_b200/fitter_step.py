"""Multi-GPU shared-parameter fitting step (BASELINE configs[3]).

``SMALFitter`` (/root/reference/smal_fitter/fitter.py:150-296) optimises per-frame parameters
(``global_rotation``, ``joint_rotations``, ``trans``) together with parameters shared by every
frame of the sequence: ONE ``betas`` vector that is ``expand``-ed over the frames (fitter.py:244),
the camera ``fov`` (fitter.py:231) and, when enabled, shared ``log_beta_scales`` / ``betas_trans``
(fitter.py:183-189).  Frames are independent through the SMAL call and the joint loss, so the
frames are sharded contiguously over the ranks (one process per GPU) and the ONLY exchange is the
gradient of the shared parameters: they are packed into one fp32 buffer (nB + 1 (+ 6J) floats =
24 B ... 1.4 KB) and summed with a single ``all_reduce`` per optimiser step (NCCL over
NVLink/NVSwitch on GPUs, gloo in the CPU tests).  Per-frame parameters never leave their rank.

Nothing here launches kernels itself: the SMAL forward/backward and the loss are the C-ABI CUDA
path (``smal_torch.SMAL``, ``losses.fitter_joint_loss``); the step function is injectable so the
world-size-2 gloo tests can drive the same host logic with a CPU stand-in.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Sequence

import torch
import torch.distributed as dist


def shard_range(num_frames: int, rank: int, world: int):
    """Contiguous frame shard of ``rank`` (the first ``num_frames % world`` ranks get one extra)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside [0,%d)" % (rank, world))
    base, extra = divmod(num_frames, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_grads(tensors: Sequence[Optional[torch.Tensor]], like: Sequence[torch.Tensor]) -> torch.Tensor:
    """Flatten the gradients of the shared parameters into one fp32 buffer (``None`` -> zeros)."""
    parts = []
    for g, p in zip(tensors, like):
        parts.append(torch.zeros(p.numel(), dtype=torch.float32, device=p.device) if g is None
                     else g.detach().reshape(-1).to(torch.float32))
    return torch.cat(parts) if parts else torch.zeros(0)


def unpack_grads(buf: torch.Tensor, like: Sequence[torch.Tensor]):
    out, o = [], 0
    for p in like:
        n = p.numel()
        out.append(buf[o:o + n].reshape(p.shape).to(p.dtype))
        o += n
    return out


def allreduce_shared_grads(shared: Sequence[torch.Tensor], group=None) -> int:
    """Sum ``p.grad`` of every shared parameter over the ranks with ONE collective.
    Returns the number of bytes in the packed buffer (0 when not distributed)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return 0
    buf = pack_grads([p.grad for p in shared], shared)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    for p, g in zip(shared, unpack_grads(buf, shared)):
        p.grad = g.clone() if p.grad is None else p.grad.copy_(g)
    return buf.numel() * 4


def pack_boundary_frames(frame_params: Dict[str, torch.Tensor]) -> torch.Tensor:
    """``[first frame | last frame]`` of a shard, each ``[global(3) | joints(3(J-1)) | trans(3)]`` (the halo layout of
    ``smil_fitter_prior_terms``)."""
    def one(i):
        return torch.cat([frame_params["global_rotation"][i].reshape(-1), frame_params["joint_rotations"][i].reshape(-1),
                          frame_params["trans"][i].reshape(-1)])
    return torch.cat([one(0), one(-1)]).detach().to(torch.float32)


def exchange_halos(frame_params: Dict[str, torch.Tensor], group=None):
    """The 1-frame halo of SURVEY.md section 8(e) for the temporal term (fitter.py:337-350 couples frame i and i+1):
    every rank contributes its first and last frame to ONE ``all_gather`` (2 (3J + 3) floats per rank) and picks the
    last frame of the rank before it and the first frame of the rank after it.  Returns ``(halo_prev, halo_next)``,
    ``None`` at the ends of the sequence (and for a single process)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None, None
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = pack_boundary_frames(frame_params)
    n = mine.numel() // 2
    allb = torch.empty(world * 2 * n, dtype=mine.dtype, device=mine.device)
    dist.all_gather_into_tensor(allb, mine, group=group)
    allb = allb.view(world, 2, n)
    return (allb[rank - 1, 1] if rank > 0 else None), (allb[rank + 1, 0] if rank + 1 < world else None)


def with_prior_terms(loss_fn, num_frames_total: int, frame_offset: int, weights: Dict[str, float], window: int = 10,
                     group=None, prior_fn=None, **consts):
    """``loss_fn`` + the fitter's parameter-space terms (fitter.py:298-350) for :class:`ShardedFitterStep`.

    The step divides what ``loss_fn`` returns by ``num_frames_total`` (the joint loss is a mean over all frames); the
    prior terms carry their own per-window normalisation, so they enter multiplied by ``num_frames_total``.  With
    ``w_temp > 0`` the shard boundaries are exchanged by :func:`exchange_halos` (one small all_gather per step, inside
    the captured graph).  ``prior_fn`` defaults to the CUDA kernel (``fitter_terms.fitter_prior_terms``); the gloo CPU
    tests inject a stand-in with the same signature."""
    if prior_fn is None:
        from .fitter_terms import fitter_prior_terms as prior_fn

    def fn(fp, sp):
        hp = hn = None
        if weights.get("w_temp", 0.0) > 0.0:
            hp, hn = exchange_halos(fp, group)
        prior = prior_fn(fp["global_rotation"], fp["joint_rotations"], fp["trans"], sp.get("betas"), weights, window=window,
                         frame_offset=frame_offset, total_frames=num_frames_total, halo_prev=hp, halo_next=hn, **consts)
        return loss_fn(fp, sp) + prior * float(num_frames_total)

    return fn


class ShardedFitterStep:
    """One optimisation step of a frame-sharded joint fit.

    ``frame_params``  dict name -> per-frame leaf tensor of THIS rank's shard, shape (n_local, ...)
    ``shared_params`` dict name -> leaf tensor replicated on every rank (``betas`` (nB,), ``fov`` (1,), ...)
    ``loss_fn(frame_params, shared_params) -> scalar``  the *sum* over the local frames of the
        per-frame loss terms (so that the global objective is ``sum_ranks loss / num_frames_total``,
        i.e. the same ``mean`` the single-process fitter takes, fitter.py:296).
    """

    def __init__(self, frame_params: Dict[str, torch.Tensor], shared_params: Dict[str, torch.Tensor],
                 loss_fn: Callable[[Dict[str, torch.Tensor], Dict[str, torch.Tensor]], torch.Tensor],
                 num_frames_total: int, group=None):
        self.frame_params, self.shared_params = frame_params, shared_params
        self.loss_fn, self.group = loss_fn, group
        self.num_frames_total = int(num_frames_total)
        self.collective_bytes = 0

    def __call__(self) -> torch.Tensor:
        """Backward through the local shard, then the one packed all-reduce.  Returns the GLOBAL loss
        (identical on every rank); gradients are left in ``.grad`` of every leaf."""
        for p in list(self.frame_params.values()) + list(self.shared_params.values()):
            p.grad = None
        local = self.loss_fn(self.frame_params, self.shared_params) / self.num_frames_total
        local.backward()
        shared = list(self.shared_params.values())
        distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1
        if not distributed:
            return local.detach()
        # the loss value rides in the same buffer: still ONE collective per step
        buf = torch.cat([pack_grads([p.grad for p in shared], shared), local.detach().reshape(1).to(torch.float32)])
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
        self.collective_bytes = buf.numel() * 4
        for p, g in zip(shared, unpack_grads(buf[:-1], shared)):
            p.grad = g.clone() if p.grad is None else p.grad.copy_(g)
        return buf[-1].clone()


class GraphedShardedFitterStep(ShardedFitterStep):
    """:class:`ShardedFitterStep` with the whole step -- loss, backward through the CUDA path, packing of the shared
    gradients, the NCCL all-reduce and the write-back into ``.grad`` -- captured ONCE in a CUDA graph and replayed
    per step.  An optimiser loop over a fixed window of frames (fitter.py:150-296 runs the same shapes for hundreds
    of epochs) then costs one graph launch instead of ~40 Python-side launches plus an eager collective, which is
    what bounds the step once the frames are spread over 8 GPUs (8 192 frames per GPU take ~0.45 ms of kernels).
    NCCL collectives are stream-capturable once the communicator exists, so one eager all-reduce is issued before the
    capture; ``capture_collective=False`` (or a backend that cannot be captured) keeps the all-reduce as an eager call
    after the replay.  Leaves keep their storage: update them in place (``p.data.add_``, a captured or foreach
    optimiser); ``.grad`` tensors are allocated during capture and rewritten by every replay.  CUDA only."""

    def __init__(self, frame_params, shared_params, loss_fn, num_frames_total, group=None, warmup=3,
                 capture_collective=True, optimizer=None):
        super().__init__(frame_params, shared_params, loss_fn, num_frames_total, group)
        # optimizer: an object with a graph-capturable ``step()`` over the leaves (``fitter_terms.FusedAdam``): the update
        # then runs inside the captured step, after the shared gradients are final (the whole of optimize_to_joints.py:
        # 153-175 is one graph replay).  NOTE: the warm-up iterations below do NOT step it.
        self.optimizer = optimizer
        leaves = list(frame_params.values()) + list(shared_params.values())
        if not leaves or not all(p.is_cuda for p in leaves):
            raise RuntimeError("GraphedShardedFitterStep needs CUDA leaf tensors (there is no CPU fallback)")
        dev = leaves[0].device
        self._shared = list(shared_params.values())
        self._distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                buf = self._local()
                if self._distributed:
                    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)   # also creates the communicator
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.collective_in_graph = False
        if self._distributed and capture_collective:
            try:
                self._capture(True)
                self.collective_in_graph = True
            except Exception:   # a backend / build whose collectives cannot be stream-captured
                torch.cuda.synchronize(dev)
                self._capture(False)
        else:
            self._capture(False)
        self._leaves = leaves
        self._grads = [p.grad for p in leaves]   # static: every replay rewrites these tensors
        if self._distributed:
            self.collective_bytes = self._buf.numel() * 4

    def _capture(self, with_collective):
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._buf = self._local()
            if with_collective:
                dist.all_reduce(self._buf, op=dist.ReduceOp.SUM, group=self.group)
                for p, g in zip(self._shared, unpack_grads(self._buf[:-1], self._shared)):
                    p.grad.copy_(g)
            self._opt_in_graph = False
            if self.optimizer is not None and (with_collective or not self._distributed):
                self.optimizer.step()
                self._opt_in_graph = True

    def _local(self):
        for p in list(self.frame_params.values()) + list(self.shared_params.values()):
            p.grad = None
        local = self.loss_fn(self.frame_params, self.shared_params) / self.num_frames_total
        local.backward()
        return torch.cat([pack_grads([p.grad for p in self._shared], self._shared),
                          local.detach().reshape(1).to(torch.float32)])

    def __call__(self) -> torch.Tensor:
        """Returns the GLOBAL loss as a view of the step's static buffer (valid until the next call)."""
        self.graph.replay()
        for p, g in zip(self._leaves, self._grads):   # re-attach (a caller may have reset .grad since the capture)
            p.grad = g
        buf = self._buf
        if self._distributed and not self.collective_in_graph:
            dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
            with torch.no_grad():
                for p, g in zip(self._shared, unpack_grads(buf[:-1], self._shared)):
                    p.grad.copy_(g)
        if self.optimizer is not None and not self._opt_in_graph:
            self.optimizer.step()
        return buf[-1]


def smal_joint_loss_fn(model, cam_R, cam_T, target_yx, visibility, joint_ids=None, image_size=512, w_j2d=1.0,
                       propagate_scaling=False, get_skin=False):
    """``loss_fn`` for :class:`ShardedFitterStep` built from the CUDA path: SMAL forward (joints) +
    the fitter's 2-D joint loss (fitter.py:262-296).  ``frame_params``: ``global_rotation`` (n,3),
    ``joint_rotations`` (n,J-1,3), ``trans`` (n,3);  ``shared_params``: ``betas`` (nB,), ``fov`` (1,),
    optional ``log_beta_scales`` / ``betas_trans`` (1,J,3).  Returns the SUM over the local frames."""
    from . import losses

    def fn(fp, sp):
        n = fp["trans"].shape[0]
        theta = torch.cat([fp["global_rotation"].unsqueeze(1), fp["joint_rotations"]], dim=1)
        ls = sp["log_beta_scales"].expand(n, -1, -1) if "log_beta_scales" in sp else None
        bt = sp["betas_trans"].expand(n, -1, -1) if "betas_trans" in sp else None
        out = model(sp["betas"].unsqueeze(0).expand(n, -1), theta, betas_logscale=ls, betas_trans=bt,
                    propagate_scaling=propagate_scaling, get_skin=get_skin)   # get_skin=True also poses the mesh (fitter.py:262)
        joints = out[1] if get_skin else out
        joints = joints + fp["trans"].unsqueeze(1)                     # p3d_renderer.py:131
        if joint_ids is not None:
            joints = joints[:, joint_ids]                              # fitter.py:283
        loss, _ = losses.fitter_joint_loss(joints, cam_R, cam_T, sp["fov"].expand(n), target_yx, visibility,
                                           w_j2d=w_j2d, image_size=image_size)
        # fitter_joint_loss is a mean over n*Jc*2 elements: turn it back into the per-frame sum
        return loss * n

    return fn
