"""CUDA-graph capture of one SMAL forward+backward step.

At B=4096 the four kernels of a fused step take ~0.1-0.5 ms on a B200 while the Python side
of an eager ``SMAL.__call__`` + ``autograd`` round trip (ctypes marshalling, output allocation,
autograd bookkeeping) costs about as much, so an optimisation loop that replays the same shapes
(SMALFitter epochs, fitter_3d stages) should capture the step once and replay it:

    step = GraphedStep(model, B, with_scale=False)          # captures on the current device
    step.beta.copy_(...); step.theta.copy_(...); step.trans.copy_(...)
    step.g_verts.copy_(...); step.g_joints.copy_(...)       # upstream gradients
    step.replay()                                           # verts/joints/Rs/v_shaped + grads updated in place
    step.grad_theta ...

The captured work is exactly what the eager call launches (the same ``SmalFunction``); nothing here
changes numerics."""
from __future__ import annotations

import torch


class GraphedStep:
    def __init__(self, model, batch, num_betas=None, with_scale=False, with_trans_offsets=False,
                 propagate_scaling=False, shared_betas=False, warmup=3):
        dev = model._device
        J, V = model._tables["J"], model._tables["V"]
        nb = model.num_betas if num_betas is None else num_betas
        opts = dict(device=dev, dtype=torch.float32)
        self.model, self.batch = model, batch
        self.shared_betas = shared_betas
        self.beta = torch.zeros((1 if shared_betas else batch, nb), **opts).requires_grad_(True)
        self.theta = torch.zeros((batch, J, 3), **opts).requires_grad_(True)
        self.trans = torch.zeros((batch, 3), **opts).requires_grad_(True)
        self.logscale = torch.zeros((batch, J, 3), **opts).requires_grad_(True) if with_scale else None
        self.btrans = torch.zeros((batch, J, 3), **opts).requires_grad_(True) if with_trans_offsets else None
        self.g_verts = torch.zeros((batch, V, 3), **opts)
        self.g_joints = torch.zeros((batch, J, 3), **opts)
        self._propagate = propagate_scaling
        self._leaves = [t for t in (self.beta, self.theta, self.trans, self.logscale, self.btrans) if t is not None]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._run()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            outs, grads = self._run()
        self.verts, self.joints, self.Rs, self.v_shaped = outs
        it = iter(grads)
        self.grad_beta = next(it)
        self.grad_theta = next(it)
        self.grad_trans = next(it)
        self.grad_logscale = next(it) if with_scale else None
        self.grad_btrans = next(it) if with_trans_offsets else None

    def _run(self):
        beta = self.beta.expand(self.batch, -1) if self.shared_betas else self.beta
        outs = self.model(beta, self.theta, trans=self.trans, betas_logscale=self.logscale, betas_trans=self.btrans,
                          propagate_scaling=self._propagate)
        grads = torch.autograd.grad([outs[0], outs[1]], self._leaves, [self.g_verts, self.g_joints])
        return outs, grads

    def replay(self):
        self.graph.replay()


class HostPipeline:
    """Streams HOST-resident parameter batches through graphed steps.

    ``depth`` device buffer sets, each a :class:`GraphedStep` on its own CUDA stream: the host->device
    copy of batch i+1 and the device->host copy of batch i-1 overlap the kernels of batch i, so a
    stream of batches runs at ``max(copy, compute)`` per batch instead of their sum.  Upstream gradients
    (``slot.g_verts`` / ``slot.g_joints``) stay on the device -- in a fit they come from a loss computed there.

        pipe = HostPipeline(model, B)
        for batch in batches:                       # pinned CPU tensors
            done = pipe.submit(batch.beta, batch.theta, batch.trans)
            if done is not None: consume(done)      # dict of pinned CPU tensors of an EARLIER batch
        for done in pipe.drain(): consume(done)
    """

    OUT = ("grad_beta", "grad_theta", "grad_trans", "joints")

    def __init__(self, model, batch, depth=2, **step_kwargs):
        dev = model._device
        self.slots = []
        for _ in range(depth):
            st = torch.cuda.Stream(device=dev)
            with torch.cuda.stream(st):
                gs = GraphedStep(model, batch, **step_kwargs)
            st.synchronize()
            host = {k: torch.empty(getattr(gs, k).shape, dtype=torch.float32).pin_memory() for k in self.OUT}
            self.slots.append(dict(step=gs, stream=st, event=torch.cuda.Event(), host=host, busy=False))
        self._n = 0
        self.h2d_bytes = 0
        self.d2h_bytes = sum(t.numel() * 4 for t in self.slots[0]["host"].values())

    def submit(self, beta, theta, trans):
        s = self.slots[self._n % len(self.slots)]
        self._n += 1
        done = None
        if s["busy"]:
            s["event"].synchronize()      # the batch that used this slot is complete: its outputs are in s["host"]
            done = s["host"]
        gs = s["step"]
        with torch.cuda.stream(s["stream"]), torch.no_grad():
            gs.beta.copy_(beta, non_blocking=True)
            gs.theta.copy_(theta, non_blocking=True)
            gs.trans.copy_(trans, non_blocking=True)
            gs.replay()
            for k in self.OUT:
                s["host"][k].copy_(getattr(gs, k), non_blocking=True)
            s["event"].record(s["stream"])
        s["busy"] = True
        self.h2d_bytes = (beta.numel() + theta.numel() + trans.numel()) * 4
        return done

    def drain(self):
        out = []
        n = len(self.slots)
        for i in range(n):   # oldest first
            s = self.slots[(self._n + i) % n]
            if s["busy"]:
                s["event"].synchronize()
                s["busy"] = False
                out.append(s["host"])
        return out
