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
                 propagate_scaling=False, shared_betas=False, warmup=3, packed_io=False, pack_joints=True):
        dev = model._device
        J, V = model._tables["J"], model._tables["V"]
        nb = model.num_betas if num_betas is None else num_betas
        opts = dict(device=dev, dtype=torch.float32)
        self.model, self.batch = model, batch
        self.shared_betas = shared_betas
        nbeta = (1 if shared_betas else batch) * nb
        # packed_io: beta | theta | trans are views of ONE device buffer (one host->device copy feeds a step) and the
        # parameter gradients + posed joints are gathered into ONE device buffer inside the captured graph
        self.packed_in = torch.zeros(nbeta + batch * J * 3 + batch * 3, **opts) if packed_io else None
        if packed_io:
            o1, o2 = nbeta, nbeta + batch * J * 3
            self.beta = self.packed_in[:o1].view(1 if shared_betas else batch, nb).requires_grad_(True)
            self.theta = self.packed_in[o1:o2].view(batch, J, 3).requires_grad_(True)
            self.trans = self.packed_in[o2:].view(batch, 3).requires_grad_(True)
        else:
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
        self.packed_out = None
        self.out_layout = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            outs, grads = self._run()
            if packed_io:   # (grad_beta | grad_theta | grad_trans | joints) contiguous, gathered on the device
                # pack_joints=False: only the parameter gradients travel back (the posed joints are consumed by the
                # loss on the device; a fit that wants them on the host keeps the default)
                names = ["grad_beta", "grad_theta", "grad_trans"] + (["joints"] if pack_joints else [])
                parts = [grads[0], grads[1], grads[2]] + ([outs[1]] if pack_joints else [])
                self.packed_out = torch.cat([t.reshape(-1) for t in parts])
                self.out_layout = [(n, tuple(t.shape)) for n, t in zip(names, parts)]
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
    stream of batches runs at ``max(copy, compute)`` per batch instead of their sum.  Per step there is ONE
    host->device copy (beta | theta | trans packed in one pinned buffer) and ONE device->host copy (parameter
    gradients | posed joints, gathered on the device inside the captured graph).  Upstream gradients
    (``slot.g_verts`` / ``slot.g_joints``) and the posed vertices stay on the device -- in a fit the loss lives there.

        pipe = HostPipeline(model, B)
        buf, (beta, theta, trans) = pipe.new_input() # pinned, packed: fill the views per batch ...
        for batch in batches:
            done = pipe.submit_packed(buf)          # ... or pipe.submit(beta, theta, trans) with three pinned tensors
            if done is not None: consume(done)      # dict of pinned CPU views of an EARLIER batch
        for done in pipe.drain(): consume(done)

    A dict handed out by ``submit`` / ``drain`` stays valid until the SECOND following ``submit`` that reuses its slot:
    every slot owns two host output sets and alternates between them, so the copies queued by the submit that returns
    ``done`` never touch ``done``'s memory.
    """

    OUT = ("grad_beta", "grad_theta", "grad_trans", "joints")

    def __init__(self, model, batch, depth=2, **step_kwargs):
        dev = model._device
        self.slots = []
        for _ in range(depth):
            st = torch.cuda.Stream(device=dev)
            with torch.cuda.stream(st):
                gs = GraphedStep(model, batch, packed_io=True, **step_kwargs)
            st.synchronize()
            n_out = gs.packed_out.numel()
            hosts = []
            for _h in range(2):
                flat = torch.empty(n_out, dtype=torch.float32).pin_memory()
                flat.zero_()      # first touch after the rank's CPU affinity is set: pages land on the local NUMA node
                views, o = {}, 0
                for name, shape in gs.out_layout:
                    n = 1
                    for d in shape:
                        n *= d
                    views[name] = flat[o:o + n].view(shape)
                    o += n
                hosts.append((flat, views))
            self.slots.append(dict(step=gs, stream=st, event=torch.cuda.Event(), hosts=hosts, cur=0, busy=False))
        self._n = 0
        self.h2d_bytes = self.slots[0]["step"].packed_in.numel() * 4
        self.d2h_bytes = self.slots[0]["hosts"][0][0].numel() * 4

    def new_input(self):
        """A pinned host buffer in the packed input layout plus its ``(beta, theta, trans)`` views: fill the views,
        hand the buffer to :meth:`submit_packed` (ONE host->device copy per step, nothing staged on the host)."""
        gs = self.slots[0]["step"]
        flat = torch.empty(gs.packed_in.numel(), dtype=torch.float32).pin_memory()
        flat.zero_()
        o1 = gs.beta.numel()
        o2 = o1 + gs.theta.numel()
        return flat, (flat[:o1].view(gs.beta.shape), flat[o1:o2].view(gs.theta.shape), flat[o2:].view(gs.trans.shape))

    def _launch(self, copy_in):
        s = self.slots[self._n % len(self.slots)]
        self._n += 1
        done = None
        if s["busy"]:
            s["event"].synchronize()      # the batch that used this slot is complete: its outputs are in hosts[cur]
            done = s["hosts"][s["cur"]][1]
            s["cur"] ^= 1                 # the new batch writes the OTHER host set
        gs = s["step"]
        with torch.cuda.stream(s["stream"]), torch.no_grad():
            copy_in(gs)
            gs.replay()
            s["hosts"][s["cur"]][0].copy_(gs.packed_out, non_blocking=True)
            s["event"].record(s["stream"])
        s["busy"] = True
        return done

    def submit_packed(self, packed):
        """``packed``: a pinned buffer from :meth:`new_input` (the caller keeps it untouched until the step has run)."""
        return self._launch(lambda gs: gs.packed_in.copy_(packed, non_blocking=True))

    def submit(self, beta, theta, trans):
        """Three separate (ideally pinned) host tensors: three host->device copies into the packed device buffer."""
        def copy_in(gs):
            gs.beta.copy_(beta, non_blocking=True)
            gs.theta.copy_(theta, non_blocking=True)
            gs.trans.copy_(trans, non_blocking=True)
        return self._launch(copy_in)

    def drain(self):
        out = []
        n = len(self.slots)
        for i in range(n):   # oldest first
            s = self.slots[(self._n + i) % n]
            if s["busy"]:
                s["event"].synchronize()
                s["busy"] = False
                out.append(s["hosts"][s["cur"]][1])
        return out
