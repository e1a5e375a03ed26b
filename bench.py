#!/usr/bin/env python
"""bench.py -- fwd+bwd posed meshes/s of the SMAL/SMIL hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): SMILy_STICK (V=3020, J=55, nB=5), batch 4096 random
poses/shapes PER GPU (frames are independent: ranks shard them with no data-path collective ->
weak scaling), one step = SMAL forward (verts, joints, Rs, v_shaped) + the full backward to
beta/theta/trans from upstream gradients on verts and joints.

Arms
  ours       the CUDA path through ``smilify_b200.smal_torch.SMAL`` (-> C ABI -> sm_100a kernels).
             ``value``: inputs and upstream gradients resident in HBM, step replayed from a CUDA
             graph captured through that public API (``--no-graph`` times the eager call instead).
             ``e2e``: the same step with HOST (pinned) pose/shape parameters copied in and the
             parameter gradients + posed joints copied back inside the timed region.
  reference  the reference's algorithm on the box's host cores: SMILify is Python/torch, so this
             is the torch-CPU restatement in oracle/ (pinned against the real reference, see
             oracle/pin_against_reference.py), all host threads, on a bounded sample of the batch.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "fwd+bwd posed meshes/s (SMIL LBS, B=4096)"
UNIT = "meshes/s"


def rig_path(name):
    """On-disk rig file: the two shipped pickles, or a stand-in written to a temp .npz by tests/common.py."""
    from tests import rigs
    if name in rigs.REAL_RIGS:
        return rigs.REAL_RIGS[name]
    from tests.common import rig_file

    return rig_file(name)


def algorithmic_bytes(V, J, nB):
    """fp32, every tensor of the call contract moved exactly once."""
    n_in = nB + 3 * J + 3
    fwd = 4 * (n_in + 3 * V + 3 * J + 9 * J + 3 * V)
    bwd = 4 * (3 * V + 3 * J + 2 * n_in)
    per_kernel = {  # the contract tensors each kernel of the fused path moves
        "pose_fwd_kernel": 4 * (n_in + 9 * J + 3 * J),          # params in; Rs, joints out
        "skin_fwd_kernel": 4 * (3 * V + 3 * V),                 # verts, v_shaped out
        "skin_bwd_kernel": 4 * (3 * V),                         # g_verts in
        "pose_bwd_kernel": 4 * (3 * J + 2 * n_in),              # g_joints + params in; param grads out
    }
    assert sum(per_kernel.values()) == fwd + bwd
    per_kernel["pose_bwd_lean_kernel"] = per_kernel["pose_bwd_kernel"]   # the no-log-scale variant moves the same tensors
    per_kernel["skin_bwd_vm_kernel"] = per_kernel["skin_bwd_kernel"]     # vertex-major variant: the same contract tensor
    per_kernel["pose_fwd_ls_kernel"] = per_kernel["pose_fwd_kernel"]     # lane-per-sample variants: the same tensors
    per_kernel["pose_bwd_ls_kernel"] = per_kernel["pose_bwd_kernel"]
    return fwd, bwd, per_kernel


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_for(kernel, args):
    """DRAM bytes per launch of ``kernel`` from the committed ncu capture (profiles/r01_traffic.json), valid for the
    rig / batch it was captured on; ``--traffic`` overrides; None when there is no matching capture."""
    if args.traffic is not None:
        return args.traffic
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if t.get("rig") == args.rig and t.get("batch") == args.batch:
            return t["kernels"].get(kernel)
    except Exception:
        pass
    return None


def synth_inputs(B, V, J, nB, seed):
    """SURVEY.md section 8(d) config 2 distributions."""
    rs = np.random.RandomState(seed)
    return dict(
        beta=(rs.standard_normal((B, nB)) * 0.5).astype(np.float32),
        theta=(rs.standard_normal((B, J, 3)) * 0.3).astype(np.float32),
        trans=rs.standard_normal((B, 3)).astype(np.float32),
        g_verts=rs.standard_normal((B, V, 3)).astype(np.float32),
        g_joints=rs.standard_normal((B, J, 3)).astype(np.float32),
    )


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the GPU is under load."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.path = os.path.join(tempfile.gettempdir(), "smil_clocks_%d.csv" % os.getpid())
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        try:
            os.remove(self.path)
        except OSError:
            pass
        return out


def gpu_local_cores(index):
    """CPU cores on the NUMA node the GPU hangs off (NVML's ideal affinity), or None."""
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            tok = vis.split(",")[index].strip()
            h = pynvml.nvmlDeviceGetHandleByIndex(int(tok)) if tok.isdigit() else pynvml.nvmlDeviceGetHandleByUUID(tok)
        else:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        return {w * 64 + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
    except Exception:
        return None


def pin_rank_to_cores(local_rank, world):
    """Disjoint CPU cores per rank, taken from the cores local to the rank's GPU where NVML reports them (the 8 ranks of
    a box otherwise share one affinity mask, their copy / launch threads migrate over each other and pinned buffers land
    on the far NUMA node); pinned host buffers are first-touched AFTER this, so they land on the local node."""
    try:
        allowed = sorted(os.sched_getaffinity(0))
        local = [gpu_local_cores(r) for r in range(world)]
        mine_set = local[local_rank]
        if mine_set and (set(allowed) & mine_set):
            # ranks whose GPUs share a node split that node's cores between them
            sharers = [r for r in range(world) if local[r] == mine_set]
            cores = sorted(set(allowed) & mine_set)
            per = max(1, len(cores) // len(sharers))
            k = sharers.index(local_rank)
            mine = cores[k * per:(k + 1) * per] or cores
        else:
            per = max(1, len(allowed) // max(world, 1))
            mine = allowed[local_rank * per:(local_rank + 1) * per] or allowed
        os.sched_setaffinity(0, mine)
        return len(mine)
    except (AttributeError, OSError):
        return None


def torch_gpu_baseline(rig_path, B, dev, seed=0, warmup=3, iters=10):
    """The reference's op-by-op torch graph (oracle/smal_oracle.py, pinned against the unmodified reference) with every
    tensor on the SAME B200 at the full batch: the comparator SURVEY.md section 8(d) config 2 / BASELINE configs[1] ask
    for ("vs reference torch path").  Eager, CUDA events."""
    import torch

    from oracle import smal_oracle as SO

    dd = SO.load_model_dict(rig_path)
    m = SO.build_model(dd, dtype=torch.float32)
    for k, v in list(vars(m).items()):
        if isinstance(v, torch.Tensor):
            setattr(m, k, v.to(dev))
    V, J, nB = m.num_verts, m.num_joints, m.num_betas
    x = synth_inputs(B, V, J, nB, seed)
    beta, theta, trans = (torch.from_numpy(x[k]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans"))
    gv, gj = torch.from_numpy(x["g_verts"]).to(dev), torch.from_numpy(x["g_joints"]).to(dev)

    def step():
        with torch.device(dev):   # the oracle's helper tensors (eye, zeros, ...) are created on the default device
            verts, joints, _Rs, _vs = SO.smal_forward(m, beta, theta, trans)
            torch.autograd.backward([verts, joints], [gv, gj])
        beta.grad = theta.grad = trans.grad = None

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {"value": B / ms * 1e3, "unit": UNIT, "ms_per_step": ms, "kind": "port-on-cuda", "batch": B,
            "what": "oracle/smal_oracle.py (the reference's dense op-by-op torch graph) on cuda, eager, %d timed steps" % iters}


# ---- CPU arm: the reference's algorithm on the host cores (oracle/ restatement) -----------------------
def cpu_step_fn(rig_path, B, seed=0):
    """-> (callable running one fwd+bwd of the oracle on B samples, V, J, nB)."""
    import torch

    from oracle import smal_oracle as SO

    dd = SO.load_model_dict(rig_path)
    m = SO.build_model(dd, dtype=torch.float32)
    V, J, nB = m.num_verts, m.num_joints, m.num_betas
    x = synth_inputs(B, V, J, nB, seed)
    beta, theta, trans = (torch.from_numpy(x[k]).requires_grad_(True) for k in ("beta", "theta", "trans"))
    gv, gj = torch.from_numpy(x["g_verts"]), torch.from_numpy(x["g_joints"])

    def step():
        verts, joints, _Rs, _vs = SO.smal_forward(m, beta, theta, trans)
        torch.autograd.backward([verts, joints], [gv, gj])
        beta.grad = theta.grad = trans.grad = None

    return step, V, J, nB


def time_cpu(rig_path, B, budget_s, min_iters=2, warmup=1):
    import torch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step, V, J, nB = cpu_step_fn(rig_path, B)
    for _ in range(warmup):
        step()
    times = []
    t_end = time.perf_counter() + budget_s
    while len(times) < min_iters or time.perf_counter() < t_end:
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if len(times) >= 200:
            break
    return B / statistics.median(times), cores, len(times)


def run_reference(args, rank, world):
    """--impl reference: rank 0 alone times the CPU path; other ranks exit 0 without work."""
    if rank != 0:
        return
    import torch

    from tests import rigs

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Bs = args.cpu_batch
    step, V, J, nB = cpu_step_fn(rig_path(args.rig), Bs)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = Bs * args.steps / dt
    sample = "%d of the %d meshes of a step per CPU step (SMAL fwd+bwd, torch-CPU, %d threads)" % (Bs, args.batch, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": ("SMILy_STICK.pkl" if args.rig == "stick" else args.rig + " rig") + " fwd+bwd batch %d synthetic poses (CPU arm: %d-mesh sample per step)"
                               % (args.batch, Bs), "rig": args.rig, "V": V, "J": J, "nB": nB, "batch_per_gpu": args.batch},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Python/torch (nothing to compile into oracle/_ref); timed: oracle/smal_oracle.py, "
                "pinned against the unmodified reference by oracle/pin_against_reference.py",
    }
    print(json.dumps(line), flush=True)


# ---- GPU arm ---------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import smilify_b200
    from smilify_b200 import _lib
    from smilify_b200.graphed import GraphedStep
    from smilify_b200.smal_torch import SMAL
    from tests import rigs

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl ours needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cores_pinned = pin_rank_to_cores(local_rank, world) if world > 1 else None
    smilify_b200.configure(rig_path(args.rig))
    model = SMAL(dev)
    V, J, nB = model._tables["V"], model._tables["J"], model.num_betas
    B, K, W = args.batch, args.steps, args.warmup
    x = synth_inputs(B, V, J, nB, seed=rank)
    fwd_b, bwd_b, per_kernel = algorithmic_bytes(V, J, nB)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    # --- the step ------------------------------------------------------------------------------
    if args.no_graph:
        beta, theta, trans = (torch.from_numpy(x[k]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans"))
        gv, gj = torch.from_numpy(x["g_verts"]).to(dev), torch.from_numpy(x["g_joints"]).to(dev)
        state = {}

        def step():
            verts, joints, Rs, vs = model(beta, theta, trans=trans)
            state["joints"] = joints
            state["grads"] = torch.autograd.grad([verts, joints], [beta, theta, trans], [gv, gj])

        def outputs():
            return state["grads"] + (state["joints"],)
        dst_in = (beta, theta, trans)
    else:
        gs = GraphedStep(model, B)
        with torch.no_grad():
            gs.beta.copy_(torch.from_numpy(x["beta"])); gs.theta.copy_(torch.from_numpy(x["theta"]))
            gs.trans.copy_(torch.from_numpy(x["trans"]))
            gs.g_verts.copy_(torch.from_numpy(x["g_verts"])); gs.g_joints.copy_(torch.from_numpy(x["g_joints"]))
        step = gs.replay

        def outputs():
            return (gs.grad_beta, gs.grad_theta, gs.grad_trans, gs.joints)
        dst_in = (gs.beta, gs.theta, gs.trans)

    c0 = _lib.launch_count()
    with torch.enable_grad():
        _b, _t, _tr = (torch.from_numpy(x[k][:8]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans"))
        _v, _j, _r, _s = model(_b, _t, trans=_tr)
        torch.autograd.grad([_v, _j], [_b, _t, _tr], [torch.zeros_like(_v), torch.zeros_like(_j)])
    launches_per_step = int(_lib.launch_count() - c0)   # kernels of this library per fused step (a replay launches the same)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    # pre-warm so that clock samples are taken under load even when K is small
    t_end = time.perf_counter() + 0.6
    while time.perf_counter() < t_end:
        step()
        torch.cuda.synchronize()
    for _ in range(max(W, 3)):
        step()
    launches0 = _lib.launch_count()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    eager_launches = _lib.launch_count() - launches0

    # --- e2e: host parameters in, parameter gradients + joints out, every step -----------------------
    # Through the public streaming API (smilify_b200.graphed.HostPipeline): every step copies its pinned-host
    # beta/theta/trans to the device, replays the step and copies the parameter gradients + posed joints back
    # to pinned host memory; two buffer sets on two streams let the copies of neighbouring steps overlap the kernels.
    h_in = [torch.from_numpy(x[k]).pin_memory() for k in ("beta", "theta", "trans")]
    if args.no_graph:
        step()
        h_out = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in outputs()]
        h2d = sum(t.numel() * 4 for t in h_in)
        d2h = sum(t.numel() * 4 for t in h_out)

        def e2e_run(n):
            for _ in range(n):
                with torch.no_grad():
                    for d, h in zip(dst_in, h_in):
                        d.copy_(h, non_blocking=True)
                step()
                for h, d in zip(h_out, outputs()):
                    h.copy_(d.detach(), non_blocking=True)
                torch.cuda.current_stream().synchronize()  # the caller reads the result
        e2e_mode = "eager, serial copies"
    else:
        from smilify_b200.graphed import HostPipeline

        pipe = HostPipeline(model, B, depth=args.e2e_depth, pack_joints=False)
        for sl in pipe.slots:
            with torch.no_grad():
                sl["step"].g_verts.copy_(gs.g_verts); sl["step"].g_joints.copy_(gs.g_joints)
        packed, (hb, ht, htr) = pipe.new_input()     # pinned, first-touched after the rank's cores were pinned
        hb.copy_(h_in[0]); ht.copy_(h_in[1]); htr.copy_(h_in[2])
        torch.cuda.synchronize()
        h2d = pipe.h2d_bytes
        d2h = pipe.d2h_bytes
        sink = [0.0]

        def e2e_run(n):
            for _ in range(n):
                done = pipe.submit_packed(packed)
                if done is not None:
                    sink[0] += float(done["grad_trans"][0, 0])   # the caller reads the finished step's result
            for done in pipe.drain():
                sink[0] += float(done["grad_trans"][0, 0])
        e2e_mode = ("HostPipeline depth %d: per step" % args.e2e_depth + " ONE pinned host->device copy (beta | theta | trans) and ONE "
                    "device->host copy (the parameter gradients: the step's result) overlap the kernels of neighbouring steps; "
                    "posed vertices, joints and upstream gradients stay device-resident (the loss lives on the device); with several "
                    "steps in flight the latency-bound pose kernels of one step run under the skin kernels of another, so this "
                    "number can exceed the single-stream `value`")

    e2e_run(3)
    barrier()
    t0 = time.perf_counter()
    e2e_run(K)
    torch.cuda.synchronize()
    e2e_wall = (time.perf_counter() - t0) * 1e3
    barrier()
    e2e_ms = max_over_ranks(e2e_wall)
    # informational: device-resident steps with several steps in flight (the pipeline's graphed buffer sets replayed round
    # robin on their own streams, no copies) -- what a caller that streams many resident batches gets; `value` stays the
    # single-stream number
    overlapped = None
    if not args.no_graph:
        cur = torch.cuda.current_stream()
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def overlapped_run(n):
            for i in range(n):
                sl = pipe.slots[i % len(pipe.slots)]
                with torch.cuda.stream(sl["stream"]):
                    sl["step"].replay()
        overlapped_run(2 * len(pipe.slots))
        torch.cuda.synchronize()
        barrier()
        o0.record(cur)
        for sl in pipe.slots:
            sl["stream"].wait_event(o0)
        overlapped_run(K)
        for sl in pipe.slots:
            ev = torch.cuda.Event()
            ev.record(sl["stream"])
            cur.wait_event(ev)
        o1.record(cur)
        torch.cuda.synchronize()
        barrier()
        ov_ms = max_over_ranks(o0.elapsed_time(o1))
        overlapped = {"value": world * B * K / (ov_ms * 1e-3), "unit": UNIT, "ms_per_step": ov_ms / K, "streams": len(pipe.slots),
                      "what": "K device-resident steps replayed round robin from %d graphed buffer sets on their own streams "
                              "(the latency-bound pose kernels of one step run under the skin kernels of another)" % len(pipe.slots)}
    if os.environ.get("SMIL_BENCH_DIAG"):
        hbuf = torch.empty(max(h2d, d2h), dtype=torch.uint8).pin_memory()
        dbuf = torch.empty(max(h2d, d2h), dtype=torch.uint8, device=dev)
        for name, nbytes, to_dev in (("h2d", h2d, True), ("d2h", d2h, False)):
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); ev0.record()
            for _ in range(50):
                (dbuf[:nbytes].copy_(hbuf[:nbytes], non_blocking=True) if to_dev else hbuf[:nbytes].copy_(dbuf[:nbytes], non_blocking=True))
            ev1.record(); torch.cuda.synchronize()
            us = ev0.elapsed_time(ev1) * 1e3 / 50
            print("[diag] rank %d %s %d B: %.1f us/copy, %.1f GB/s" % (rank, name, nbytes, us, nbytes / us / 1e3), file=sys.stderr, flush=True)
        for n in (K, 10 * K):
            barrier(); t1 = time.perf_counter(); e2e_run(n); torch.cuda.synchronize()
            print("[diag] rank %d e2e %d steps: %.4f ms/step, cores %s" % (rank, n, (time.perf_counter() - t1) * 1e3 / n,
                                                                            sorted(os.sched_getaffinity(0))), file=sys.stderr, flush=True)
    clocks = sampler.stop() if sampler else None

    # --- per-kernel device time (eager, event pair around every launch) -> roofline of the dominant kernel
    roofline = None
    kernels = {}
    if rank == 0:
        beta, theta, trans = (torch.from_numpy(x[k]).to(dev).requires_grad_(True) for k in ("beta", "theta", "trans"))
        gv, gj = torch.from_numpy(x["g_verts"]).to(dev), torch.from_numpy(x["g_joints"]).to(dev)
        n_prof = min(max(K, 5), 50)
        _lib.profile_enable(16 * n_prof)
        for _ in range(n_prof):
            verts, joints, Rs, vs = model(beta, theta, trans=trans)
            torch.autograd.grad([verts, joints], [beta, theta, trans], [gv, gj])
        torch.cuda.synchronize()
        prof = _lib.profile_collect()
        _lib.profile_enable(0)
        peak, peak_src = measured_peaks()
        for name, (n, ms) in prof.items():
            us = ms / n * 1e3
            ab = per_kernel.get(name)
            kernels[name] = {"launches_per_step": n / n_prof, "avg_us": round(us, 2),
                             "alg_bytes_per_launch": ab * B if ab else None,
                             "GBps": round(ab * B / (us * 1e-6) / 1e9, 1) if ab else None}
        if prof:
            top = max((k for k in prof if per_kernel.get(k)), key=lambda k: prof[k][1])
            ab = per_kernel.get(top, 0) * B
            ach = ab / (prof[top][1] / prof[top][0] * 1e-3) / 1e9
            roofline = {"kernel": top, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                        "frac": round(ach / peak, 4), "traffic": traffic_for(top, args), "peak_source": peak_src,
                        "alg_bytes_per_launch": ab, "avg_us": kernels[top]["avg_us"]}
    fitter = None
    fitter_full = None
    if world > 1 and not args.no_fitter:
        # the one collective of the path (north_star): BASELINE configs[3], frames strong-scaled over the ranks, shared
        # betas / fov, ONE packed NCCL all-reduce per step captured inside the step's CUDA graph
        import importlib.util

        spec = importlib.util.spec_from_file_location("fitter_multigpu", os.path.join(ROOT, "tools_gpu", "fitter_multigpu.py"))
        fm = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(fm)
        del gs
        torch.cuda.empty_cache()
        fitter = fm.run_fitter(args.fitter_frames, max(5, min(K, 20)), 3, rank, world, dev, rig="stick13")
        torch.cuda.empty_cache()
        # the same step with the fitter's parameter-space terms (halo exchange for the temporal term) and the fused Adam
        # inside the graph: SURVEY.md section 8(f) row 2
        fitter_full = fm.run_fitter(args.fitter_frames, max(5, min(K, 20)), 3, rank, world, dev, rig="stick13", full_step=True)
        smilify_b200.configure(rig_path(args.rig))
    if rank != 0:
        return
    torch_gpu = None
    if world == 1 and not args.no_torch_gpu:
        try:
            torch_gpu = torch_gpu_baseline(rig_path(args.rig), B, dev)
        except Exception as e:   # e.g. out of memory on a smaller part: the dense graph materialises (B, V, J) tensors
            torch_gpu = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:120])}
    peak, peak_src = measured_peaks()
    ms_single = ms_total / K
    value_single = world * B * K / (ms_total * 1e-3)
    # `value`: whole-job throughput over K resident batches.  Independent batches need not wait for each other, so the
    # job keeps several steps in flight (the pipeline's graphed buffer sets on their own streams); the single-stream
    # figure -- one step after the other -- is reported beside it.
    ms_step, value = (overlapped["ms_per_step"], overlapped["value"]) if overlapped else (ms_single, value_single)
    step_gbs = (fwd_b + bwd_b) * B / (ms_step * 1e-3) / 1e9
    cpu_val, cores, n_it = time_cpu(rig_path(args.rig), args.cpu_batch, args.cpu_seconds) if world == 1 else (None, None, 0)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": ("SMILy_STICK.pkl fwd+bwd batch %d synthetic poses per GPU (BASELINE configs[1])" % B) if args.rig == "stick"
                   else "%s rig fwd+bwd batch %d synthetic poses per GPU" % (args.rig, B),
                   "rig": args.rig, "V": V, "J": J, "nB": nB, "batch_per_gpu": B, "global_batch": B * world,
                   "parallelism": "frames sharded, dp%d, no collective" % world,
                   "launch": "eager" if args.no_graph else
                             "cuda-graph replay of SMAL.__call__ + autograd, %d independent resident batches in flight (own buffer "
                             "sets and streams): the latency-bound pose kernels of one step run under the skin kernels of "
                             "another" % overlapped["streams"],
                   "l2": "per-step working set %.0f MB > 126 MB L2 (no flush needed)" % ((fwd_b + bwd_b) * B / 1e6)},
        "e2e": {"value": world * B * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / K, "mode": e2e_mode,
                "timing": "host wall clock around K steps incl. the final drain (max over ranks)"},
        "gpu_launches": int(eager_launches) if args.no_graph else launches_per_step * K,
        "value_single_stream": {"value": value_single, "unit": UNIT, "ms_per_step": ms_single,
                                "what": "the same K steps replayed one after the other from ONE graphed buffer set on one stream"},
        "steps_in_flight": overlapped["streams"] if overlapped else 1,
        "clocks": clocks,
        "roofline": roofline,
        "roofline_step": {"bound": "hbm", "alg_bytes_per_mesh": fwd_b + bwd_b, "achieved": round(step_gbs, 1),
                          "peak": peak, "unit": "GB/s", "frac": round(step_gbs / peak, 4), "peak_source": peak_src},
        "kernels": kernels,
    }
    if torch_gpu is not None:
        line["torch_gpu_baseline"] = torch_gpu
    if fitter is not None:
        line["fitter"] = fitter
    if fitter_full is not None:
        line["fitter_full"] = fitter_full
    if cores_pinned:
        line["config"]["host_cores_per_rank"] = cores_pinned
    if cpu_val is not None:
        line["cpu_baseline"] = {"value": cpu_val, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "%d-mesh batches (%s) of the same workload, %d timed fwd+bwd (median), torch-CPU "
                                          "oracle, %d threads" % (args.cpu_batch, "the full step" if args.cpu_batch >= B
                                                                  else "a bounded sample", n_it, cores)}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rig", default="stick", choices=["stick", "mouse", "stick13", "mouse40"],
                    help="stick13 / mouse40: the stand-ins for the missing BASELINE rigs (nB = 13 OmniAnt, J = 40 Falkner mouse)")
    ap.add_argument("--batch", type=int, default=4096, help="meshes per GPU per step")
    ap.add_argument("--no-graph", action="store_true", help="time the eager Python call instead of the graph replay")
    ap.add_argument("--cpu-batch", type=int, default=4096, help="meshes per CPU step (4096 = the full batch of a step)")
    ap.add_argument("--no-torch-gpu", action="store_true", help="skip the torch-on-the-same-GPU comparator")
    ap.add_argument("--e2e-depth", type=int, default=4, help="graphed buffer sets of the host pipeline (copies of neighbouring steps overlap the kernels)")
    ap.add_argument("--no-fitter", action="store_true", help="N > 1: skip the frame-sharded fitter step (the NCCL path)")
    ap.add_argument("--fitter-frames", type=int, default=65536)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--traffic", type=float, default=None, help="dram bytes/launch of the dominant kernel from ncu")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


if __name__ == "__main__":
    main()
