"""HostPipeline (graphed steps fed from pinned host memory on alternating streams) returns, batch by
batch and in order, exactly what the eager ``SMAL`` call + autograd returns."""
import numpy as np
import pytest
import torch

from tests import rigs

pytestmark = pytest.mark.gpu


def test_host_pipeline_matches_eager():
    import smilify_b200
    from smilify_b200.graphed import HostPipeline
    from smilify_b200.smal_torch import SMAL

    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    dev = torch.device("cuda:0")
    model = SMAL(dev)
    B, V, J = 37, model._tables["V"], model._tables["J"]
    dd = rigs.get_rig("stick")
    gv = torch.randn(B, V, 3, device=dev)
    gj = torch.randn(B, J, 3, device=dev)
    pipe = HostPipeline(model, B, depth=2)
    for sl in pipe.slots:
        with torch.no_grad():
            sl["step"].g_verts.copy_(gv); sl["step"].g_joints.copy_(gj)
    batches = []
    for i in range(5):
        inp = rigs.make_inputs(dd, B, seed=100 + i)
        batches.append({k: torch.from_numpy(inp[k]).pin_memory() for k in ("beta", "theta", "trans")})
    got = []
    # a slow consumer: the previous round's dict is read only AFTER the following submit has queued its copies and a
    # long GPU operation has run (a slot must never write into the host set it has just handed out)
    pending = None
    big = torch.randn(4096, 4096, device=dev)
    bufs = [pipe.new_input() for _ in range(2)]
    for i, b in enumerate(batches):
        if i % 2 == 0:
            done = pipe.submit(b["beta"], b["theta"], b["trans"])
        else:   # the packed route: one host->device copy
            flat, (hb, ht, htr) = bufs[(i // 2) % 2]
            hb.copy_(b["beta"]); ht.copy_(b["theta"]); htr.copy_(b["trans"])
            done = pipe.submit_packed(flat)
        if pending is not None:
            (big @ big).sum().item()
            got.append({k: v.clone() for k, v in pending.items()})
        pending = done
    if pending is not None:
        got.append({k: v.clone() for k, v in pending.items()})
    got += [{k: v.clone() for k, v in d.items()} for d in pipe.drain()]
    assert len(got) == len(batches)
    for b, g in zip(batches, got):
        t = {k: v.to(dev).requires_grad_(True) for k, v in b.items()}
        verts, joints, _, _ = model(t["beta"], t["theta"], trans=t["trans"])
        gb, gt, gtr = torch.autograd.grad([verts, joints], [t["beta"], t["theta"], t["trans"]], [gv, gj])
        for name, ref in (("grad_beta", gb), ("grad_theta", gt), ("grad_trans", gtr), ("joints", joints.detach())):
            # same kernels, same inputs: equal up to the summation order of the backward's vector reductions
            assert torch.allclose(g[name], ref.cpu(), rtol=2e-6, atol=2e-6 * float(ref.abs().max())), name
