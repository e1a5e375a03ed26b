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
    for b in batches:
        done = pipe.submit(b["beta"], b["theta"], b["trans"])
        if done is not None:
            got.append({k: v.clone() for k, v in done.items()})
    got += [{k: v.clone() for k, v in d.items()} for d in pipe.drain()]
    assert len(got) == len(batches)
    for b, g in zip(batches, got):
        t = {k: v.to(dev).requires_grad_(True) for k, v in b.items()}
        verts, joints, _, _ = model(t["beta"], t["theta"], trans=t["trans"])
        gb, gt, gtr = torch.autograd.grad([verts, joints], [t["beta"], t["theta"], t["trans"]], [gv, gj])
        for name, ref in (("grad_beta", gb), ("grad_theta", gt), ("grad_trans", gtr), ("joints", joints.detach())):
            assert torch.equal(g[name], ref.cpu()), name
