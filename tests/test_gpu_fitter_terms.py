"""SURVEY.md section 8(f) row 2, first half: the fitter's parameter-space terms (limits, pose prior, splay, betas prior,
temporal smoothness) and the fused Adam, through the C ABI, against the outputs of the reference's OWN code
(tests/golden/fitter_prior_pins.npz: fitter.py statements extracted by AST and executed window by window, see
oracle/pin_against_reference.py::pin_fitter_priors) and against torch.optim.Adam."""
import os

import numpy as np
import pytest
import torch

from tests import rigs
from tests.common import rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-5   # fp32 kernel against the fp64 run of the reference code (same bar as the SMAL path)
PINS = np.load(os.path.join(rigs.GOLDEN, "fitter_prior_pins.npz"))
CASES = sorted({k.split("/")[0] for k in PINS.files})


def _case(name):
    c = {k.split("/", 1)[1][3:]: PINS[k] for k in PINS.files if k.startswith(name + "/in_")}
    t = lambda k, grad=False: torch.from_numpy(np.ascontiguousarray(c[k])).to(DEV).requires_grad_(grad)
    wl, wp, ws, wb, wt = (float(x) for x in c["weights"])
    weights = dict(w_limit=wl, w_pose=wp, w_splay=ws, w_betas=wb, w_temp=wt)
    consts = dict(max_limits=t("max_limits"), min_limits=t("min_limits"), betas_mean=t("betas_mean"), betas_prec=t("betas_prec"),
                  rotation_mask=t("rotation_mask"), global_mask=t("global_mask"),
                  pose_prior=dict(mean=t("pose_mean"), precs=t("pose_precs"), use=t("pose_use")))
    return c, t, weights, consts


@pytest.mark.parametrize("name", CASES)
def test_terms_match_the_reference_code(name):
    from smilify_b200.fitter_terms import fitter_prior_terms

    c, t, weights, consts = _case(name)
    leaves = {k: t(k, True) for k in ("global_rot", "joint_rot", "trans", "betas")}
    loss = fitter_prior_terms(leaves["global_rot"], leaves["joint_rot"], leaves["trans"], leaves["betas"], weights,
                              window=int(c["window"]), **consts)
    loss.backward()
    ref = float(PINS[name + "/ref64_total"])
    assert abs(float(loss) - ref) <= TOL * abs(ref), (float(loss), ref)
    for k, v in leaves.items():
        e = rel(v.grad.cpu().numpy(), PINS["%s/ref64_grad_%s" % (name, k)])
        assert e <= TOL, (k, e)


def test_identity_prior_defaults_equal_the_explicit_identity():
    from smilify_b200.fitter_terms import fitter_prior_terms

    name = "fp_n23_j55_w10"   # pinned with the identity prior of fitter.py:25-52 passed explicitly
    c, t, weights, consts = _case(name)
    consts.pop("pose_prior")
    leaves = {k: t(k, True) for k in ("global_rot", "joint_rot", "trans", "betas")}
    loss = fitter_prior_terms(leaves["global_rot"], leaves["joint_rot"], leaves["trans"], leaves["betas"], weights,
                              window=int(c["window"]), **consts)
    loss.backward()
    ref = float(PINS[name + "/ref64_total"])
    assert abs(float(loss) - ref) <= TOL * abs(ref)
    assert rel(leaves["joint_rot"].grad.cpu().numpy(), PINS[name + "/ref64_grad_joint_rot"]) <= TOL


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("cut", [1, 9, 10])
def test_sharded_frames_with_halos_sum_to_the_whole(name, cut):
    """Frames [0, cut) and [cut, N) as two shards with the 1-frame halos of SURVEY.md section 8(e): the losses add up to
    the single-process value and the gradients concatenate to the single-process gradients."""
    from smilify_b200.fitter_terms import fitter_prior_terms

    c, t, weights, consts = _case(name)
    N = c["global_rot"].shape[0]
    cut = min(cut, N - 1)
    full = {k: t(k) for k in ("global_rot", "joint_rot", "trans")}
    halo = lambda i: torch.cat([full["global_rot"][i].reshape(-1), full["joint_rot"][i].reshape(-1), full["trans"][i].reshape(-1)])
    total, grads, gb = 0.0, {k: [] for k in full}, 0.0
    for lo, hi in ((0, cut), (cut, N)):
        leaves = {k: full[k][lo:hi].clone().requires_grad_(True) for k in full}
        betas = t("betas", True)
        loss = fitter_prior_terms(leaves["global_rot"], leaves["joint_rot"], leaves["trans"], betas, weights, window=int(c["window"]),
                                  frame_offset=lo, total_frames=N, halo_prev=halo(lo - 1) if lo > 0 else None,
                                  halo_next=halo(hi) if hi < N else None, **consts)
        loss.backward()
        total += float(loss)
        gb = gb + betas.grad
        for k in full:
            grads[k].append(leaves[k].grad)
    ref = float(PINS[name + "/ref64_total"])
    assert abs(total - ref) <= TOL * abs(ref), (total, ref)
    for k in full:
        assert rel(torch.cat(grads[k]).cpu().numpy(), PINS["%s/ref64_grad_%s" % (name, k)]) <= TOL, k
    assert rel(gb.cpu().numpy(), PINS[name + "/ref64_grad_betas"]) <= TOL


def test_missing_halo_is_an_error():
    from smilify_b200.fitter_terms import fitter_prior_terms

    g, j, tr = torch.zeros(4, 3, device=DEV), torch.zeros(4, 10, 3, device=DEV), torch.zeros(4, 3, device=DEV)
    with pytest.raises(RuntimeError, match="halo_prev"):
        fitter_prior_terms(g, j, tr, None, dict(w_temp=1.0), frame_offset=2, total_frames=6, halo_next=torch.zeros(36, device=DEV))


def test_fused_adam_matches_torch_adam_eager_and_replayed():
    from smilify_b200.fitter_terms import FusedAdam

    rs = np.random.RandomState(0)
    shapes = [(64, 3), (64, 54, 3), (64, 3), (5,), (1,)]
    init = [torch.from_numpy(rs.standard_normal(s).astype(np.float32)).to(DEV) for s in shapes]
    ours = [p.clone() for p in init]
    theirs = [p.clone().requires_grad_(True) for p in init]
    groups = lambda ps: [{"params": ps[:4], "lr": 5e-3}, {"params": ps[4:], "lr": 1.0}]   # optimize_to_joints.py:117-126
    opt_o, opt_t = FusedAdam(groups(ours), lr=5e-3), torch.optim.Adam(groups(theirs), lr=5e-3)
    grads = [torch.empty_like(p) for p in ours]
    graph = None
    for it in range(7):
        new = [torch.from_numpy(rs.standard_normal(s).astype(np.float32)).to(DEV) for s in shapes]
        for g, n, p in zip(grads, new, theirs):
            g.copy_(n)
            p.grad = n.clone()
        opt_t.step()
        if it < 3:
            opt_o.step(grads)                       # eager
        else:
            if graph is None:                        # capture once, replay for the remaining steps
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    snap = [p.clone() for p in ours], [m.clone() for m in opt_o.exp_avg], [v.clone() for v in opt_o.exp_avg_sq], opt_o.step_count.clone()
                    with torch.cuda.graph(graph, stream=side):
                        opt_o.step(grads)
                    # the capture itself executes nothing: state unchanged
                torch.cuda.current_stream().wait_stream(side)
                assert int(opt_o.step_count) == int(snap[3])
            graph.replay()
        for a, b in zip(ours, theirs):
            assert rel(a.cpu().numpy(), b.detach().cpu().numpy()) < 2e-6, it
    assert int(opt_o.step_count) == 7
