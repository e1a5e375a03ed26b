"""Memory-safety and race evidence without compute-sanitizer (the tool is closed on this GPU pool, see
profiles/r02_sanitizer.md): every buffer the host side hands to the library -- outputs, saved state, scratch -- is
allocated by a test allocator that (a) poisons the contents with NaN bit patterns, so that a kernel reading scratch it
has not written, or leaving part of an output unwritten, shows up as NaN against the oracle, and (b) surrounds it with
guard bands whose bytes must be untouched afterwards (out-of-bounds writes).  Races in the persistent TMA / mbarrier
rings show up as run-to-run differences: the non-atomic kernel variants must be bit-identical over repeated launches
at batch sizes that wrap the rings with odd tails."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import rigs  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 8192      # bytes either side
POISON = 0xFF     # float32 0xFFFFFFFF = NaN


class GuardAllocator:
    def __init__(self):
        self.live = []

    def __call__(self, *size, **kw):
        if len(size) == 1 and not isinstance(size[0], int):
            size = tuple(size[0])
        dtype = kw.get("dtype", torch.float32)
        dev = kw["device"]
        n = int(np.prod(size)) if len(size) else 1
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        padded = (nbytes + 255) // 256 * 256
        raw = torch.full((GUARD + padded + GUARD,), POISON, dtype=torch.uint8, device=dev)
        self.live.append((raw, nbytes))
        return raw[GUARD:GUARD + nbytes].view(dtype).reshape(tuple(size))

    def check(self):
        torch.cuda.synchronize()
        bad = []
        for i, (raw, nbytes) in enumerate(self.live):
            lo, hi = raw[:GUARD], raw[GUARD + nbytes:]
            if not bool((lo == POISON).all()) or not bool((hi == POISON).all()):
                bad.append((i, nbytes, int((lo != POISON).sum()), int((hi != POISON).sum())))
        n = len(self.live)
        self.live.clear()
        return n, bad


@pytest.fixture
def guard(monkeypatch):
    from smilify_b200 import functions

    g = GuardAllocator()
    monkeypatch.setattr(functions, "_empty", g)
    return g


# batches that make the persistent passes wrap their rings with odd tails (the fused kernels take 32 samples per warp
# pass and 148 CTAs): big enough for several groups per CTA on some CTAs and none on others
BIG = [
    rigs._c("guard_stick_1201", "stick", 1201, dict(with_scale=True, with_trans_off=True)),
    rigs._c("guard_stick_2377", "stick", 2377),
    rigs._c("guard_mouse_301", "mouse", 301, dict(with_trans_off=True)),
    rigs._c("guard_stick13_333", "stick13", 333, dict(with_scale=True)),
]
SMALL = ["stick_base", "stick_full", "stick_rotmat", "stick_delv", "stick_jointsonly", "stick_b1_nobeta", "stick_odd5",
         "mouse_full", "mouse_regress", "mouse_odd7", "syn_k3_delv_prop", "syn_pose_full", "syn_wide", "syn_chain",
         "stick13_full", "mouse40_base"]


def _check_against_oracle(case, force_generic):
    from tests.common import TOL, rel, run_cuda_case, run_oracle_case

    c_out, c_grad = run_cuda_case(case, DEV, force_generic)
    o_out, o_grad = run_oracle_case(case)
    assert set(c_grad) == set(o_grad)
    bad = [(k, rel(c_out[k], v)) for k, v in o_out.items() if not rel(c_out[k], v) <= TOL]
    bad += [(k, rel(c_grad[k], v)) for k, v in o_grad.items() if not rel(c_grad[k], v) <= TOL]
    assert not bad, bad   # a NaN (poisoned scratch read, unwritten output) fails the comparison too


@pytest.mark.parametrize("force_generic", [False, True], ids=["fused", "generic"])
@pytest.mark.parametrize("name", SMALL)
def test_guard_bands_small_cases(name, force_generic, guard):
    _check_against_oracle(rigs.CASE_BY_NAME[name], force_generic)
    n, bad = guard.check()
    assert n >= 4 and not bad, "buffers with touched guard bands (index, bytes, low, high): %s" % bad


@pytest.mark.parametrize("case", BIG, ids=[c["name"] for c in BIG])
def test_guard_bands_ring_wrapping_batches(case, guard):
    _check_against_oracle(case, False)
    n, bad = guard.check()
    assert n >= 4 and not bad, bad


@pytest.mark.parametrize("variant", ["item_major_bwd", "vm_slots_summed", "vm_slots_gathered", "pose_general_bwd",
                                     "pose_samples_per_warp"])
def test_guard_bands_kernel_variants(variant, guard, monkeypatch):
    from tests.test_gpu_parity import VARIANTS

    for k, v in VARIANTS[variant].items():
        monkeypatch.setenv(k, v)
    _check_against_oracle(BIG[0], False)
    n, bad = guard.check()
    assert n >= 4 and not bad, bad


def test_guard_bands_losses_and_triangulation(guard):
    from smilify_b200 import losses, triangulation

    sc = rigs.loss_scene(5, 4, 17, 3, False)
    f = lambda k, grad=False: torch.from_numpy(np.ascontiguousarray(sc[k])).to(DEV).requires_grad_(grad)
    j, R, T, fov = f("joints", True), f("R", True), f("T", True), f("fov", True)
    loss, _ = losses.multiview_keypoint_loss(j, R, T, fov, f("aspect"), f("target"), f("vis"), f("view_mask"),
                                             f("joint_weights"), 512.0)
    loss.backward()
    tri, valid = triangulation.triangulate_joints_dlt(f("target"), f("vis"), R, T, fov, f("aspect"), f("view_mask"), 512)
    (tri * valid[..., None]).sum().backward()
    d6 = torch.randn(67, 6, device=DEV, requires_grad=True)
    triangulation.rotation_6d_to_matrix(d6).sum().backward()
    for t in (j, R, T, fov, d6):
        assert bool(torch.isfinite(t.grad).all())
    n, bad = guard.check()
    assert n >= 8 and not bad, bad


@pytest.mark.parametrize("rig,B", [("stick", 1201), ("stick", 4096), ("mouse", 301)])
@pytest.mark.parametrize("variant", ["vm_slots_summed", "item_major_bwd"])
def test_repeated_launches_are_bit_identical(rig, B, variant, monkeypatch):
    """Race evidence: with the atomic accumulation switched off every sum has a fixed order, so any run-to-run
    difference over 12 launches (ring slots reused before their consumer finished, a missing fence between the TMA
    write and the reader) would be a data race."""
    import smilify_b200
    from smilify_b200.smal_torch import SMAL
    from tests.common import rig_file
    from tests.test_gpu_parity import VARIANTS

    for k, v in VARIANTS[variant].items():
        monkeypatch.setenv(k, v)
    dd = rigs.get_rig(rig)
    smilify_b200.configure(rig_file(rig))
    model = SMAL(torch.device(DEV))
    inp = rigs.make_inputs(dd, B, seed=3, with_scale=True, with_trans_off=True)
    t = {k: torch.from_numpy(v).to(DEV).requires_grad_(True) for k, v in inp.items()}
    V, J = model._tables["V"], model._tables["J"]
    gv, gj = torch.randn(B, V, 3, device=DEV), torch.randn(B, J, 3, device=DEV)
    first = None
    for it in range(12):
        verts, joints, Rs, vs = model(t["beta"], t["theta"], trans=t["trans"], betas_logscale=t["betas_logscale"],
                                      betas_trans=t["betas_trans"])
        grads = torch.autograd.grad([verts, joints], list(t.values()), [gv, gj])
        cur = [verts, joints] + list(grads)
        if first is None:
            first = [c.clone() for c in cur]
        else:
            for a, b in zip(first, cur):
                assert torch.equal(a, b), "launch %d differs from launch 0" % it


def test_atomic_accumulation_spread_is_rounding_only():
    """The default backward adds per-CTA partial sums with red.global.add: the order varies, so repeated launches may
    differ -- by rounding only."""
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    dd = rigs.get_rig("stick")
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    model = SMAL(torch.device(DEV))
    B = 1201
    inp = rigs.make_inputs(dd, B, seed=3)
    t = {k: torch.from_numpy(inp[k]).to(DEV).requires_grad_(True) for k in ("beta", "theta", "trans")}
    gv, gj = torch.randn(B, 3020, 3, device=DEV), torch.randn(B, 55, 3, device=DEV)
    runs = []
    for it in range(8):
        verts, joints, Rs, vs = model(t["beta"], t["theta"], trans=t["trans"])
        runs.append(torch.autograd.grad([verts, joints], list(t.values()), [gv, gj]))
    for r in runs[1:]:
        for a, b in zip(runs[0], r):
            assert float((a - b).abs().max() / a.abs().max()) < 2e-6


def test_guard_bands_fitter_terms_and_adam(guard):
    """The fitter objective kernel (dense pose-prior path, shard with both halos) and the fused Adam: no write outside
    their buffers, no unwritten output (a NaN from the poisoned allocations would show in the results)."""
    from smilify_b200.fitter_terms import FusedAdam, fitter_prior_terms

    rs = np.random.RandomState(3)
    n, J, nB = 37, 35, 6
    f = lambda *s: torch.from_numpy(rs.standard_normal(s).astype(np.float32)).to(DEV)
    g, j, tr, be = f(n, 3).requires_grad_(True), f(n, J - 1, 3).requires_grad_(True), f(n, 3).requires_grad_(True), f(nB).requires_grad_(True)
    P = 3 * J
    loss = fitter_prior_terms(g, j, tr, be, dict(w_limit=1.0, w_pose=0.5, w_splay=0.1, w_betas=0.3, w_temp=0.2), window=10,
                              max_limits=f(J - 1, 3).abs(), min_limits=-f(J - 1, 3).abs(), betas_mean=f(nB), betas_prec=f(nB, nB),
                              pose_prior=dict(mean=f(P), precs=f(P, P) * 0.1, use=(f(P) > 0).float()),
                              frame_offset=20, total_frames=100, halo_prev=f(P + 3), halo_next=f(P + 3))
    loss.backward()
    for t in (g, j, tr, be):
        assert bool(torch.isfinite(t.grad).all())
    assert bool(torch.isfinite(loss))
    nbuf, bad = guard.check()
    assert nbuf >= 5 and not bad, bad
    # Adam updates caller tensors in place: surround them with guard bands by hand
    raw = torch.full((3, 4096 + 1001 + 4096), float("nan"), device=DEV)
    p, m_, v_ = (raw[i, 4096:4096 + 1001] for i in range(3))
    p.copy_(f(1001)); m_.zero_(); v_.zero_()
    opt = FusedAdam([p.detach()], lr=1e-2)
    opt.exp_avg[0], opt.exp_avg_sq[0] = m_, v_
    opt.params[0] = p
    opt.step([f(1001)])
    torch.cuda.synchronize()
    assert bool(torch.isfinite(raw[:, 4096:4096 + 1001]).all())
    assert bool(torch.isnan(raw[:, :4096]).all()) and bool(torch.isnan(raw[:, 4096 + 1001:]).all())
