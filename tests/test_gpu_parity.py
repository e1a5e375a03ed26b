"""GPU parity (the first gate): the CUDA path, called through the reference-shaped ``SMAL``
class and the C ABI, against the fp64 oracle and the frozen outputs of the real reference, on
every case of tests/rigs.CASES -- fused kernels and unfused kernels both.

Bar (north_star): 1e-5 relative (max-norm) on fp32 vertices, joints and gradients."""
import numpy as np
import pytest
import torch

from tests import rigs
from tests.common import TOL, compare_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("force_generic", [False, True], ids=["fused", "generic"])
@pytest.mark.parametrize("case", rigs.CASES, ids=[c["name"] for c in rigs.CASES])
def test_case(case, force_generic):
    rows = compare_case(case, force_generic=force_generic)
    bad = [(n, e, g) for n, e, g in rows if not (e <= TOL) or (g is not None and not (g <= 2 * TOL))]
    assert not bad, "tolerance %g exceeded: %s" % (TOL, bad)


def test_parent_indices_bit_exact():
    from tests.common import make_cuda_model

    for name in ("stick_base", "mouse_base"):
        model, dd = make_cuda_model(rigs.CASE_BY_NAME[name])
        ref = np.asarray(dd["kintree_table"])[0].astype(np.int32)
        assert model.parents.dtype == np.int32 and np.array_equal(model.parents, ref)
        assert np.array_equal(model._tables["parents"][1:], ref[1:])
        assert model.faces.dtype == torch.int64


VARIANTS = {
    # per-model tuning switches (SmilTuning, read from SMIL_* at model creation): every kernel variant kept in the
    # library meets the same bar as the default path
    "item_major_bwd": {"SMIL_BWD_VM": "0"},
    "vm_slots_summed": {"SMIL_BWD_ATOMIC": "0"},
    "vm_slots_gathered": {"SMIL_BWD_ATOMIC": "0", "SMIL_BWD_SLOT_REDUCE": "0"},
    "pose_samples_per_warp": {"SMIL_POSE_V2": "1"},
    "pose_general_bwd": {"SMIL_POSE_LEAN": "0"},
    "item_major_double_buffered": {"SMIL_BWD_VM": "0", "SMIL_BWD_STAGES": "2"},
    "fwd_topsync_on": {"SMIL_FWD_TOPSYNC": "1"},
    "fwd_topsync_off": {"SMIL_FWD_TOPSYNC": "0"},
}


@pytest.mark.parametrize("variant", list(VARIANTS))
@pytest.mark.parametrize("name", ["stick_base", "stick_full", "stick_beta3", "mouse_full", "stick_b1_nobeta", "syn_wide",
                                  "syn_chain"])
def test_kernel_variants(name, variant, monkeypatch):
    for k, v in VARIANTS[variant].items():
        monkeypatch.setenv(k, v)   # per MODEL (the drop-in builds a fresh model per case), nothing process-wide
    rows = compare_case(rigs.CASE_BY_NAME[name])
    bad = [(n, e, g) for n, e, g in rows if not (e <= TOL) or (g is not None and not (g <= 2 * TOL))]
    assert not bad, "tolerance %g exceeded: %s" % (TOL, bad)
