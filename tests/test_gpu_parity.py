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


@pytest.mark.parametrize("name", ["stick_base", "stick_full", "stick_beta3", "mouse_full", "stick_b1_nobeta"])
def test_lane_per_sample_forward_variant(name, monkeypatch):
    """The alternative forward skin pass (SMIL_FWD_V3=1, read at model creation) meets the same bar."""
    monkeypatch.setenv("SMIL_FWD_V3", "1")
    try:
        rows = compare_case(rigs.CASE_BY_NAME[name])
    finally:
        monkeypatch.setenv("SMIL_FWD_V3", "0")
        from tests.common import make_cuda_model
        make_cuda_model(rigs.CASE_BY_NAME["stick_base"])   # the switch is process-wide: put it back
    bad = [(n, e, g) for n, e, g in rows if not (e <= TOL) or (g is not None and not (g <= 2 * TOL))]
    assert not bad, "tolerance %g exceeded: %s" % (TOL, bad)
