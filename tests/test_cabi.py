"""CPU: the C-ABI library builds for sm_100a, loads without a GPU and exports every symbol that
include/smil_b200.h declares; host-side argument checks fail loudly; no CPU fallback exists."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import smilify_b200
from smilify_b200 import _lib
from tests import rigs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from smilify_b200 import build

    build.build()
    return _lib.lib()


def test_exports_match_header(lib):
    hdr = open(os.path.join(ROOT, "include", "smil_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int|void|size_t|long long|const char\*)\s+(smil_\w+)\s*\(", hdr, re.M))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.smil_abi_version() == _lib.ABI_VERSION


def test_struct_sizes_match_header(lib):
    """ctypes mirrors of the POD argument structs have the C layout (checked against nvcc's view
    through sizes that the library reports indirectly: a NULL-argument call must fail cleanly)."""
    assert C.sizeof(_lib.SmilFwdArgs) % 8 == 0 and C.sizeof(_lib.SmilBwdArgs) > C.sizeof(_lib.SmilFwdArgs)
    assert lib.smil_forward(None, None, None) == -1
    assert b"NULL" in lib.smil_last_error()
    assert lib.smil_project_loss_forward(None, None) == -1


def test_no_cpu_fallback():
    smilify_b200.configure(rigs.REAL_RIGS["stick"])
    from smilify_b200.smal_torch import SMAL

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        SMAL(torch.device("cpu"))
    from smilify_b200.batch_lbs import batch_rodrigues

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        batch_rodrigues(torch.zeros(4, 3))


def test_missing_library_is_loud(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libsmil_b200.so")
    with pytest.raises(RuntimeError, match="missing"):
        _lib.lib()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "smilify_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), fn


def test_packer_index_tables_bit_exact():
    from smilify_b200 import packer
    from smilify_b200.model_io import load_rig, reference_layout

    for name in ("stick", "mouse"):
        dd = load_rig(rigs.REAL_RIGS[name])
        lay = reference_layout(dd, static_joints=bool(dd.get("static_joint_locs", False)))
        ref = np.asarray(dd["kintree_table"])[0].astype(np.int32)
        assert lay["parents"].dtype == np.int32 and np.array_equal(lay["parents"], ref)
        t = packer.tree_tables(lay["parents"])
        assert np.array_equal(t["parents"][1:], ref[1:]) and t["parents"][0] == -1
        # every joint appears once, parents before children
        pos = np.empty(len(ref), dtype=np.int64)
        pos[t["order"]] = np.arange(len(ref))
        assert sorted(t["order"].tolist()) == list(range(len(ref)))
        assert all(pos[ref[i]] < pos[i] for i in range(1, len(ref)))
