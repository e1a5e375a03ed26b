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


def test_dropin_reexports_and_real_package_path():
    """After install(): the names other reference modules import from ``smal_model.smal_torch`` exist
    (smal_fitter/priors/pose_prior_35.py:8 imports CustomUnpickler), and the REST of the real ``smal_model`` package
    stays importable when the reference is on sys.path (``smal_model.smal_basics``, needed for ignore_sym=False)."""
    import io
    import os
    import pickle
    import sys

    import numpy as np

    import smilify_b200

    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] == "smal_model"}
    saved_path = list(sys.path)
    for k in saved:
        del sys.modules[k]
    try:
        ref = "/root/reference"
        have_ref = os.path.isdir(os.path.join(ref, "smal_model"))
        if have_ref:
            sys.path.insert(0, ref)
        smilify_b200.install(force=True)
        from smal_model.smal_torch import CustomUnpickler, load_smal_model, undo_chumpy  # noqa: F401

        buf = io.BytesIO()
        pickle.dump({"a": np.arange(3)}, buf, protocol=2)
        buf.seek(0)
        assert np.array_equal(CustomUnpickler(buf).load()["a"], np.arange(3))
        w = CustomUnpickler.ChumpyWrapper(np.arange(4.0))
        assert np.array_equal(undo_chumpy(w), np.arange(4.0)) and np.array_equal(np.asarray(w), w.r)
        assert undo_chumpy(3) == 3
        if have_ref:
            from oracle.pin_against_reference import install_config_shim

            saved_cfg = sys.modules.get("config")
            install_config_shim(ref, os.path.join(ref, "3D_model_prep", "SMILy_STICK.pkl"))   # its default rig is a missing blob
            try:
                import smal_model.smal_basics as sb   # not replaced: must resolve to the reference's file
            finally:
                if saved_cfg is None:
                    sys.modules.pop("config", None)
                else:
                    sys.modules["config"] = saved_cfg
            assert sb.__file__.startswith(ref)
    finally:
        sys.path[:] = saved_path
        for k in [k for k in sys.modules if k.split(".")[0] == "smal_model"]:
            del sys.modules[k]
        sys.modules.update(saved)


def test_rig_schema_validation_and_pack_cache(tmp_path):
    """(f)3: the loader rejects rigs whose arrays disagree (instead of failing somewhere inside a kernel), and the packed
    tables survive the flat on-disk cache bit for bit."""
    import numpy as np
    import pytest

    from smilify_b200 import model_io, packer
    from tests import rigs

    dd = rigs.synthetic_rig(**rigs.SYNTH_SPECS["syn_k3"])
    model_io.validate_rig(dd)
    for key, mutate, msg in (
        ("weights", lambda a: a[:, :-1], "weights has shape"),
        ("kintree_table", lambda a: np.stack([np.r_[a[0][:3], [7], a[0][4:]], a[1]]), "must precede"),
        ("shapedirs", lambda a: a[:-1], "shapedirs has shape"),
        ("weights", lambda a: a * 0.5, "sum to 1"),
        ("v_template", lambda a: np.where(np.arange(a.size).reshape(a.shape) == 5, np.nan, a), "non-finite"),
    ):
        broken = dict(dd)
        broken[key] = mutate(np.asarray(dd[key]).copy())
        with pytest.raises(ValueError, match=msg):
            model_io.validate_rig(broken, "broken.pkl")
    lay = model_io.reference_layout(dd, static_joints=False)
    args = (lay["v_template"], lay["shapedirs"], lay["posedirs"], lay["J_regressor"], lay["weights"], lay["parents"], lay["J"])
    first = packer.pack_model_cached(*args, cache_dir=str(tmp_path))
    again = packer.pack_model_cached(*args, cache_dir=str(tmp_path))     # from the file
    direct = packer.pack_model(*args)
    assert len(list(tmp_path.iterdir())) == 1
    for k, v in direct.items():
        for other in (first, again):
            a, b = np.ascontiguousarray(np.asarray(v)), np.ascontiguousarray(np.asarray(other[k]))
            assert a.shape == b.shape and a.dtype == b.dtype and a.tobytes() == b.tobytes(), k


def test_flat_rig_file_roundtrip_and_embedded_pack(tmp_path, monkeypatch):
    """a rig exported as one pickle-free .npz loads to the same arrays, carries its packed
    tables, and a stale embedded pack (another packer version) is ignored."""
    import numpy as np

    from smilify_b200 import model_io, packer
    from tests import rigs

    src = rigs.REAL_RIGS["stick"]
    dst = model_io.export_flat(src, str(tmp_path / "stick_flat.npz"))
    a, b = model_io.load_rig(src), model_io.load_rig(dst)
    for k in ("v_template", "shapedirs", "posedirs", "J_regressor", "weights", "kintree_table", "f"):
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k
    assert [str(x) for x in a["J_names"]] == list(b["J_names"])
    with np.load(dst, allow_pickle=False) as z:      # nothing in the file needs pickle
        assert "pack_key" in z.files and any(k.startswith("pack/") for k in z.files)
    lay = model_io.reference_layout(b, static_joints=bool(b.get("static_joint_locs", False)))
    args = (lay["v_template"], lay["shapedirs"], lay["posedirs"], lay["J_regressor"], lay["weights"], lay["parents"], lay["J"])
    monkeypatch.setattr(packer, "pack_model", lambda *a_, **k_: (_ for _ in ()).throw(AssertionError("packed again")))
    t = packer.pack_model_cached(*args, cache_dir="", source=dst)      # served from the file: the packer is not called
    assert t["V"] == 3020 and t["vm_ok"]
    monkeypatch.setattr(packer, "PACK_FORMAT", packer.PACK_FORMAT + 1)
    assert model_io.embedded_pack(dst, packer.pack_key(*args)) is None  # other packer version: the embedded pack is stale
