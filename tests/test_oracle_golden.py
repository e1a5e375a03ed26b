"""CPU: the oracle (oracle/smal_oracle.py, fp32 and fp64) against the frozen outputs and gradients
of the REAL reference (tests/golden/case_*.npz, written by oracle/pin_against_reference.py from the
unmodified /root/reference/smal_model/smal_torch.py), and the projection oracle against the
reference's own numpy OpenCV projection (tests/golden/projection_opencv.npz)."""
import os

import numpy as np
import pytest
import torch

from oracle import projection_oracle as PO
from tests import rigs
from tests.common import rel, run_oracle_case


@pytest.mark.parametrize("case", rigs.CASES, ids=[c["name"] for c in rigs.CASES])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64], ids=["o32", "o64"])
def test_oracle_matches_frozen_reference(case, dtype):
    gold = np.load(os.path.join(rigs.GOLDEN, "case_%s.npz" % case["name"]))
    outs, grads = run_oracle_case(case, dtype=dtype)
    s = case["stride"]
    checked = 0
    for k, v in outs.items():
        if "out_" + k in gold.files:
            cv = v[:, ::s] if k in ("verts", "v_shaped") else v
            assert rel(cv, gold["out_" + k]) < 2e-6, k
            checked += 1
    gold_grads = {f[5:] for f in gold.files if f.startswith("grad_")}
    assert gold_grads == set(grads), (sorted(gold_grads), sorted(grads))
    for k, v in grads.items():
        cv = v[:, ::s] if k == "del_v" else v
        assert rel(cv, gold["grad_" + k]) < 2e-5, k
        checked += 1
    assert checked >= 2


def test_projection_oracle_matches_reference_numpy():
    z = np.load(os.path.join(rigs.GOLDEN, "projection_opencv.npz"))
    S = 512.0
    n = z["X"].shape[0]
    scr = PO.fov_project_screen(torch.from_numpy(z["X"]), torch.from_numpy(z["R_p3d"]), torch.from_numpy(z["T_p3d"]),
                                torch.from_numpy(z["fov"]), torch.from_numpy(z["aspect"]), S).numpy()
    mine = np.stack([scr[..., 0] / S * z["wh"][:, None, 0], scr[..., 1] / S * z["wh"][:, None, 1]], axis=-1)
    ok = np.isfinite(z["px"]).all(axis=-1)
    assert ok.sum() > n * 8
    assert np.abs(mine[ok] - z["px"][ok]).max() < 5e-3  # R/T were rounded to fp32 by the reference helper


def test_projection_default_camera():
    """look_at_view_transform(2.7, 0, 0), fov 60: the origin lands on the image centre and +x
    (world) moves the point to smaller screen x (PyTorch3D's screen x grows to the left of +x view)."""
    R, T = PO.look_at_default()
    p = PO.project_yx(torch.tensor([[[0.0, 0.0, 0.0], [0.1, 0.0, 0.0], [0.0, 0.1, 0.0]]]), R, T,
                      torch.tensor([60.0]), torch.ones(1), 512.0)[0]
    assert torch.allclose(p[0], torch.tensor([256.0, 256.0]), atol=1e-4)
    assert p[1, 1] > 256.0 and abs(p[1, 0] - 256.0) < 1e-4   # R = diag(-1,1,-1) flips x back
    assert p[2, 0] < 256.0 and abs(p[2, 1] - 256.0) < 1e-4
