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


# ---- rows 9-12: oracle vs outputs of the reference's OWN loss / PCA code (tests/golden/loss_pins.npz, frozen by
# oracle/pin_against_reference.py::pin_losses from AST-extracted reference functions) --------------------------------
def _pins(prefix):
    z = np.load(os.path.join(rigs.GOLDEN, "loss_pins.npz"))
    names = sorted({k.split("/")[1] for k in z.files if k.startswith(prefix + "/")})
    return z, names


def _opt(z, key, dtype=torch.float64):
    if key not in z.files:
        return None
    a = z[key]
    return torch.from_numpy(a) if a.dtype == bool else torch.from_numpy(a).to(dtype)


def test_multiview_loss_oracle_matches_reference_code():
    z, names = _pins("multiview")
    assert len(names) >= 12
    for n in names:
        p = "multiview/%s/" % n
        pred = _opt(z, p + "pred").requires_grad_(True)
        vm = _opt(z, p + "view_mask")
        loss = PO.multiview_keypoint_loss(pred, _opt(z, p + "target"), _opt(z, p + "vis"),
                                          None if vm is None else vm.double(), _opt(z, p + "joint_weights"))
        assert abs(float(loss) - float(z[p + "loss"])) <= 1e-12 * abs(float(z[p + "loss"])), n
        if loss.grad_fn is not None:
            loss.backward()
            assert rel(pred.grad.numpy(), z[p + "grad_pred"]) < 1e-12, n
        else:
            assert np.abs(z[p + "grad_pred"]).max() == 0.0, n


def test_singleview_and_3d_loss_oracle_matches_reference_code():
    z, names = _pins("singleview")
    assert len(names) >= 6
    for n in names:
        p = "singleview/%s/" % n
        sm = _opt(z, p + "sample_mask")
        jw = _opt(z, p + "joint_weights")
        pred = _opt(z, p + "pred").requires_grad_(True)
        loss = PO.singleview_keypoint_loss(pred, _opt(z, p + "target"), _opt(z, p + "vis"), jw, sm)
        loss.backward()
        assert abs(float(loss) - float(z[p + "loss"])) <= 1e-12 * abs(float(z[p + "loss"])), n
        assert rel(pred.grad.numpy(), z[p + "grad_pred"]) < 1e-12, n
        j3 = _opt(z, p + "j3").requires_grad_(True)
        loss3 = PO.keypoint3d_loss(j3, _opt(z, p + "t3"), _opt(z, p + "vis"), jw, sm)
        loss3.backward()
        if np.isnan(float(z[p + "loss3"])):   # a non-finite predicted joint: inf * 0 in the reference -> NaN loss
            assert np.isnan(float(loss3)), n
            assert np.array_equal(np.isnan(j3.grad.numpy()), np.isnan(z[p + "grad_j3"])), n
            ok = ~np.isnan(z[p + "grad_j3"])
            assert rel(j3.grad.numpy()[ok], z[p + "grad_j3"][ok]) < 1e-12, n
        else:
            assert abs(float(loss3) - float(z[p + "loss3"])) <= 1e-12 * abs(float(z[p + "loss3"])), n
            assert rel(j3.grad.numpy(), z[p + "grad_j3"]) < 1e-12, n


def test_fitter_objective_oracle_matches_reference_code():
    z, names = _pins("fitter")
    for n in names:
        p = "fitter/%s/" % n
        pred = _opt(z, p + "pred").requires_grad_(True)
        vis = _opt(z, p + "vis")
        # fitter.py:292-296 on already-projected points == projection_oracle.fitter_joint_loss after its projection
        v = vis.bool()
        pm = torch.where(v[..., None], pred, torch.full_like(pred, -1.0))
        tm = torch.where(v[..., None], _opt(z, p + "target"), torch.full_like(pred, -1.0))
        loss = float(z[p + "w_j2d"]) * torch.mean((pm - tm) ** 2)
        loss.backward()
        assert abs(float(loss) - float(z[p + "loss"])) <= 1e-12 * max(abs(float(z[p + "loss"])), 1e-30), n
        assert rel(pred.grad.numpy(), z[p + "grad_pred"]) < 1e-12 or np.abs(z[p + "grad_pred"]).max() == 0.0, n


def test_pca_joint_values_oracle_matches_reference_code():
    from oracle import smal_oracle as SO

    z, names = _pins("pca")
    assert len(names) == 3
    for n in names:
        p = "pca/%s/" % n
        sw = _opt(z, p + "scale_w").requires_grad_(True)
        tw = _opt(z, p + "trans_w")
        tw = tw.requires_grad_(True) if tw.numel() else None
        ls, bt = SO.pca_weights_to_joint_values(_opt(z, p + "scaledirs"), _opt(z, p + "transdirs"), sw, tw)
        ((ls * _opt(z, p + "g_ls")).sum() + ((bt * _opt(z, p + "g_bt")).sum() if tw is not None else 0.0)).backward()
        assert rel(ls.detach().numpy(), z[p + "log_scales"]) < 1e-12, n
        assert rel(bt.detach().numpy(), z[p + "betas_trans"]) < 1e-12, n
        assert rel(sw.grad.numpy(), z[p + "grad_scale_w"]) < 1e-12, n
        if tw is not None:
            assert rel(tw.grad.numpy(), z[p + "grad_trans_w"]) < 1e-12, n


def test_loss_chain_oracle_matches_frozen_reference_values():
    """Projection oracle -> the reference's multi-view loss code, on the scenes the GPU tests use."""
    z = np.load(os.path.join(rigs.GOLDEN, "loss_chain_pins.npz"))
    for name, (B, V, J, seed, shared) in rigs.LOSS_SCENES.items():
        sc = rigs.loss_scene(B, V, J, seed, shared)
        t = {k: torch.from_numpy(np.ascontiguousarray(v)).double() for k, v in sc.items() if k != "image_size"}
        leaf = {k: t[k].clone().requires_grad_(True) for k in ("joints", "R", "T", "fov")}
        pred = PO.multiview_project(leaf["joints"], leaf["R"].expand(B, V, 3, 3), leaf["T"].expand(B, V, 3),
                                    leaf["fov"].expand(B, V), t["aspect"].expand(B, V), float(sc["image_size"]))
        loss = PO.multiview_keypoint_loss(pred, t["target"], t["vis"], t["view_mask"], t["joint_weights"])
        loss.backward()
        assert rel(pred.detach().numpy(), z[name + "/pred"]) < 1e-12
        assert abs(float(loss) - float(z[name + "/loss"])) <= 1e-12 * abs(float(z[name + "/loss"]))
        for k, v in leaf.items():
            assert rel(v.grad.numpy(), z[name + "/grad_" + k]) < 1e-11, (name, k)


def test_fitter_prior_oracle_matches_the_frozen_reference_code():
    """(f)2: oracle/fitter_oracle.py against the outputs of the reference's own fitter.py statements (fp64, exact)."""
    import torch

    from oracle import fitter_oracle as FO

    z = np.load(os.path.join(rigs.GOLDEN, "fitter_prior_pins.npz"))
    for name in sorted({k.split("/")[0] for k in z.files}):
        c = {k.split("/", 1)[1][3:]: z[k] for k in z.files if k.startswith(name + "/in_")}
        t = lambda k: torch.from_numpy(c[k]).double()
        leaves = {k: t(k).clone().requires_grad_(True) for k in ("global_rot", "joint_rot", "trans", "betas")}
        wl, wp, ws, wb, wt = (float(x) for x in c["weights"])
        total = FO.epoch_objective(leaves["global_rot"], leaves["joint_rot"], leaves["trans"], leaves["betas"],
                                   dict(w_limit=wl, w_pose=wp, w_splay=ws, w_betas=wb, w_temp=wt), int(c["window"]),
                                   max_limits=t("max_limits"), min_limits=t("min_limits"),
                                   pose_prior=dict(mean=t("pose_mean"), precs=t("pose_precs"), use=t("pose_use")),
                                   betas_mean=t("betas_mean"), betas_prec=t("betas_prec"),
                                   rotation_mask=t("rotation_mask"), global_mask=t("global_mask"))
        total.backward()
        assert abs(float(total) - float(z[name + "/ref64_total"])) <= 1e-12 * abs(float(total))
        for k, v in leaves.items():
            assert np.abs(v.grad.numpy() - z["%s/ref64_grad_%s" % (name, k)]).max() <= 1e-12
