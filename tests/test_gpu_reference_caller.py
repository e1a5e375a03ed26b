"""an UNMODIFIED reference caller driven through the shadowed module.

``/root/reference/fitter_3d/trainer.py`` does ``from smal_model.smal_torch import SMAL`` at import time (:21) and
``SMAL3DFitter.forward`` (:169-245) calls it with per-mesh betas, log-scales and joint translation offsets.  The test
imports that file twice -- once over the reference's own ``SMAL`` on the CPU, once after ``smilify_b200.install()`` on
the GPU -- with the same parameter values, and compares outputs and parameter gradients.  Only what the file needs at
IMPORT time and cannot be had here is stubbed: the ``pytorch3d`` / ``matplotlib`` names and ``nibabel.eulerangles``
(one function, restated).  Skipped where ``/root/reference`` does not exist (the GPU box of the round-end run)."""
import importlib
import os
import sys
import types

import numpy as np
import pytest
import torch

from tests import rigs
from tests.common import rel

pytestmark = pytest.mark.gpu
REF = "/root/reference"
DEV = "cuda:0"


class _Anything(types.ModuleType):
    """A module whose every attribute exists (callable no-op / sub-module): enough for ``from x import y`` lines."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        sub = _Anything(self.__name__ + "." + name)
        sub.__call__ = lambda *a, **k: None
        setattr(self, name, sub)
        return sub

    def __call__(self, *a, **k):
        return None


def _euler2angle_axis(z=0, y=0, x=0):
    """nibabel.eulerangles.euler2angle_axis (static frame, z then y then x) through scipy."""
    from scipy.spatial.transform import Rotation

    rv = Rotation.from_euler("xyz", [x, y, z]).as_rotvec()   # extrinsic x, y, z = R_z R_y R_x
    ang = float(np.linalg.norm(rv))
    return (ang, rv / ang) if ang > 0 else (0.0, np.array([1.0, 0.0, 0.0]))


STUBS = ("pytorch3d", "pytorch3d.ops", "pytorch3d.ops.knn", "pytorch3d.loss", "pytorch3d.structures", "pytorch3d.io",
         "pytorch3d.renderer", "pytorch3d.transforms", "matplotlib", "matplotlib.pyplot", "matplotlib.tri", "nibabel",
         "nibabel.eulerangles")
PURGE = ("smal_model", "fitter_3d", "smal_fitter", "config")


def caller_inputs(B=3, seed=17):
    """Seeded parameter values / upstream gradients shared by the test and oracle/pin_against_reference.py::pin_caller."""
    rs = np.random.RandomState(seed)
    dd = rigs.get_rig("stick")
    V, J, nB = np.asarray(dd["v_template"]).shape[0], np.asarray(dd["kintree_table"]).shape[1], np.asarray(dd["shapedirs"]).shape[-1]
    f = lambda a: np.asarray(a, dtype=np.float32)
    vals = dict(betas=f(rs.normal(0, 0.5, (B, nB))), global_rot=f(rs.normal(0, 0.3, (B, 3))), joint_rot=f(rs.normal(0, 0.3, (B, J - 1, 3))),
                trans=f(rs.normal(0, 1, (B, 3))), log_beta_scales=f(rs.normal(0, 0.05, (B, J, 3))),
                betas_trans=f(rs.normal(0, 0.01, (B, J, 3))), deform_verts=f(rs.normal(0, 1e-3, (B, V, 3))))
    return vals, f(rs.standard_normal((B, V, 3))), f(rs.standard_normal((B, J, 3)))


def _purge():
    for k in [k for k in sys.modules if k.split(".")[0] in PURGE]:
        del sys.modules[k]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "fitter_3d")), reason="needs the reference tree")
def test_smal3dfitter_forward_through_the_shadowed_module():
    import multiprocessing

    import smilify_b200
    from oracle.pin_against_reference import install_config_shim

    saved_modules = dict(sys.modules)
    saved_path = list(sys.path)
    saved_cwd = os.getcwd()
    start_method = multiprocessing.get_start_method(allow_none=True)
    try:
        for name in STUBS:
            sys.modules[name] = _Anything(name)
        sys.modules["nibabel.eulerangles"].euler2angle_axis = _euler2angle_axis
        sys.modules["nibabel"].eulerangles = sys.modules["nibabel.eulerangles"]
        sys.path.insert(0, REF)
        os.chdir(REF)
        vals, gv, gj = caller_inputs()
        B = gv.shape[0]

        def run(device, install):
            _purge()
            install_config_shim(REF, os.path.join(REF, "3D_model_prep", "SMILy_STICK.pkl"))
            if install:
                cls = smilify_b200.install(force=True)
            trainer = importlib.import_module("fitter_3d.trainer")     # the unmodified reference file
            if install:
                assert trainer.SMAL is cls and trainer.SMAL.__module__ == "smilify_b200.smal_torch"
            else:
                assert trainer.SMAL.__module__ == "smal_model.smal_torch" and "reference" in trainer.SMAL.__init__.__code__.co_filename
            fit = trainer.SMAL3DFitter(batch_size=B, device=device)
            with torch.no_grad():
                for k, v in vals.items():
                    getattr(fit, k).copy_(torch.from_numpy(v).float())
            verts, joints = fit.forward(return_joints=True)
            ((verts * torch.from_numpy(gv).to(device)).sum() + (joints * torch.from_numpy(gj).to(device)).sum()).backward()
            out = dict(verts=verts, joints=joints)
            out.update({"g_" + k: getattr(fit, k).grad for k in vals})
            return {k: v.detach().cpu().numpy() for k, v in out.items()}

        ref = run("cpu", install=False)
        ours = run(DEV, install=True)
        for k, v in ref.items():
            assert rel(ours[k], v) < 1e-5, (k, rel(ours[k], v))
    finally:
        os.chdir(saved_cwd)
        sys.path[:] = saved_path
        for k in list(sys.modules):
            if k not in saved_modules:
                del sys.modules[k]
        sys.modules.update(saved_modules)
        if start_method is not None:
            multiprocessing.set_start_method(start_method, force=True)


def test_dropin_matches_frozen_smal3dfitter_outputs():
    """The GPU box has no reference tree: the same parameter values through the drop-in ``SMAL`` with the call sequence
    of ``SMAL3DFitter.forward`` (fitter_3d/trainer.py:224-240), against what the UNMODIFIED caller produced over the
    reference's own ``SMAL`` (tests/golden/caller_smal3dfitter.npz, frozen by pin_caller)."""
    import smilify_b200
    from smilify_b200.smal_torch import SMAL

    z = np.load(os.path.join(rigs.GOLDEN, "caller_smal3dfitter.npz"))
    vals, gv, gj = caller_inputs()
    cfg = smilify_b200.configure(rigs.REAL_RIGS["stick"])
    cfg.ALLOW_LIMB_SCALING = True
    cfg.ignore_hardcoded_body = True
    model = SMAL(torch.device(DEV))
    p = {k: torch.from_numpy(v).to(DEV).requires_grad_(True) for k, v in vals.items()}
    verts, joints, Rs, v_shaped = model(p["betas"], torch.cat([p["global_rot"].unsqueeze(1), p["joint_rot"]], dim=1),
                                        betas_logscale=p["log_beta_scales"], betas_trans=p["betas_trans"])
    verts = verts + p["trans"].unsqueeze(1)
    joints = joints + p["trans"].unsqueeze(1)
    verts = verts + p["deform_verts"]
    ((verts * torch.from_numpy(gv).to(DEV)).sum() + (joints * torch.from_numpy(gj).to(DEV)).sum()).backward()
    assert rel(verts.detach().cpu().numpy(), z["out_verts"]) < 1e-5
    assert rel(joints.detach().cpu().numpy(), z["out_joints"]) < 1e-5
    for k, v in p.items():
        assert rel(v.grad.cpu().numpy(), z["grad_" + k]) < 1e-5, (k, rel(v.grad.cpu().numpy(), z["grad_" + k]))
