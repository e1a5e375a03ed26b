"""Make the reference pick up the CUDA path without editing it.

``fitter.py``, ``fitter_3d/trainer.py`` and the neuralSMIL regressors do
``from smal_model.smal_torch import SMAL`` at import time
(/root/reference/smal_fitter/fitter.py:15, fitter_3d/trainer.py:21).  ``install()`` registers
this package's modules under those names in ``sys.modules`` *before* they are imported, so the
reference's own call sites resolve to the B200 kernels."""
from __future__ import annotations

import sys
import types


def install(force: bool = False):
    from . import batch_lbs, smal_torch

    if not force and "smal_model.smal_torch" in sys.modules and \
            getattr(sys.modules["smal_model.smal_torch"], "__smilify_b200__", False) is False and \
            any(m in sys.modules for m in ("smal_fitter.fitter", "fitter_3d.trainer")):
        raise RuntimeError("install() must run before smal_fitter / fitter_3d are imported (or pass force=True)")
    pkg = sys.modules.get("smal_model")
    if pkg is None:
        # keep the REAL package importable for everything that is not replaced (smal_model.smal_basics, ...): a bare
        # namespace module only when the reference is not on sys.path at all
        import importlib.util
        import os

        pkg = types.ModuleType("smal_model")
        pkg.__path__ = []
        for entry in sys.path:
            cand = os.path.join(entry or ".", "smal_model")
            if os.path.isdir(cand) and os.path.exists(os.path.join(cand, "smal_torch.py")):
                pkg.__path__.append(cand)
                init = os.path.join(cand, "__init__.py")
                if os.path.exists(init):
                    pkg.__file__ = init
                break
        sys.modules["smal_model"] = pkg
    smal_torch.__smilify_b200__ = True
    batch_lbs.__smilify_b200__ = True
    sys.modules["smal_model.smal_torch"] = smal_torch
    sys.modules["smal_model.batch_lbs"] = batch_lbs
    pkg.smal_torch = smal_torch
    pkg.batch_lbs = batch_lbs
    return smal_torch.SMAL
