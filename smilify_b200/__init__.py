"""smilify_b200 -- B200-native (sm_100a) SMAL/SMIL forward+backward hot path.

Drop-in for one path of FabianPlum/SMILify: ``smal_model.smal_torch.SMAL`` (+ the 2D joint
reprojection losses built on it).  Host side is Python/PyTorch (device memory, streams,
autograd, torch.distributed); the arithmetic is hand-written CUDA behind a C ABI
(``include/smil_b200.h``, ``smilify_b200/csrc``).  No CPU fallback, no Triton.
"""
from . import config  # noqa: F401
from .config import configure  # noqa: F401

__all__ = ["config", "configure", "SMAL", "batch_rodrigues", "batch_global_rigid_transformation", "install"]
__version__ = "0.1.0"


def __getattr__(name):  # lazy: importing the package must not require torch.cuda or the .so
    if name == "SMAL":
        from .smal_torch import SMAL
        return SMAL
    if name in ("batch_rodrigues", "batch_global_rigid_transformation", "batch_skew"):
        from . import batch_lbs
        return getattr(batch_lbs, name)
    if name == "install":
        from .dropin import install
        return install
    raise AttributeError(name)
