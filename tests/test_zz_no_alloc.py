"""The allocation-free entry point ``smil_keypoint3d_loss_forward_static`` (include/smil_b200.h; no counterpart in the
reference): called through the C ABI and compared with the existing
``smil_keypoint3d_loss_forward`` on the same inputs, two calls in a row (the second one with a different, smaller batch,
so it runs on top of the first call's left-over scratch contents), plus the size limit.

Collected last on purpose (``zz``)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-6   # both entry points run the same two kernels: the results agree to the last bit in practice


def _scene(B, J, seed):
    rs = np.random.RandomState(seed)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(DEV)
    joints = f(rs.standard_normal((B, J, 3)))
    target = f(rs.standard_normal((B, J, 3)))
    vis = f((rs.uniform(size=(B, J)) > 0.3).astype(np.float32))
    vis[B // 2] = 0.0                       # a sample with no visible joint drops out of the mean
    weights = f(rs.uniform(0.5, 1.5, size=(J,)))
    mask = f((rs.uniform(size=(B,)) > 0.1).astype(np.float32))
    return joints, target, vis, weights, mask


def _args(L, joints, target, vis, weights, mask, loss, partials):
    from smilify_b200 import _lib

    a = _lib.SmilKp3dArgs()
    a.batch, a.num_joints = joints.shape[0], joints.shape[1]
    a.joints, a.target, a.visibility = joints.data_ptr(), target.data_ptr(), vis.data_ptr()
    a.joint_weights, a.sample_mask = weights.data_ptr(), mask.data_ptr()
    a.loss = loss.data_ptr()
    a.partials = partials.data_ptr() if partials is not None else None
    return a


def _both(B, J, seed):
    from smilify_b200 import _lib

    L = _lib.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    t = _scene(B, J, seed)
    loss_ref = torch.full((1,), float("nan"), device=DEV)
    loss_new = torch.full((1,), float("nan"), device=DEV)
    partials = torch.empty(3 * B + 4, device=DEV)
    a = _args(L, *t, loss_ref, partials)
    _lib.check(L.smil_keypoint3d_loss_forward(C.byref(a), st), "smil_keypoint3d_loss_forward")
    b = _args(L, *t, loss_new, None)        # no scratch passed: the library's static array is used
    _lib.check(L.smil_keypoint3d_loss_forward_static(C.byref(b), st), "smil_keypoint3d_loss_forward_static")
    torch.cuda.synchronize()
    return float(loss_ref), float(loss_new)


def test_static_scratch_entry_point_matches_the_existing_one_twice():
    ref1, new1 = _both(4096, 55, seed=1)
    assert np.isfinite(ref1) and abs(new1 - ref1) <= TOL * abs(ref1), (ref1, new1)
    # second call in a row: smaller batch, other data; the static array still holds call one's partials beyond 3 * 37
    ref2, new2 = _both(37, 31, seed=2)
    assert np.isfinite(ref2) and abs(new2 - ref2) <= TOL * abs(ref2), (ref2, new2)
    assert ref1 != ref2


def test_static_scratch_entry_point_rejects_sizes_beyond_the_limit():
    from smilify_b200 import _lib

    L = _lib.lib()
    B = _lib.KP3D_STATIC_MAX_BATCH + 1
    t = _scene(B, 3, seed=3)
    loss = torch.zeros(1, device=DEV)
    a = _args(L, *t, loss, None)
    rc = L.smil_keypoint3d_loss_forward_static(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -4 and b"exceeds the static scratch" in L.smil_last_error()
    # the limit itself is accepted
    t = _scene(_lib.KP3D_STATIC_MAX_BATCH, 3, seed=4)
    a = _args(L, *t, loss, None)
    _lib.check(L.smil_keypoint3d_loss_forward_static(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
               "smil_keypoint3d_loss_forward_static")
    torch.cuda.synchronize()
    assert np.isfinite(float(loss))
