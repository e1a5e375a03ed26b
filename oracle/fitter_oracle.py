"""TEST INFRASTRUCTURE ONLY (never imported by the product): torch-CPU restatement of the parameter-space terms of the
reference fitter objective and of its optimiser step.

Follows /root/reference/smal_fitter/fitter.py:298-350 (joint limits, pose prior, splay, betas prior inside
``SMALFitter.forward``; ``get_temporal``) as they are driven by /root/reference/smal_fitter/optimize_to_joints.py:150-175:
the per-window objective is evaluated for windows of ``config.WINDOW_SIZE`` consecutive frames and SUMMED over the
windows (every ``mean`` runs over one window), the temporal terms run over all frames once, then ONE ``Adam`` step
(optimize_to_joints.py:117-126, torch.optim.Adam defaults) updates the parameters.

Pinned against the reference's own code by ``oracle/pin_against_reference.py::pin_fitter_priors`` (the statements are
extracted from fitter.py by AST and executed): tests/golden/fitter_prior_pins.npz."""
from __future__ import annotations

import torch


def identity_pose_prior(num_joints, dtype=torch.float64):
    """The prior the reference builds for rigs without hard-coded body data (fitter.py:25-52): zero mean, identity
    precision, the root joint's three angles ignored."""
    P = 3 * num_joints
    use = torch.ones(P, dtype=dtype)
    use[:3] = 0
    return dict(mean=torch.zeros(P, dtype=dtype), precs=torch.eye(P, dtype=dtype), use=use)


def window_objective(global_rot, joint_rot, betas, w, max_limits=None, min_limits=None, pose_prior=None,
                     betas_mean=None, betas_prec=None, rotation_mask=None, global_mask=None):
    """One window (fitter.py:298-331).  ``global_rot`` (n, 3), ``joint_rot`` (n, J-1, 3), ``betas`` (nB,) shared by the
    frames (fitter.py:240-241 expands it).  ``w``: dict with w_limit / w_pose / w_splay / w_betas.  Returns a dict of terms."""
    n = global_rot.shape[0]
    g = global_rot * (global_mask if global_mask is not None else 1.0)       # fitter.py:242-243
    jr = joint_rot * (rotation_mask if rotation_mask is not None else 1.0)
    objs = {}
    if w.get("w_limit", 0) > 0 and max_limits is not None:                   # fitter.py:301-308
        zeros = torch.zeros_like(jr)
        objs["limit"] = w["w_limit"] * torch.mean(torch.max(jr - max_limits, zeros) + torch.max(min_limits - jr, zeros))
    if w.get("w_pose", 0) > 0 and pose_prior is not None:                    # fitter.py:310-317, Prior.__call__
        x = torch.cat([g.unsqueeze(1), jr], dim=1).reshape(n, -1)
        res = torch.tensordot(x - pose_prior["mean"].unsqueeze(0), pose_prior["precs"], dims=([1], [0])) * pose_prior["use"]
        objs["pose"] = w["w_pose"] * (res ** 2).mean()
    if w.get("w_splay", 0) > 0:                                              # fitter.py:319-320
        objs["splay"] = w["w_splay"] * torch.sum(jr[:, :, [0, 2]] ** 2)
    if w.get("w_betas", 0) > 0 and betas_prec is not None:                   # fitter.py:322-331
        all_betas = betas.unsqueeze(0).expand(n, -1)
        res = torch.tensordot(all_betas - betas_mean.unsqueeze(0), betas_prec, dims=([1], [0]))
        objs["betas"] = w["w_betas"] * (res ** 2).mean()
    return objs


def temporal_terms(global_rot, joint_rot, trans, w_temp, rotation_mask=None, global_mask=None):
    """fitter.py:337-350: sum over consecutive frame pairs of the three mean-squared differences."""
    import torch.nn.functional as F

    g = global_rot * (global_mask if global_mask is not None else 1.0)
    jr = joint_rot * (rotation_mask if rotation_mask is not None else 1.0)
    # the reference accumulates into fp32 scalars IN PLACE (fitter.py:341-348), whatever the dtype of the parameters
    jl, gl, tl = torch.tensor(0.0).float(), torch.tensor(0.0).float(), torch.tensor(0.0).float()
    for i in range(global_rot.shape[0] - 1):
        gl += F.mse_loss(g[i], g[i + 1]) * w_temp
        jl += F.mse_loss(jr[i], jr[i + 1]) * w_temp
        tl += F.mse_loss(trans[i], trans[i + 1]) * w_temp
    return jl, gl, tl


def epoch_objective(global_rot, joint_rot, trans, betas, w, window, **consts):
    """optimize_to_joints.py:154-172 without the image terms: windows summed, then the temporal terms."""
    N = global_rot.shape[0]
    total = torch.zeros((), dtype=global_rot.dtype)
    for j in range(0, N, window):
        objs = window_objective(global_rot[j:j + window], joint_rot[j:j + window], betas, w, **consts)
        for v in objs.values():
            total = total + v
    if w.get("w_temp", 0) > 0:
        jl, gl, tl = temporal_terms(global_rot, joint_rot, trans, w["w_temp"], consts.get("rotation_mask"), consts.get("global_mask"))
        total = total + jl.to(total.dtype) + gl.to(total.dtype) + tl.to(total.dtype)
    return total
