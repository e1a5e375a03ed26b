"""float64 numpy emulation of the *algorithms* the CUDA kernels implement, driven by the
packed tables of ``smilify_b200.packer``.  Test infrastructure only: it lets the CPU test
suite prove that the sparse / closed-form / item-major reformulation (and the hand-derived
backward of the chain and of Rodrigues) equals the dense oracle and its autograd, before any
GPU time is spent.  The CUDA kernels are transcriptions of these functions."""
from __future__ import annotations

import numpy as np

FLIP = np.array([1.0, -1.0, 1.0])


def rodrigues_fwd(theta):
    """theta (...,3) -> R (...,3,3); /root/reference/smal_model/batch_lbs.py:31-50."""
    a = np.sqrt(((theta + 1e-8) ** 2).sum(-1, keepdims=True))
    r = theta / a
    c, s = np.cos(a)[..., None], np.sin(a)[..., None]
    K = np.zeros(theta.shape[:-1] + (3, 3))
    K[..., 0, 1], K[..., 0, 2] = -r[..., 2], r[..., 1]
    K[..., 1, 0], K[..., 1, 2] = r[..., 2], -r[..., 0]
    K[..., 2, 0], K[..., 2, 1] = -r[..., 1], r[..., 0]
    return c * np.eye(3) + (1 - c) * r[..., :, None] * r[..., None, :] + s * K


def rodrigues_bwd(theta, G):
    tp = theta + 1e-8
    a = np.sqrt((tp ** 2).sum(-1, keepdims=True))
    r = theta / a
    c, s = np.cos(a), np.sin(a)
    kG = np.stack([G[..., 2, 1] - G[..., 1, 2], G[..., 0, 2] - G[..., 2, 0], G[..., 1, 0] - G[..., 0, 1]], -1)
    Gr = np.einsum("...ij,...j->...i", G + np.swapaxes(G, -1, -2), r)
    rGr = np.einsum("...i,...ij,...j->...", r, G, r)[..., None]
    tr = (G[..., 0, 0] + G[..., 1, 1] + G[..., 2, 2])[..., None]
    gc = tr - rGr
    gs = (r * kG).sum(-1, keepdims=True)
    gr = (1 - c) * Gr + s * kG
    ga = -s * gc + c * gs - (gr * theta).sum(-1, keepdims=True) / (a * a)
    return gr / a + ga * tp / a


def chain_fwd(t, R, Jrest, ls, bt, propagate):
    """R (B,J,3,3), Jrest (B,J,3) -> dict with GR,Gt,M,tl,s,pinv (all (B,J,...))."""
    B, J = R.shape[:2]
    par = t["parents"]
    s = np.exp(ls) if ls is not None else np.ones((B, J, 3))
    GR = np.zeros((B, J, 3, 3))
    Gt = np.zeros((B, J, 3))
    M = np.zeros((B, J, 3, 3))
    tl = np.zeros((B, J, 3))
    pinv = np.ones((B, J, 3))
    GR[:, 0], Gt[:, 0], M[:, 0] = R[:, 0], Jrest[:, 0], R[:, 0]
    for i in t["order"][1:]:
        p = par[i]
        if not propagate:
            pinv[:, i] = 1.0 / s[:, p]
        M[:, i] = pinv[:, i][:, :, None] * R[:, i] * s[:, i][:, None, :]
        tl[:, i] = Jrest[:, i] - Jrest[:, p] + (bt[:, i] * FLIP if bt is not None else 0.0)
        GR[:, i] = GR[:, p] @ M[:, i]
        Gt[:, i] = np.einsum("bij,bj->bi", GR[:, p], tl[:, i]) + Gt[:, p]
    At = Gt - np.einsum("bjik,bjk->bji", GR, Jrest)
    return dict(GR=GR, Gt=Gt, M=M, tl=tl, s=s, pinv=pinv, At=At)


def chain_bwd(t, R, Jrest, ls, bt, propagate, f, gAR, gAt, g_newJ):
    """VJP of chain_fwd.  gAR (B,J,3,3), gAt (B,J,3) grads of A=[GR | Gt-GR Jrest]; g_newJ of Gt."""
    B, J = R.shape[:2]
    par = t["parents"]
    gGR = gAR - gAt[..., :, None] * Jrest[..., None, :]
    gGt = gAt + (g_newJ if g_newJ is not None else 0.0)
    gJ = -np.einsum("bjki,bjk->bji", f["GR"], gAt)
    gR = np.zeros((B, J, 3, 3))
    gls = np.zeros((B, J, 3))
    gbt = np.zeros((B, J, 3))
    for i in t["order"][:0:-1]:
        p = par[i]
        gM = np.einsum("bki,bkj->bij", f["GR"][:, p], gGR[:, i])
        gtl = np.einsum("bki,bk->bi", f["GR"][:, p], gGt[:, i])
        gGR[:, p] += gGR[:, i] @ np.swapaxes(f["M"][:, i], -1, -2) + gGt[:, i][:, :, None] * f["tl"][:, i][:, None, :]
        gGt[:, p] += gGt[:, i]
        pin, s = f["pinv"][:, i], f["s"][:, i]
        gR[:, i] = pin[:, :, None] * gM * s[:, None, :]
        gls[:, i] += (pin[:, :, None] * R[:, i] * gM).sum(1) * s
        if not propagate:
            gls[:, p] -= (gM * R[:, i] * s[:, None, :]).sum(2) * pin
        gJ[:, i] += gtl
        gJ[:, p] -= gtl
        gbt[:, i] = gtl * FLIP
    gR[:, 0] = gGR[:, 0]
    gJ[:, 0] += gGt[:, 0]
    return gR, gJ, (gls if ls is not None else None), (gbt if bt is not None else None)


def fast_forward(t, beta, theta, trans, ls, bt, propagate, static, rotmat=False):
    """The fused path: closed-form rest joints and posed joints, ELL skinning."""
    B = theta.shape[0]
    V, J, nB = t["V"], t["J"], t["nB"]
    nb = beta.shape[1]
    R = theta if rotmat else rodrigues_fwd(theta)
    if static:
        Jrest = np.broadcast_to(t["J_static"].astype(np.float64).reshape(1, J, 3), (B, J, 3)).copy()
    else:
        Jrest = t["JT"].astype(np.float64)[None] + np.einsum("bm,mjc->bjc", beta, t["JS"].astype(np.float64)[:nb])
    f = chain_fwd(t, R, Jrest, ls, bt, propagate)
    vs = t["v_template"].astype(np.float64)[None] + (beta @ t["shapedirs"].astype(np.float64)[:nb]).reshape(B, V, 3)
    verts = np.zeros((B, V, 3))
    for k in range(t["K"]):
        b_id = t["skin_idx"][:, k]
        w = t["skin_w"][:, k].astype(np.float64)
        verts += w[None, :, None] * (np.einsum("bvij,bvj->bvi", f["GR"][:, b_id], vs) + f["At"][:, b_id])
    tr = trans if trans is not None else np.zeros((B, 3))
    verts = verts + tr[:, None]
    if static:
        joints = f["Gt"].copy()
    else:
        joints = t["jreg_rowsum"].astype(np.float64)[None, :, None] * tr[:, None, :]
        for pi in range(t["npairs"]):
            j, k = t["pj_joint"][pi], t["pj_bone"][pi]
            c0 = t["pj_c0"][pi].astype(np.float64)
            q = c0[None, :3] + beta @ t["pj_cm"][pi].astype(np.float64)[:nb]
            joints[:, j] += np.einsum("bij,bj->bi", f["GR"][:, k], q) + f["At"][:, k] * c0[3]
    return dict(verts=verts, joints=joints, Rs=R, v_shaped=vs, J_transformed=f["Gt"], chain=f, Jrest=Jrest)


def skin_backward_vm(t, beta, GR, gv):
    """Vertex-major backward skin pass (csrc/skin_bwd_vm.cuh) driven by the ``vm_*`` tables: per tile and warp share a
    list of segments; two resident bones (slots X, Y) with a 3x4 accumulator each; per vertex the shape blend,
    u = sum_k w_k R_k^T g and the dL/dbeta contraction once; flushes write bone-major numbered partial slots."""
    B = beta.shape[0]
    J, nB = t["J"], t["nB"]
    nb = beta.shape[1]
    recf = t["vm_recf"]
    rec = t["vm_rec"].astype(np.float64)
    seg = t["vm_seg"]
    slots = np.zeros((t["vm_nslots"], B, 3, 4))
    written = np.zeros(t["vm_nslots"], dtype=int)
    g_beta = np.zeros((B, nb))
    g_trans = np.zeros((B, 3))
    C = t["vm_chunk"]
    seen = np.zeros(t["V"], dtype=int)
    NPH, NW = t["vm_nphase"], t["vm_warps"]
    PW = C // NPH
    wsp = t["vm_warp_seg_ptr"]
    for ti in range(t["vm_ntile"]):
        rbase = t["vm_tile_rec_ptr"][ti]
        tile_seg0 = wsp[ti * NPH * NW]
        for w in range(NW):
            acc = np.zeros((2, B, 3, 4))
            bone = [None, None]
            prev = None
            for ph in range(NPH):
                for sg in range(wsp[(ti * NPH + ph) * NW + w], wsp[(ti * NPH + ph) * NW + w + 1]):
                    lx, ly, first, cnt, fx, fy, mode, nxt = (int(x) for x in seg[sg])
                    if prev is not None:
                        assert prev == sg - tile_seg0, "segment links do not follow the warp's processing order"
                    prev = nxt if nxt >= 0 else None
                    if lx >= 0:
                        assert not acc[0].any(), "slot X reloaded with an unflushed accumulator"
                        bone[0] = lx
                    if ly >= 0:
                        assert not acc[1].any(), "slot Y reloaded with an unflushed accumulator"
                        bone[1] = ly
                    for it in range(rbase + first, rbase + first + cnt):
                        r = rec[it]
                        wX, wY = r[0], r[1]
                        lv = int(t["vm_rec"][it, 2].view(np.int32)) // 3
                        assert 0 <= lv < PW
                        v = int(rbase) + ph * PW + lv   # tiles are consecutive vertex ranges, one record per vertex
                        seen[v] += 1
                        g = gv[:, v]
                        S_ = r[6:6 + 3 * nB].reshape(nB, 3)[:nb]
                        vs = r[3:6][None] + beta @ S_
                        g_trans += g
                        u = np.zeros((B, 3))
                        for s_, wgt in ((0, wX), (1, wY)):
                            if mode == 1 and s_ == 1:
                                assert wgt == 0.0
                                continue
                            if wgt == 0.0:
                                continue
                            wg = wgt * g
                            acc[s_][:, :, :3] += wg[:, :, None] * vs[:, None, :]
                            acc[s_][:, :, 3] += wg
                            u += np.einsum("bki,bk->bi", GR[:, bone[s_]], wg)
                        g_beta += u @ S_.T
                    for s_, fl in ((0, fx), (1, fy)):
                        if fl >= 0:
                            assert t["vm_slot_bone"][fl] == bone[s_], "flush slot belongs to another bone"
                            slots[fl] = acc[s_]
                            written[fl] += 1
                            acc[s_] = 0.0
            assert prev is None
            assert not acc.any(), "a warp share ended with an unflushed accumulator"
    assert (written == 1).all() and (seen == 1).all()
    gA = np.zeros((B, J, 3, 4))
    for k in range(J):
        for si in range(t["vm_sb_ptr"][k], t["vm_sb_ptr"][k + 1]):
            gA[:, k] += slots[si]
    return gA, g_beta, g_trans


def fast_backward(t, beta, theta, trans, ls, bt, propagate, static, fw, g_verts, g_joints, g_Rs=None, rotmat=False,
                  vertex_major=False):
    """Item-major backward of the fused path (what skin_bwd + pose_bwd compute)."""
    B = theta.shape[0]
    V, J, nB = t["V"], t["J"], t["nB"]
    nb = beta.shape[1]
    f = fw["chain"]
    C = t["bwd_chunk"]
    rec = t["vrec"].astype(np.float64)
    g_beta = np.zeros((B, nb))
    g_trans = np.zeros((B, 3))
    nun = t["nunits"]
    slots = np.zeros((B, nun, 3, 4))
    gv = g_verts if g_verts is not None else np.zeros((B, V, 3))
    for u in range(nun):
        k = t["unit_bone"][u]
        base = t["unit_tile"][u] * C
        AkR = f["GR"][:, k]
        for it in range(t["unit_ptr"][u], t["unit_ptr"][u + 1]):
            v = base + t["it_v"][it]
            w = float(t["it_w"][it])
            g = gv[:, v]
            T_, S_ = rec[v, :3], rec[v, 3:3 + 3 * nB].reshape(nB, 3)[:nb]
            vp = T_[None] + beta @ S_
            uvec = np.einsum("bki,bk->bi", AkR, g)
            g_beta += w * (uvec @ S_.T)
            wg = w * g
            slots[:, t["unit_slot"][u], :, :3] += wg[:, :, None] * vp[:, None, :]
            slots[:, t["unit_slot"][u], :, 3] += wg
            g_trans += float(t["it_first"][it]) * g
    gA = np.zeros((B, J, 3, 4))
    for k in range(J):
        for si in range(t["sb_ptr"][k], t["sb_ptr"][k + 1]):
            gA[:, k] += slots[:, si]
    if vertex_major:
        gA, g_beta, g_trans = skin_backward_vm(t, beta, f["GR"], gv)
    g_newJ = None
    if static:
        g_newJ = g_joints
    elif g_joints is not None:
        g_trans += np.einsum("j,bjc->bc", t["jreg_rowsum"].astype(np.float64), g_joints)
        for pi in range(t["npairs"]):
            j, k = t["pj_joint"][pi], t["pj_bone"][pi]
            c0 = t["pj_c0"][pi].astype(np.float64)
            cm = t["pj_cm"][pi].astype(np.float64)[:nb]
            q = c0[None, :3] + beta @ cm
            gj = g_joints[:, j]
            gA[:, k, :, :3] += gj[:, :, None] * q[:, None, :]
            gA[:, k, :, 3] += gj * c0[3]
            g_beta += np.einsum("bki,bk->bi", f["GR"][:, k], gj) @ cm.T
    gR, gJ, gls, gbt = chain_bwd(t, fw["Rs"], fw["Jrest"], ls, bt, propagate, f, gA[..., :3], gA[..., 3], g_newJ)
    if g_Rs is not None:
        gR = gR + g_Rs
    if not static:
        g_beta += np.einsum("bjc,mjc->bm", gJ, t["JS"].astype(np.float64)[:nb])
    g_theta = gR if rotmat else rodrigues_bwd(theta, gR)
    return dict(beta=g_beta, theta=g_theta, trans=g_trans, betas_logscale=gls, betas_trans=gbt)
