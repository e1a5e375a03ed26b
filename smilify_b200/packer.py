"""Load-time packer: reference-layout model arrays -> compact tables for the CUDA kernels.

Input is exactly what ``SMAL.__init__`` holds (/root/reference/smal_model/smal_torch.py:120-196):
``v_template (V,3)``, ``shapedirs (nB,3V)``, ``posedirs (P,3V)``, ``J_regressor (V,J)``,
``weights (V,J)``, ``parents int32 (J,)``, optional static ``J (J,3)``.  Output is a dict of
flat numpy arrays (fp32 / int32) that ``smil_model_create`` uploads verbatim.

What the packing exploits ("hard parts"):
  * skinning weights have <= K non-zeros per vertex (K = 2 for every SMIL rig shipped)
    -> ELL ``(bone, weight)`` pairs instead of the dense (V,J) matrix the reference
    multiplies with (smal_torch.py:320-323);
  * ``J_regressor`` is sparse (STICK: 1 544 of 166 100) -> CSC/CSR;
  * rest joints are linear in beta:  J = Jreg^T v_template + sum_m beta_m Jreg^T S_m;
  * posed joints are linear in the bone transforms:
        joints_j = sum_k A_k [c0_jk + sum_m beta_m cm_jkm ; .] + rowsum_j * trans
    with c0_jk = sum_v Jreg[v,j] w[v,k] [T_v;1], cm_jkm = sum_v Jreg[v,j] w[v,k] S_m[v]
    -> both regressions leave the per-vertex pass entirely (tables ``pj_*``);
  * the kinematic tree is walked level by level -> ``order`` / ``level_ptr`` / children CSR;
  * all-zero ``posedirs`` (every SMIL rig) is detected and the 9(J-1) x 3V GEMM skipped.

All index arrays are int32 and bit-exact with the reference's
``kintree_table[0].astype(np.int32)`` (smal_torch.py:193).
"""
from __future__ import annotations

import numpy as np

FWD_CHUNK = 256  # vertices per CTA of the fused forward skinning pass (one per thread)
import os as _os
BWD_CHUNK = int(_os.environ.get("SMIL_BWD_CHUNK", "192"))  # upper bound of the vertices per CTA tile of the fused backward pass


def backward_chunk(V: int, limit: int = None) -> int:
    """Tile size of the fused backward pass: the fewest tiles of at most ``limit`` vertices, made equal
    (STICK: 16 tiles of 189 instead of 15 x 192 + 140) so that every CTA column carries the same work."""
    limit = BWD_CHUNK if limit is None else limit
    ntile = max(1, -(-V // limit))
    return max(1, -(-V // ntile))
BWD_WARPS = 4    # warps per CTA of the fused backward pass
MAX_FUSED_K = 4
MAX_FUSED_NB = 16
NBT_VALUES = (1, 2, 3, 4, 5, 6, 8, 10, 13, 16)  # beta-count template instances of the fused kernels


def nbt_of(nb: int) -> int:
    """Smallest compiled beta-count instance >= nb (0 if the fused kernels do not cover nb)."""
    for v in NBT_VALUES:
        if v >= nb:
            return v
    return 0


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def tree_tables(parents: np.ndarray):
    """Levels, topological order and children lists of the kinematic tree.

    ``parents[i] < i`` is required for i >= 1 -- the reference indexes a growing Python list
    with it (batch_lbs.py:157,179) and would raise IndexError otherwise; the root is joint 0
    whatever ``parents[0]`` holds (batch_lbs.py:151-152)."""
    J = int(parents.shape[0])
    par = parents.astype(np.int64).copy()
    par[0] = -1
    for i in range(1, J):
        if not (0 <= par[i] < i):
            raise IndexError("parents[%d]=%d: the kinematic chain needs 0 <= parents[i] < i" % (i, par[i]))
    depth = np.zeros(J, dtype=np.int64)
    for i in range(1, J):
        depth[i] = depth[par[i]] + 1
    order = np.argsort(depth, kind="stable")
    nlev = int(depth.max()) + 1
    level_ptr = np.zeros(nlev + 1, dtype=np.int64)
    for d in depth:
        level_ptr[d + 1] += 1
    level_ptr = np.cumsum(level_ptr)
    child_ptr = np.zeros(J + 1, dtype=np.int64)
    for i in range(1, J):
        child_ptr[par[i] + 1] += 1
    child_ptr = np.cumsum(child_ptr)
    child_idx = np.zeros(max(J - 1, 1), dtype=np.int64)
    fill = child_ptr[:-1].copy()
    for i in range(1, J):
        child_idx[fill[par[i]]] = i
        fill[par[i]] += 1
    widths = np.diff(level_ptr)
    return dict(
        parents=_i32(par), order=_i32(order), level_ptr=_i32(level_ptr), child_ptr=_i32(child_ptr),
        child_idx=_i32(child_idx), depth=_i32(depth), num_levels=nlev, max_width=int(widths.max()),
    )


PACK_FORMAT = 9   # bump when the meaning of a packed table changes: invalidates cached packs


def pack_key(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static=None) -> str:
    """Hash of the rig arrays, the pack format and the packer's tuning switches: names a pack in the on-disk cache and
    inside a flat rig file (model_io.export_flat)."""
    import hashlib

    h = hashlib.sha1(repr((PACK_FORMAT, FWD_CHUNK, BWD_CHUNK, VM_CHUNK, VM_PHASES, VM_WARPS, VM_CUT_TOL, VM_SEG_COST, VM_FLUSH_COST, VM_TILE_BALANCE, VM_CHUNK_MAX,
                           sorted((k, v) for k, v in _os.environ.items() if k.startswith("SMIL_")))).encode())
    for a in (v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static):
        a = np.ascontiguousarray(np.asarray(a)) if a is not None else np.zeros(0)
        h.update(str((a.shape, a.dtype.str)).encode())
        h.update(a.tobytes())
    return h.hexdigest()[:24]


def pack_model_cached(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static=None, cache_dir=None, source=None):
    """:func:`pack_model` through an on-disk cache: the packed tables of a rig as ONE flat ``.npz`` (every table a
    plain fp32 / int32 array, scalars as 0-d arrays), keyed by a hash of the rig arrays, the pack format and the
    packer's tuning switches.  The anchor search and the run ordering take seconds per rig; the 8 ranks of a launch
    and every later process load the file instead.  ``SMILIFY_B200_CACHE`` names the directory ("" disables)."""
    import tempfile

    key = pack_key(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static)
    if source is not None:      # a flat rig file that carries its own pack (model_io.export_flat)
        from .model_io import embedded_pack

        t = embedded_pack(source, key)
        if t is not None:
            return t
    cache_dir = _os.environ.get("SMILIFY_B200_CACHE", _os.path.join(_os.path.expanduser("~"), ".cache", "smilify_b200")) \
        if cache_dir is None else cache_dir
    if not cache_dir:
        return pack_model(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static)
    path = _os.path.join(cache_dir, "pack_%s.npz" % key)
    if _os.path.exists(path):
        try:
            with np.load(path, allow_pickle=False) as z:
                return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}
        except Exception:   # a torn or stale file: repack
            pass
    t = pack_model(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static)
    try:
        _os.makedirs(cache_dir, exist_ok=True)
        fd, tmp = tempfile.mkstemp(dir=cache_dir, suffix=".npz")
        _os.close(fd)
        np.savez(tmp, **{k: np.asarray(v) for k, v in t.items()})
        _os.replace(tmp, path)    # atomic: concurrent ranks either see the whole file or none
    except OSError:
        pass
    return t


def pack_model(v_template, shapedirs, posedirs, J_regressor, weights, parents, J_static=None):
    vt = np.asarray(v_template, dtype=np.float64)
    V = vt.shape[0]
    sd = np.asarray(shapedirs, dtype=np.float64)
    nB = sd.shape[0]
    assert sd.shape == (nB, 3 * V), sd.shape
    W = np.asarray(weights, dtype=np.float64)
    J = W.shape[1]
    jr = np.asarray(J_regressor, dtype=np.float64)
    assert jr.shape == (V, J) and W.shape == (V, J)
    t = dict(V=V, J=J, nB=nB)

    # ---- posedirs ------------------------------------------------------------------------
    P = 0
    pd = None
    if posedirs is not None:
        pd = np.asarray(posedirs)
        if pd.ndim == 2 and pd.shape[1] == 3 * V and pd.shape[0] == 9 * (J - 1) and np.any(pd != 0):
            P = pd.shape[0]
        elif pd.size and np.any(pd != 0):
            raise ValueError("posedirs shape %s is not (9(J-1), 3V)" % (pd.shape,))
    t["P"] = P
    t["posedirs"] = _f32(pd) if P else np.zeros(0, np.float32)

    t["v_template"] = _f32(vt)
    t["shapedirs"] = _f32(sd)
    t["J_static"] = _f32(J_static) if J_static is not None else np.zeros(0, np.float32)
    t.update(tree_tables(np.asarray(parents)))

    # ---- skinning ELL ----------------------------------------------------------------------
    nnz_row = (W != 0).sum(1)
    K = max(int(nnz_row.max()), 1)
    idx = np.zeros((V, K), dtype=np.int32)
    val = np.zeros((V, K), dtype=np.float32)
    for v in range(V):
        cols = np.nonzero(W[v])[0]
        idx[v, : len(cols)] = cols
        val[v, : len(cols)] = W[v, cols]
    t["K"] = K
    t["skin_idx"] = _i32(idx)
    t["skin_w"] = _f32(val)
    t["skin_cnt"] = _i32(nnz_row)

    # ---- joint regressor CSC (by joint) / CSR (by vertex) ------------------------------------
    vv, jj = np.nonzero(jr)
    o = np.lexsort((vv, jj))
    t["jc_ptr"] = _i32(np.concatenate([[0], np.cumsum(np.bincount(jj, minlength=J))]))
    t["jc_vid"] = _i32(vv[o]) if len(o) else np.zeros(1, np.int32)
    t["jc_val"] = _f32(jr[vv[o], jj[o]]) if len(o) else np.zeros(1, np.float32)
    o = np.lexsort((jj, vv))
    t["jr_ptr"] = _i32(np.concatenate([[0], np.cumsum(np.bincount(vv, minlength=V))]))
    t["jr_jid"] = _i32(jj[o]) if len(o) else np.zeros(1, np.int32)
    t["jr_val"] = _f32(jr[vv[o], jj[o]]) if len(o) else np.zeros(1, np.float32)
    t["jreg_nnz"] = int(len(vv))

    # ---- closed-form regressions -------------------------------------------------------------
    S3 = sd.reshape(nB, V, 3)
    t["JT"] = _f32(jr.T @ vt)                                    # (J,3)
    t["JS"] = _f32(np.einsum("vj,mvc->mjc", jr, S3))              # (nB,J,3)
    t["jreg_rowsum"] = _f32(jr.sum(0))                            # (J,)
    X = np.concatenate([vt, np.ones((V, 1)), S3.transpose(1, 0, 2).reshape(V, 3 * nB)], axis=1)  # (V, 4+3nB)
    pj_ptr = [0]
    pj_bone, pj_joint, pj_c0, pj_cm = [], [], [], []
    for j in range(J):
        rows = np.nonzero(jr[:, j])[0]
        if len(rows):
            coup = (jr[rows, j][:, None, None] * W[rows][:, :, None] * X[rows][:, None, :]).sum(0)  # (J, 4+3nB)
            bones = np.nonzero((W[rows] != 0).any(0))[0]
            for k in bones:
                pj_bone.append(k)
                pj_joint.append(j)
                pj_c0.append(coup[k, :4])
                pj_cm.append(coup[k, 4:].reshape(nB, 3))
        pj_ptr.append(len(pj_bone))
    npairs = len(pj_bone)
    t["npairs"] = npairs
    t["pj_ptr"] = _i32(pj_ptr)
    t["pj_bone"] = _i32(pj_bone) if npairs else np.zeros(1, np.int32)
    t["pj_joint"] = _i32(pj_joint) if npairs else np.zeros(1, np.int32)
    t["pj_c0"] = _f32(pj_c0).reshape(-1, 4) if npairs else np.zeros((1, 4), np.float32)
    t["pj_cm"] = _f32(pj_cm).reshape(-1, nB, 3) if npairs else np.zeros((1, max(nB, 1), 3), np.float32)
    # the same pairs grouped by bone (backward accumulates g_A per bone without atomics)
    pb = np.argsort(np.asarray(pj_bone, dtype=np.int64), kind="stable") if npairs else np.zeros(0, np.int64)
    t["pb_ptr"] = _i32(np.concatenate([[0], np.cumsum(np.bincount(np.asarray(pj_bone, dtype=np.int64), minlength=J))])
                       if npairs else np.zeros(J + 1))
    t["pb_pair"] = _i32(pb) if npairs else np.zeros(1, np.int32)
    # the same pairs once more, bone-sorted and 32-padded, one contiguous record each, for the backward's
    # lane-per-pair pass: (bone | joint | c0[4] | cm[nB][3] | pad); padding records carry bone -1
    rp = (6 + 3 * nB + 3) // 4 * 4
    npad = (npairs + 31) // 32 * 32
    pbrec = np.zeros((max(npad, 1), rp), dtype=np.float32)
    pbrec[:, 0] = np.int32(-1).view(np.float32)
    for s_, pi in enumerate(pb):
        pbrec[s_, 0] = np.int32(pj_bone[pi]).view(np.float32)
        pbrec[s_, 1] = np.int32(pj_joint[pi]).view(np.float32)
        pbrec[s_, 2:6] = pj_c0[pi]
        pbrec[s_, 6:6 + 3 * nB] = np.asarray(pj_cm[pi], dtype=np.float32).reshape(-1)
    t["pbrec"] = _f32(pbrec)
    t["pbrec_floats"] = rp
    t["npairs_pad"] = npad

    # ---- fused forward skinning pass: per-vertex constants, SoA over a padded vertex axis -----
    nchunk = (V + FWD_CHUNK - 1) // FWD_CHUNK
    Vpad = nchunk * FWD_CHUNK
    assigned = forward_thread_assignment(idx, val, nnz_row, V, FWD_CHUNK)
    fwd_pos, owner = assigned[0], assigned[1]
    has = owner >= 0
    nconst = 3 + 3 * nB
    vconst = np.zeros((nconst, Vpad), dtype=np.float32)
    vconst[:3, has] = vt.T[:, owner[has]]
    vconst[3:, has] = S3.transpose(0, 2, 1).reshape(3 * nB, V)[:, owner[has]]      # row 3+3m+c = S_m[:,c]
    ell_i = np.zeros((K, Vpad), dtype=np.int32)
    ell_w = np.zeros((K, Vpad), dtype=np.float32)
    ell_i[:, has] = idx.T[:, owner[has]]
    ell_w[:, has] = val.T[:, owner[has]]
    # an empty skinning slot may name any bone (its weight is 0): name the vertex' first bone, so that the
    # slot's shared-memory read hits an address the warp already touches
    for k in range(1, K):
        empty = has & (ell_w[k] == 0)
        ell_i[k, empty] = ell_i[0, empty]
    if len(assigned) == 4:   # the assignment chose the slot order itself
        ell_i, ell_w = assigned[2], assigned[3]
    t.update(fwd_chunk=FWD_CHUNK, fwd_nchunk=nchunk, Vpad=Vpad, vconst=_f32(vconst), ell_idx=_i32(ell_i),
             ell_w=_f32(ell_w), fwd_pos=_i32(fwd_pos))

    # ---- fused backward pass: item-major tables ------------------------------------------------
    bchunk = backward_chunk(V)
    t.update(pack_backward_items(vt, S3, idx, val, nnz_row, J, nB, chunk=bchunk))
    # vertices of every backward tile that carry joint-regressor entries (the backward folds
    # Jreg g_joints into the staged g_verts tile: joints = Jreg^T verts, smal_torch.py:348-351)
    nt = t["bwd_ntile"]
    has_j = np.diff(t["jr_ptr"]) > 0
    tjv_ptr, tjv_v = [0], []
    for ti in range(nt):
        v0, v1 = ti * bchunk, min(V, (ti + 1) * bchunk)
        tjv_v.extend((np.nonzero(has_j[v0:v1])[0]).tolist())
        tjv_ptr.append(len(tjv_v))
    t["tjv_ptr"] = _i32(tjv_ptr)
    t["tjv_v"] = _i32(tjv_v + [0] * (max(t["jreg_nnz"], 1) - len(tjv_v)))  # padded to the size the library copies
    t["fused_ok"] = bool(K <= MAX_FUSED_K and nbt_of(nB) > 0 and P == 0)
    t["vm_ok"] = False
    if t["fused_ok"] and K <= 2 and _os.environ.get("SMIL_BWD_VM", "1") != "0":
        t.update(pack_backward_vm_balanced(vt, S3, idx, val, nnz_row, J, nB))
    return t


def forward_thread_assignment(idx, val, nnz_row, V, chunk):
    # default: "anchor" for rigs with at most two weights per vertex (every SMIL rig; measured 80.4 vs 81.8 us for the
    # forward skin pass of STICK at B = 4096), "uniform" otherwise; "spread" is kept for comparison
    mode = _os.environ.get("SMIL_FWD_ASSIGN", "anchor")
    if mode == "anchor" and idx.shape[1] == 2:
        return forward_thread_assignment_anchor(idx, val, nnz_row, V, chunk)
    if mode in ("anchor", "uniform"):
        return forward_thread_assignment_uniform(idx, val, nnz_row, V, chunk)
    return forward_thread_assignment_spread(idx, val, nnz_row, V, chunk)


def _lds128_cost(d):
    """Shared-memory wavefronts of one 16-byte warp load whose lanes name ``d`` distinct addresses (measured on B200,
    tools_micro/lds_gather.cu / profiles/r01_lds_gather_micro.csv)."""
    return 2 if d <= 2 else 4 if d <= 4 else 6 if d <= 8 else 8


def _refine_anchor_warps(a0, a1, members, sweeps=4, w_load=0.77, w_store=0.61):
    """Local search over the warps of one chunk: vertices that sit in a staging-store bank conflict (a second vertex of
    their warp has the same position mod 32) are swapped with a vertex of the SAME anchor in another warp when that
    lowers the modelled cost  ``w_load * 6 * (c(d0) + c(d1)) + w_store * 12 * max multiplicity``  of the two warps
    (c = :func:`_lds128_cost`; the weights are the measured us per wavefront and warp pass of loads / stores in
    ``skin_fwd``).  Same-anchor swaps leave slot 0 untouched, so the search stays small
    (about a second for STICK); deterministic."""
    nw = len(members)
    wof = {}
    for w, ms in enumerate(members):
        for i in ms:
            wof[i] = w
    c0 = [dict() for _ in range(nw)]
    c1 = [dict() for _ in range(nw)]
    cr = [dict() for _ in range(nw)]

    def mv(w, i, sgn):
        for d, key in ((c0[w], a0[i]), (cr[w], i & 31)) + (((c1[w], a1[i]),) if a1[i] >= 0 else ()):
            n = d.get(key, 0) + sgn
            if n:
                d[key] = n
            else:
                del d[key]

    for i, w in wof.items():
        mv(w, i, 1)

    def cost(w):
        if not cr[w]:
            return 0.0
        return (w_load * 6 * (_lds128_cost(len(c0[w])) + _lds128_cost(len(c1[w]))) + w_store * 12 * max(cr[w].values())
                + 1e-3 * sum(x * x for x in cr[w].values()) - 1e-4 * sum(x * x for x in c1[w].values()))

    by_anchor = {}
    for i in sorted(wof):
        by_anchor.setdefault((a0[i], a1[i] >= 0), []).append(i)   # never trade a two-bone vertex into a single-bone warp
    costs = [cost(w) for w in range(nw)]
    for _ in range(sweeps):
        improved = False
        for i in sorted(wof):
            wa = wof[i]
            if cr[wa].get(i & 31, 0) < 2:
                continue
            best = None
            for j in by_anchor[(a0[i], a1[i] >= 0)]:
                wb = wof[j]
                if wb == wa:
                    continue
                mv(wa, i, -1); mv(wb, j, -1); mv(wa, j, 1); mv(wb, i, 1)
                d = cost(wa) + cost(wb) - costs[wa] - costs[wb]
                mv(wa, j, -1); mv(wb, i, -1); mv(wa, i, 1); mv(wb, j, 1)
                if d < -1e-9 and (best is None or d < best[0]):
                    best = (d, j, wb)
            if best:
                _, j, wb = best
                mv(wa, i, -1); mv(wb, j, -1); mv(wa, j, 1); mv(wb, i, 1)
                wof[i], wof[j] = wb, wa
                costs[wa], costs[wb] = cost(wa), cost(wb)
                improved = True
        if not improved:
            break
    out = [[] for _ in range(nw)]
    for i in sorted(wof):
        out[wof[i]].append(i)
    return out


_ANCHOR_MEMO = {}


def forward_thread_assignment_anchor(idx, val, nnz_row, V, chunk):
    """Memoised per process on the skinning tables (the search below takes about a second per rig)."""
    import hashlib

    key = hashlib.sha1(np.ascontiguousarray(idx).tobytes() + np.ascontiguousarray(val).tobytes() +
                       np.ascontiguousarray(nnz_row).tobytes() + repr((V, chunk, _os.environ.get("SMIL_FWD_REFINE", "1"), _os.environ.get("SMIL_FWD_SINGLE_FIRST", "0"))).encode()
                       ).hexdigest()
    if key not in _ANCHOR_MEMO:
        _ANCHOR_MEMO[key] = _forward_thread_assignment_anchor(idx, val, nnz_row, V, chunk)
    return tuple(a.copy() for a in _ANCHOR_MEMO[key])


def _forward_thread_assignment_anchor(idx, val, nnz_row, V, chunk):
    """Thread -> vertex map for rigs with at most two weights per vertex, built on the measured cost of the bone
    transform gathers (tools_micro/lds_gather.cu: a 16-byte shared load costs 2 wavefronts while the lanes of a warp
    name <= 2 distinct addresses, 4 for 3-4, 6-8 beyond): what matters per skinning slot and warp is the NUMBER of
    distinct bones, and which of a vertex' two bones sits in slot 0 is free.  Per chunk the bones are ranked by how
    many vertices carry them; every vertex puts its higher-ranked bone (its anchor) into slot 0 and the vertices are
    dealt to the warps in (anchor, partner) order, so that slot 0 of a warp sees one or two anchors and slot 1 a short
    run of partners; the unused slot 1 of a single-bone vertex names the most common partner of its warp.  A local
    search (:func:`_refine_anchor_warps`) then trades vertices of the same anchor between warps to remove bank conflicts
    of the staging stores.
    Returns ``(fwd_pos, owner, ell_idx (2, Vpad), ell_w (2, Vpad))``."""
    nchunk = (V + chunk - 1) // chunk
    Vpad = nchunk * chunk
    pos = np.full(Vpad, -1, dtype=np.int64)
    ell_i = np.zeros((2, Vpad), dtype=np.int32)
    ell_w = np.zeros((2, Vpad), dtype=np.float32)
    for c in range(nchunk):
        v0, v1 = c * chunk, min(V, (c + 1) * chunk)
        n = v1 - v0
        bones = [tuple(int(idx[v, k]) for k in range(min(int(nnz_row[v]), 2))) or (int(idx[v, 0]),) for v in range(v0, v1)]
        wts = [tuple(float(val[v, k]) for k in range(min(int(nnz_row[v]), 2))) or (0.0,) for v in range(v0, v1)]
        freq = {}
        for t in bones:
            for b in t:
                freq[b] = freq.get(b, 0) + 1
        prio = {b: r for r, (b, _) in enumerate(sorted(freq.items(), key=lambda kv: (-kv[1], kv[0])))}
        a0, a1, w0, w1 = [], [], [], []
        for t, w in zip(bones, wts):
            k = min(range(len(t)), key=lambda q: prio[t[q]])
            a0.append(t[k]); w0.append(w[k])
            if len(t) > 1:
                a1.append(t[1 - k]); w1.append(w[1 - k])
            else:
                a1.append(-1); w1.append(0.0)
        if _os.environ.get("SMIL_FWD_SINGLE_FIRST", "0") != "0":   # measured on B200: 76.4 vs 73.7 us (STICK) -> off
            # single-bone vertices first (grouped by bone), then the two-bone ones by (anchor, partner): whole warps of
            # single-bone vertices skip the second skinning slot in the kernel (a warp-uniform branch)
            order = sorted(range(n), key=lambda i: (a1[i] >= 0, prio[a0[i]], a1[i], i))
        else:
            order = sorted(range(n), key=lambda i: (prio[a0[i]], a1[i] < 0, a1[i], i))
        members = [order[w * 32:(w + 1) * 32] for w in range((n + 31) // 32)]
        if _os.environ.get("SMIL_FWD_REFINE", "1") != "0":
            members = _refine_anchor_warps(a0, a1, members)
        for w, ms in enumerate(members):
            partners = [a1[i] for i in ms if a1[i] >= 0]
            fill = max(set(partners), key=partners.count) if partners else -1
            lanes = [-1] * 32
            left = []
            for i in ms:      # staging stores stage[3*pos + c]: distinct positions mod 32 keep them conflict-free
                r = i & 31
                if lanes[r] < 0:
                    lanes[r] = i
                else:
                    left.append(i)
            free = [q for q in range(32) if lanes[q] < 0]
            for i, q in zip(left, free):
                lanes[q] = i
            for lane, i in enumerate(lanes):
                if i < 0:
                    continue
                t = c * chunk + w * 32 + lane
                pos[t] = i
                ell_i[0, t], ell_w[0, t] = a0[i], w0[i]
                ell_i[1, t] = a1[i] if a1[i] >= 0 else (fill if fill >= 0 else a0[i])
                ell_w[1, t] = w1[i]
    owner = np.where(pos >= 0, pos + (np.arange(Vpad) // chunk) * chunk, -1)
    return pos, owner, ell_i, ell_w


def forward_thread_assignment_uniform(idx, val, nnz_row, V, chunk):
    """Thread -> vertex map of the fused forward pass that makes as many warps as possible BONE-UNIFORM.

    The pass reads a vertex' bone transforms from shared memory with 16-byte loads at per-thread addresses.
    Such a load costs four wavefronts whenever the lanes of a warp name more than one address and ONE when
    they all name the same (measured, ncu r01: the warp-uniform beta / trans loads), so the vertices of a
    chunk are grouped by their bone pair: every group first fills whole warps, the remainders are packed
    together in group order.  Inside a warp the lanes are then permuted so that their staging positions are
    as distinct as possible modulo 32 (the staging stores ``stage[3*pos + c]`` are conflict-free iff they are).
    Same return values as :func:`forward_thread_assignment_spread`."""
    nchunk = (V + chunk - 1) // chunk
    nwarp = chunk // 32
    K = idx.shape[1]
    pos = np.full(nchunk * chunk, -1, dtype=np.int64)

    def combo(v):
        b0 = int(idx[v, 0])
        b1 = int(idx[v, 1]) if K > 1 and nnz_row[v] > 1 else b0
        rest = tuple(int(idx[v, k]) for k in range(2, K) if nnz_row[v] > k)
        return (b0, b1) + rest

    for c in range(nchunk):
        v0, v1 = c * chunk, min(V, (c + 1) * chunk)
        groups = {}
        for v in range(v0, v1):
            groups.setdefault(combo(v), []).append(v)
        warps, rest = [], []
        for key in sorted(groups, key=lambda k_: (-len(groups[k_]), k_)):
            vs = groups[key]
            # whole warps of one bone pair: pick 32 vertices with distinct positions mod 32 where the group allows
            while len(vs) >= 32:
                by_res = {}
                for v in vs:
                    by_res.setdefault((v - v0) % 32, []).append(v)
                pick = [b.pop() for b in by_res.values()]
                pool = sorted((v for b in by_res.values() for v in b), key=lambda v: -len(by_res[(v - v0) % 32]))
                pick += pool[:32 - len(pick)]
                pick = pick[:32]
                warps.append(pick)
                chosen = set(pick)
                vs = [v for v in vs if v not in chosen]
            rest.append((key, vs))
        # remainders: consecutive in bone-pair order, so that a mixed warp holds as few pairs as possible
        tail = [v for key, vs in sorted(rest) for v in vs]
        for i in range(0, len(tail), 32):
            warps.append(tail[i:i + 32])
        assert len(warps) <= nwarp, (len(warps), nwarp)
        for w, vs in enumerate(warps):
            # lane order: vertices with a unique residue take the lane of that residue, the others the free lanes
            lanes = [-1] * 32
            left = []
            for v in vs:
                r = (v - v0) % 32
                if lanes[r] < 0:
                    lanes[r] = v
                else:
                    left.append(v)
            free = [i for i in range(32) if lanes[i] < 0]
            for v, i in zip(left, free):
                lanes[i] = v
            for lane, v in enumerate(lanes):
                if v >= 0:
                    pos[c * chunk + w * 32 + lane] = v - v0
    owner = np.where(pos >= 0, pos + (np.arange(nchunk * chunk) // chunk) * chunk, -1)
    return pos, owner


def forward_thread_assignment_spread(idx, val, nnz_row, V, chunk):
    """Which vertex each thread of the fused forward pass owns.

    A CTA covers ``chunk`` consecutive vertices; inside it the mapping thread -> vertex is free
    because the results are staged in shared memory before they leave.  Two things are wanted:
      * the threads of a warp should name as few distinct bones as possible: the bone transforms are
        read from shared memory with per-thread addresses, and lanes that share an address share
        the wavefront;
      * the staging stores ``stage[3*pos + c]`` must stay bank-conflict free: the 32 positions of a
        warp have to be distinct modulo 32.
    So every residue class ``pos % 32`` hands exactly one vertex to each warp; within a class the
    vertices are sorted by their bones and dealt to the warps in that order (vertices of the same
    bone quantile meet in the same warp).

    Returns ``fwd_pos[Vpad]`` (vertex index local to the chunk owned by thread t, -1 = none) and
    ``owner[Vpad]`` (the same as a global vertex index)."""
    nchunk = (V + chunk - 1) // chunk
    nwarp = chunk // 32
    K = idx.shape[1]
    pos = np.full(nchunk * chunk, -1, dtype=np.int64)

    def bones(v):
        b0 = int(idx[v, 0])
        b1 = int(idx[v, 1]) if K > 1 and nnz_row[v] > 1 else b0
        return b0, b1

    for c in range(nchunk):
        v0, v1 = c * chunk, min(V, (c + 1) * chunk)
        # grid[w][r] = vertex (or -1); start: every residue class sorted by bones and dealt to the warps in order
        grid = [[-1] * 32 for _ in range(nwarp)]
        for r in range(32):
            vs = sorted(range(v0 + r, v1, 32), key=lambda v: bones(v) + (v,))
            for w, v in enumerate(vs):
                grid[w][r] = v
        # local search: swap two warps' vertices of one residue class while that concentrates bones
        # (score = sum over warps, slots and bones of count^2; both slots weigh the same)
        cnt = [[dict(), dict()] for _ in range(nwarp)]
        for w in range(nwarp):
            for v in grid[w]:
                if v >= 0:
                    for s, b in enumerate(bones(v)):
                        cnt[w][s][b] = cnt[w][s].get(b, 0) + 1

        def gain_move(w, v, sign):
            # change of the score when vertex v is added to (sign=+1) / removed from (sign=-1) warp w
            g = 0
            if v < 0:
                return 0
            for s, b in enumerate(bones(v)):
                n = cnt[w][s].get(b, 0)
                g += (2 * n + 1) if sign > 0 else (1 - 2 * n)
            return g

        def apply(w, v, sign):
            if v < 0:
                return
            for s, b in enumerate(bones(v)):
                cnt[w][s][b] = cnt[w][s].get(b, 0) + sign

        for _sweep in range(12):
            improved = False
            for r in range(32):
                for wa in range(nwarp):
                    for wb in range(wa + 1, nwarp):
                        va, vb = grid[wa][r], grid[wb][r]
                        if va == vb or (va >= 0 and vb >= 0 and bones(va) == bones(vb)):
                            continue
                        g = gain_move(wa, va, -1) + gain_move(wb, vb, -1)
                        apply(wa, va, -1); apply(wb, vb, -1)
                        g += gain_move(wa, vb, +1) + gain_move(wb, va, +1)
                        if g > 0:
                            apply(wa, vb, +1); apply(wb, va, +1)
                            grid[wa][r], grid[wb][r] = vb, va
                            improved = True
                        else:
                            apply(wa, va, +1); apply(wb, vb, +1)
            if not improved:
                break
        for w in range(nwarp):
            for r in range(32):
                if grid[w][r] >= 0:
                    pos[c * chunk + w * 32 + r] = grid[w][r] - v0
    owner = np.where(pos >= 0, pos + (np.arange(nchunk * chunk) // chunk) * chunk, -1)
    return pos, owner


# ---------------------------------------------------------------------------------------------------
# fused backward pass, vertex-major tables (rigs with at most two skinning weights per vertex)
# ---------------------------------------------------------------------------------------------------
VM_CHUNK = int(_os.environ.get("SMIL_VM_CHUNK", "168"))   # upper bound of the vertices per tile (4 CTAs per SM on STICK)
VM_PHASES = int(_os.environ.get("SMIL_VM_PHASES", "1"))
VM_WARPS = int(_os.environ.get("SMIL_VM_WARPS", "4"))       # warps of a CTA that share one tile (1: every warp owns a small tile)   # sub-ranges of a tile with their own row buffer (double buffering)
VM_MODE_PACKED, VM_MODE_X, VM_MODE_Y = 0, 1, 2
VM_CUT_TOL = float(_os.environ.get("SMIL_VM_CUT_TOL", "0.06"))   # how far a warp cut may move from the equal-cost split
VM_SEG_COST = float(_os.environ.get("SMIL_VM_SEG_COST", "2.0"))    # a segment switch, in units of one vertex of the pass
VM_FLUSH_COST = float(_os.environ.get("SMIL_VM_FLUSH_COST", "6.0"))  # one flushed bone accumulator, same unit


def vm_rec_floats(nbt: int) -> int:
    """Floats per vertex record: (wX, wY, 3*local vertex, T0) | T1 T2 | S[nbt][3], padded to whole float4s."""
    return (6 + 3 * nbt + 3) // 4 * 4


def _vm_order_runs(combos, X=None, Y=None, lv0=0):
    """Order in which a stretch of vertices is processed by a warp that keeps TWO bones resident in registers (slots X
    and Y: rotation + 3x4 gradient accumulator each).  Vertices that share their bone set form a run; a run needs its
    bones resident; evicting a bone that is needed again later costs one more partial slot, so the runs are chained
    greedily along shared bones (the bone-pair graph of a skinned mesh is close to its skeleton tree, and a walk along
    it replaces one bone at a time).  ``X`` / ``Y``: the bones resident on entry (from the stretch before).
    Returns [(X, Y, run bones, [vertex indices lv0 + i])] and the resident pair on exit."""
    runs = {}
    for lv, c in enumerate(combos):
        runs.setdefault(tuple(sorted(c)), []).append(lv0 + lv)
    remaining = dict(runs)
    left = {}
    for c in remaining:
        for b in c:
            left[b] = left.get(b, 0) + 1
    ops = []
    while remaining:
        res = {b for b in (X, Y) if b is not None}
        best = None
        for c in sorted(remaining):
            need = [b for b in c if b not in res]
            evict = max(0, len(need) - (2 - len(res)))
            ev_cost = 0
            if evict:
                cand = sorted((b for b in res if b not in c), key=lambda b: (left.get(b, 0), b))
                ev_cost = sum(1 for b in cand[:evict] if left.get(b, 0) > 0)
            key = (ev_cost, -len(set(c) & res), evict, -len(remaining[c]), c)
            if best is None or key < best[0]:
                best = (key, c)
        c = best[1]
        for b in c:
            if b in (X, Y):
                continue
            if X is None:
                X = b
            elif Y is None:
                Y = b
            else:
                slot = min((s for s in (0, 1) if (X, Y)[s] not in c), key=lambda s: (left.get((X, Y)[s], 0), s))
                if slot == 0:
                    X = b
                else:
                    Y = b
        for b in c:
            left[b] -= 1
        ops.append((X, Y, c, remaining.pop(c)))
    return ops, (X, Y)


VM_CHUNK_MAX = int(_os.environ.get("SMIL_VM_CHUNK_MAX", "180"))       # upper bound of a tile after balancing (55.6 KB of shared memory)
VM_TILE_BALANCE = int(_os.environ.get("SMIL_VM_TILE_BALANCE", "4"))   # re-packing rounds that equalise the tiles' cost (0: equal vertex counts)


def _vm_tile_costs(t, nwarps):
    """Estimated cost of every tile of a vertex-major pack: its most loaded warp (vertices + weighted segment switches
    and flushes, the same model the warp cuts use) -- the warps of a CTA meet at a barrier after every sample group."""
    wsp, seg = t["vm_warp_seg_ptr"], t["vm_seg"].reshape(-1, 8)
    nph = t["vm_nphase"]
    out = []
    for ti in range(t["vm_ntile"]):
        worst = 0.0
        for w in range(nwarps):
            c = 0.0
            for ph in range(nph):
                s0, s1 = wsp[(ti * nph + ph) * nwarps + w], wsp[(ti * nph + ph) * nwarps + w + 1]
                sg = seg[s0:s1]
                c += float(sg[:, 3].sum()) + VM_SEG_COST * (s1 - s0) + VM_FLUSH_COST * float((sg[:, 4] >= 0).sum() + (sg[:, 5] >= 0).sum())
            worst = max(worst, c)
        out.append(worst)
    return np.asarray(out)


def pack_backward_vm_balanced(vt, S3, idx, val, nnz_row, J, nB, rounds=None):
    """:func:`pack_backward_vm` with the tile boundaries moved so that every tile has about the same estimated cost.
    All tiles of the pass are resident at once (one CTA per (tile, sample-group column)), so the pass ends with its most
    expensive tile: with equal vertex counts that tile was 14 % (STICK) / 33 % (mouse) over the mean.  The cost of a tile is
    not additive in its vertices (it depends on where the bone runs fall), so the boundaries are found by bisection on the
    cost bound T: left to right, every tile takes the most vertices (a multiple of 4, at most ``VM_CHUNK_MAX``) whose
    single-tile pack costs at most T; the smallest T for which the tiles cover the mesh wins.  Single-tile packs are
    memoised; the whole search takes a few seconds per rig and its result lives in the on-disk pack cache.
    Measured on B200 (B = 4096): skin_bwd_vm 92.0 -> 87.8 us on STICK (estimated worst tile 83 -> 74), 293.8 -> 260.2 us
    on the mouse rig (78 -> 65, and 713 -> 678 flush slots)."""
    rounds = VM_TILE_BALANCE if rounds is None else rounds
    V = vt.shape[0]
    base = pack_backward_vm(vt, S3, idx, val, nnz_row, J, nB)
    nt, nw = base["vm_ntile"], base["vm_warps"]
    if rounds <= 0 or nt < 2 or base["vm_nphase"] != 1:
        return base
    cmax = max(base["vm_chunk"], VM_CHUNK_MAX)   # light tiles may grow up to here (shared-memory bound: 4 CTAs per SM)
    memo = {}

    def tile_cost(v0, v1):
        if (v0, v1) not in memo:
            n = v1 - v0
            t1 = pack_backward_vm(vt[v0:v1], S3[:, v0:v1], idx[v0:v1], val[v0:v1], nnz_row[v0:v1], J, nB,
                                  chunk=(n + 3) // 4 * 4, tile_starts=[0, n])
            memo[(v0, v1)] = float(_vm_tile_costs(t1, nw)[0])
        return memo[(v0, v1)]

    def partition(T):
        starts = [0]
        for _ in range(nt):
            v0 = starts[-1]
            if v0 >= V:
                break
            n = min(cmax, V - v0)
            n = n if v0 + n == V else n // 4 * 4
            while n > 8 and tile_cost(v0, v0 + n) > T:
                n = (n - 1) // 4 * 4
            starts.append(v0 + n)
        return starts if starts[-1] == V else None

    base_cost = _vm_tile_costs(base, nw)
    hi, lo = float(base_cost.max()), float(base_cost.mean()) * 0.9
    best = None
    for _ in range(max(rounds, 1) + 3):
        mid = 0.5 * (lo + hi)
        st = partition(mid)
        if st is None:
            lo = mid
        else:
            best, hi = st, mid
    if best is None:
        return base
    while len(best) - 1 < nt:          # fewer tiles than CTAs columns: split the largest (keeps the grid shape)
        sizes = np.diff(best)
        i = int(np.argmax(sizes))
        best = best[:i + 1] + [best[i] + int(sizes[i]) // 8 * 4] + best[i + 1:]
    out = pack_backward_vm(vt, S3, idx, val, nnz_row, J, nB, chunk=(int(np.diff(best).max()) + 3) // 4 * 4, tile_starts=best)
    return out if _vm_tile_costs(out, nw).max() < base_cost.max() else base


def pack_backward_vm(vt, S3, idx, val, nnz_row, J, nB, chunk=None, nwarps=None, nphase=None, tile_starts=None):
    """Vertex-major tables of the fused backward skin pass (``skin_bwd_vm_kernel``).

    A tile is ``chunk`` consecutive vertices cut into ``nphase`` PHASES (sub-ranges that start on a multiple of four
    vertices).  The ``g_verts`` rows of one phase arrive by one TMA bulk copy per sample into that phase's own buffer,
    so while the CTA works on phase p the copy of phase p+1 (or of phase 0 of the next sample group) is in flight:
    the staging is double-buffered without a second tile-sized buffer.  Inside a phase the vertices are ordered by
    :func:`_vm_order_runs` and cut into ``nwarps`` shares; a warp walks its share of phase 0, then of phase 1, ... and its
    two resident bones survive the phase switch.  Every share is a list of SEGMENTS -- stretches with the same two
    resident bones.  A segment descriptor is 8 ints:
    ``(load X: bone or -1, load Y, first record, records | flush X: slot or -1, flush Y, mode, the warp's next
    segment or -1)``; a flush
    writes the slot's 3x4 partial of dL/dA to its own global slot (no atomics; slots are numbered bone-major so that
    the pose backward sums ``sb_ptr[k] .. sb_ptr[k+1]``).  Per vertex the shape blend, ``R^T g`` and the dL/dbeta
    contraction are evaluated ONCE (the item-major tables of :func:`pack_backward_items` evaluate them per bone)."""
    V = vt.shape[0]
    nphase = VM_PHASES if nphase is None else nphase
    nwarps = VM_WARPS if nwarps is None else nwarps
    chunk = backward_chunk(V, VM_CHUNK) if chunk is None else chunk
    q = 4 * nphase
    chunk = max(q, (chunk + q - 1) // q * q)   # phases start on 16-byte boundaries of a row when 3V % 4 == 0
    pw = chunk // nphase                       # vertices per phase
    ntile = (V + chunk - 1) // chunk
    if tile_starts is not None:                # explicit tile boundaries (pack_backward_vm_balanced); chunk = upper bound
        if nphase != 1:
            raise ValueError("explicit tile boundaries need nphase = 1")
        tile_starts = [int(x) for x in tile_starts]
        ntile = len(tile_starts) - 1
        if tile_starts[0] != 0 or tile_starts[-1] != V or any(b <= a or b - a > chunk for a, b in zip(tile_starts, tile_starts[1:])):
            raise ValueError("tile boundaries must cover [0, V) with 1..chunk vertices per tile")
    nbt = max(nbt_of(nB), 1)
    recf = vm_rec_floats(nbt)
    Sv = S3.transpose(1, 0, 2)  # (V, nB, 3)
    recs, segs = [], []
    tile_rec_ptr, warp_seg_ptr = [0], [0]        # warp_seg_ptr: [(tile * nphase + phase) * nwarps + warp]
    flush_bone = []          # bone of every flush, in creation order; slot ids are assigned bone-major afterwards
    max_tile_recs = max_tile_segs = 0
    for ti in range(ntile):
        v0, v1 = (ti * chunk, min(V, (ti + 1) * chunk)) if tile_starts is None else (tile_starts[ti], tile_starts[ti + 1])
        wts = [{int(idx[v, s]): float(val[v, s]) for s in range(int(nnz_row[v]))} for v in range(v0, v1)]
        tile_first_seg = len(segs)
        cur = [[None, None] for _ in range(nwarps)]       # bones in the warps' slots X / Y
        dirty = [[False, False] for _ in range(nwarps)]   # ... that hold an unflushed accumulator
        last_seg = [-1] * nwarps                          # the warp's latest segment (flushes are attached to it)
        X = Y = None
        for ph in range(nphase):
            p0, p1 = min(v1 - v0, ph * pw), min(v1 - v0, (ph + 1) * pw)
            combos = [tuple(int(idx[v0 + lv, s]) for s in range(max(int(nnz_row[v0 + lv]), 1))) for lv in range(p0, p1)]
            ops, (X, Y) = _vm_order_runs(combos, X, Y, lv0=p0)
            flat = [(lv, x, y) for x, y, _c, lvs in ops for lv in lvs]
            n = len(flat)
            # cut points between the shares: near the equal split, but preferably where the resident bones change
            # anyway (a cut inside a stretch with the same two bones costs two extra partial slots, at a boundary where
            # one bone changes one, where both change none)
            # The split is equal in estimated COST, not in vertices: a change of the resident bones costs a segment
            # switch plus one flush per replaced bone (VM_SEG_COST / VM_FLUSH_COST, in units of one vertex), and the
            # warps of a CTA meet at a barrier after every sample group -- with equal vertex counts the warp that
            # gets the many short runs at the end of a tile was 13 % (up to 22 %) over the mean on STICK.  Measured on
            # B200 (skin_bwd_vm, B = 4096): equal vertex counts 98.7 us, (seg, flush) = (1.1, 1.6) 96.5, (1.6, 2.4) 92.4,
            # (2.2, 3.2) 91.0, (2, 6) 89.2, (4, 5) 89.2, (6, 8) 89.3, (8, 10) 93.4; mouse 324 -> 293 us.  A flush weighs
            # far more than its ~90 instructions: the vector reductions queue behind each other in the LSU.
            cost = [0.0] * (n + 1)   # cost[pos]: estimated cost of flat[0:pos]
            for pos in range(1, n + 1):
                ch = 0
                if pos < n:
                    ch = int(flat[pos - 1][1] != flat[pos][1]) + int(flat[pos - 1][2] != flat[pos][2])
                cost[pos] = cost[pos - 1] + 1.0 + ((VM_SEG_COST + VM_FLUSH_COST * ch) if ch else 0.0)
            total = cost[n]
            cuts = [0]
            tol = max(1.0, VM_CUT_TOL * total / nwarps)
            for w in range(1, nwarps):
                ideal = total * w / nwarps
                best = None
                for pos in range(cuts[-1] + 1, n - (nwarps - w) + 1):
                    if abs(cost[pos] - ideal) > tol:
                        continue
                    same = int(flat[pos - 1][1] == flat[pos][1]) + int(flat[pos - 1][2] == flat[pos][2])
                    key = (same, abs(cost[pos] - ideal))
                    if best is None or key < best[0]:
                        best = (key, pos)
                if best is None:   # nothing inside the tolerance: the position closest to the ideal cost
                    cand = range(min(cuts[-1] + 1, n), max(n - (nwarps - w), cuts[-1]) + 1)
                    best = (None, min(cand, key=lambda pos: abs(cost[pos] - ideal))) if len(cand) else (None, min(cuts[-1], n))
                cuts.append(best[1])
            cuts.append(n)
            shares = [flat[cuts[k]:cuts[k + 1]] for k in range(nwarps)]
            # which warp takes which share is free: give every share to the warp that already holds its bones
            def used_bones(share):
                out = set()
                for lv, x, y in share:
                    out |= {b for b in (x, y) if b is not None and wts[lv].get(b, 0.0) != 0.0}
                return out
            ub = [used_bones(sh) for sh in shares]
            import itertools
            best_perm, best_score = tuple(range(nwarps)), -1
            for perm in itertools.permutations(range(nwarps)):
                score = sum(len(ub[perm[w]] & {b for b, d in zip(cur[w], dirty[w]) if d and b is not None}) for w in range(nwarps))
                if score > best_score:
                    best_perm, best_score = perm, score
            for w in range(nwarps):
                part = shares[best_perm[w]]
                curX, curY = cur[w]
                i = 0
                while i < len(part):
                    Xs, Ys = part[i][1], part[i][2]
                    j = i
                    while j < len(part) and part[j][1] == Xs and part[j][2] == Ys:
                        j += 1
                    # the planner's slot labels are per sequence; a warp may hold the bones in the other slots already
                    bones = [b for b in (Xs, Ys) if b is not None]
                    if curY in bones and curX not in bones and (Xs == curY or Ys == curY):
                        if Ys != curY:
                            Xs, Ys = Ys, Xs
                    elif curX in bones and Xs != curX:
                        Xs, Ys = Ys, Xs
                    seg_first = len(recs) - tile_rec_ptr[-1]
                    anyX = anyY = False
                    for lv, _x, _y in part[i:j]:
                        wX = wts[lv].get(Xs, 0.0) if Xs is not None else 0.0
                        wY = wts[lv].get(Ys, 0.0) if Ys is not None else 0.0
                        anyX |= wX != 0.0
                        anyY |= wY != 0.0
                        r = np.zeros(recf, dtype=np.float32)
                        r[0], r[1] = wX, wY
                        r[2] = np.int32(3 * (lv - ph * pw)).view(np.float32)    # float offset inside the PHASE's row
                        v = v0 + lv
                        r[3:6] = vt[v]
                        r[6:6 + 3 * nB] = Sv[v].reshape(-1)
                        recs.append(r)
                    # a slot whose bone changes is flushed by the warp's PREVIOUS segment (patched below) and loaded here
                    loadX = Xs if (Xs is not None and Xs != curX and anyX) else -1
                    loadY = Ys if (Ys is not None and Ys != curY and anyY) else -1
                    if loadX >= 0 and dirty[w][0]:
                        segs[last_seg[w]][4] = len(flush_bone); flush_bone.append(curX); dirty[w][0] = False
                    if loadY >= 0 and dirty[w][1]:
                        segs[last_seg[w]][5] = len(flush_bone); flush_bone.append(curY); dirty[w][1] = False
                    if loadX >= 0:
                        curX = Xs
                    if loadY >= 0:
                        curY = Ys
                    if anyX and Xs != curX:
                        raise AssertionError("vertex-major packer: slot X holds a stale bone")
                    if anyY and Ys != curY:
                        raise AssertionError("vertex-major packer: slot Y holds a stale bone")
                    dirty[w][0] |= anyX
                    dirty[w][1] |= anyY
                    mode = VM_MODE_X if not anyY else VM_MODE_PACKED   # a Y-only stretch (rare): packed path with w_X = 0
                    if last_seg[w] >= 0:
                        segs[last_seg[w]][7] = len(segs) - tile_first_seg    # link: the warp's next segment (tile-local)
                    last_seg[w] = len(segs)
                    segs.append([loadX, loadY, seg_first, j - i, -1, -1, mode, -1])
                    i = j
                cur[w] = [curX, curY]
                if ph == nphase - 1:   # end of the warp's work on this tile: what is still accumulating goes out with its
                    if dirty[w][0]:    # latest segment (which may sit in an earlier phase when the last share is empty)
                        segs[last_seg[w]][4] = len(flush_bone); flush_bone.append(curX)
                    if dirty[w][1]:
                        segs[last_seg[w]][5] = len(flush_bone); flush_bone.append(curY)
                warp_seg_ptr.append(len(segs))
        tile_rec_ptr.append(len(recs))
        max_tile_recs = max(max_tile_recs, tile_rec_ptr[-1] - tile_rec_ptr[-2])
        max_tile_segs = max(max_tile_segs, len(segs) - tile_first_seg)
    fb = np.asarray(flush_bone, dtype=np.int64)
    order = np.argsort(fb, kind="stable")
    slot_of = np.empty(len(fb), dtype=np.int64)
    slot_of[order] = np.arange(len(fb))
    seg_arr = np.asarray(segs, dtype=np.int64).reshape(-1, 8)
    for col in (4, 5):
        m = seg_arr[:, col] >= 0
        seg_arr[m, col] = slot_of[seg_arr[m, col]]
    sb_ptr = np.concatenate([[0], np.cumsum(np.bincount(fb, minlength=J))])
    return dict(
        vm_rec=_f32(np.stack(recs)) if recs else np.zeros((0, recf), np.float32), vm_recf=recf, vm_chunk=chunk,
        vm_nphase=nphase, vm_warps=nwarps, vm_ntile=ntile, vm_seg=_i32(seg_arr), vm_nseg=int(seg_arr.shape[0]),
        vm_tile_rec_ptr=_i32(tile_rec_ptr), vm_warp_seg_ptr=_i32(warp_seg_ptr), vm_sb_ptr=_i32(sb_ptr),
        vm_nslots=int(len(fb)), vm_slot_bone=_i32(fb[order]), vm_max_tile_recs=int(max_tile_recs),
        vm_max_tile_segs=int(max_tile_segs), vm_ok=True,
    )


def _interleave_for_streams(items):
    """Order of a unit's items ``(local vertex, weight, first)`` for the two item-pair streams of the fused backward.

    Items 4t, 4t+1 form the pair stream 0 handles in step t, items 4t+2, 4t+3 the pair of stream 1, and the two streams
    read their ``g_verts`` entries with the same instruction: lane (sample g, stream s) reads ``row_g[3 v_s + c]`` with a
    row stride of 4 (mod 32) floats, so the 16 samples of one stream already share banks two by two, and the two
    streams fall onto the SAME banks (4-way conflict) exactly when ``v_0 = v_1 (mod 4)``.  Items are therefore dealt
    so that the vertices at offsets 0 / 2 and 1 / 3 of a group of four differ mod 4 whenever the unit allows.
    Any order is valid (the kernel only sums over a unit's items)."""
    if _os.environ.get("SMIL_BWD_INTERLEAVE", "1") == "0" or len(items) < 4:
        return items
    buckets = [[], [], [], []]
    for it in items:
        buckets[it[0] & 3].append(it)
    for b_ in buckets:
        b_.reverse()   # pop() hands the items out in vertex order

    def take(avoid):
        cand = [r for r in range(4) if buckets[r] and r != avoid] or [r for r in range(4) if buckets[r]]
        r = max(cand, key=lambda q: (len(buckets[q]), -q))
        return buckets[r].pop()

    out = []
    left = len(items)
    while left:
        grp = [None] * 4
        for slot, partner in ((0, None), (2, 0), (1, None), (3, 1)):
            if not left:
                break
            avoid = grp[partner][0] & 3 if partner is not None and grp[partner] is not None else -1
            grp[slot] = take(avoid)
            left -= 1
        # a short last group keeps the pair layout: offsets (0, 1) are stream 0's pair, (2, 3) stream 1's
        got = [g for g in grp if g is not None]
        out.extend(grp if len(got) == 4 else got)
    return out


def pack_backward_items(vt, S3, idx, val, nnz_row, J, nB, chunk=BWD_CHUNK, nwarps=BWD_WARPS):
    """Items ``(vertex, bone, weight)`` grouped per vertex tile and, inside a tile, per bone.

    A tile is ``chunk`` consecutive vertices (the unit that is staged through shared memory by
    one bulk copy per sample).  Inside a tile, every bone with at least one item forms a *unit*;
    a unit is processed by one warp and produces one 3x4 partial of g_A written to its own
    *slot* (no atomics, deterministic).  Slots are numbered bone-major (``unit_slot``), so the
    slots of bone k are the contiguous range ``sb_ptr[k] .. sb_ptr[k+1]``."""
    V = vt.shape[0]
    ntile = (V + chunk - 1) // chunk
    # floats per vertex record {T, S_0..}: sized (and zero padded) for the kernel instance that
    # serves nB, so that instance never reads past a record
    rec = 4 * ((3 + 3 * max(nbt_of(nB), nB) + 3) // 4)
    vrec = np.zeros((ntile * chunk, rec), dtype=np.float32)
    vrec[:V, :3] = vt
    vrec[:V, 3:3 + 3 * nB] = S3.transpose(1, 0, 2).reshape(V, 3 * nB)   # [m][c]
    unit_bone, unit_ptr, unit_tile = [], [0], []
    it_v, it_w, it_first = [], [], []
    tile_unit_ptr = [0]
    warp_unit_ptr, warp_unit = [], []
    for ti in range(ntile):
        v0, v1 = ti * chunk, min(V, (ti + 1) * chunk)
        per_bone = {}
        for v in range(v0, v1):
            for s in range(int(nnz_row[v])):
                per_bone.setdefault(int(idx[v, s]), []).append((v - v0, float(val[v, s]), 1.0 if s == 0 else 0.0))
        first_unit = len(unit_bone)
        # a bone that owns more than one warp's share of the tile's items cannot be balanced as a single unit:
        # it is cut into equal parts, each a unit with its own slot (the slots of a bone stay contiguous)
        share = max(2, -(-sum(len(v) for v in per_bone.values()) // nwarps))
        for b in sorted(per_bone):
            items = per_bone[b]
            parts = max(1, -(-len(items) // share))
            size = -(-len(items) // parts)
            size += size & 1          # whole item pairs per part (only the last part may be odd)
            for p0 in range(0, len(items), size):
                unit_bone.append(b)
                unit_tile.append(ti)
                for (lv, w, f) in _interleave_for_streams(items[p0:p0 + size]):
                    it_v.append(lv)
                    it_w.append(w)
                    it_first.append(f)
                unit_ptr.append(len(it_v))
        tile_unit_ptr.append(len(unit_bone))
        # longest-processing-time assignment of the tile's units to warps
        units = list(range(first_unit, len(unit_bone)))
        units.sort(key=lambda u: -(unit_ptr[u + 1] - unit_ptr[u]))
        load = [0] * nwarps
        lists = [[] for _ in range(nwarps)]
        for u in units:
            w = int(np.argmin(load))
            lists[w].append(u)
            load[w] += unit_ptr[u + 1] - unit_ptr[u] + 8
        for w in range(nwarps):
            warp_unit_ptr.append(len(warp_unit))
            warp_unit.extend(lists[w])
    warp_unit_ptr.append(len(warp_unit))
    nunits = len(unit_bone)
    tup = np.asarray(tile_unit_ptr, dtype=np.int64)
    uptr = np.asarray(unit_ptr, dtype=np.int64)
    max_tile_items = int((uptr[tup[1:]] - uptr[tup[:-1]]).max()) if ntile else 0
    ub = np.asarray(unit_bone, dtype=np.int64)
    order = np.argsort(ub, kind="stable")
    unit_slot = np.empty(nunits, dtype=np.int64)   # slots are numbered bone-major: bone k owns
    unit_slot[order] = np.arange(nunits)           # slots sb_ptr[k] .. sb_ptr[k+1]
    sb_ptr = np.concatenate([[0], np.cumsum(np.bincount(ub, minlength=J))])
    # ---- item-PAIR records for the f32x2 kernel: two consecutive items of a unit share one record, so that
    # every constant arrives as an (item i, item i+1) register pair.  Units with an odd item count get a
    # zero-weight twin of their last item.  Record = T (3 pairs) | S (3*NBT pairs) | w | first | 3*lv (int bits) | pad.
    nbt = max(nbt_of(nB), 1) if nbt_of(nB) > 0 else nB
    rec2 = (6 + 6 * nbt + 8 + 3) // 4 * 4
    recs, unit_pptr = [], [0]
    Sv = S3.transpose(1, 0, 2)  # (V, nB, 3)
    for u in range(nunits):
        base = unit_tile[u] * chunk
        its = list(range(unit_ptr[u], unit_ptr[u + 1]))
        if len(its) % 2:
            its.append(-its[-1] - 1)  # zero-weight twin
        for a_, b_ in zip(its[0::2], its[1::2]):
            r = np.zeros(rec2, dtype=np.float32)
            for h, it in enumerate((a_, b_)):
                twin = it < 0
                it = -it - 1 if twin else it
                lv = it_v[it]
                v = base + lv
                r[0 + h], r[2 + h], r[4 + h] = vt[v]
                for m in range(nB):
                    for c in range(3):
                        r[6 + (m * 3 + c) * 2 + h] = Sv[v, m, c]
                o = 6 + 6 * nbt
                r[o + h] = 0.0 if twin else it_w[it]
                r[o + 2 + h] = 0.0 if twin else it_first[it]
                r[o + 4 + h] = np.int32(3 * lv).view(np.float32)
            recs.append(r)
        unit_pptr.append(len(recs))
    rec2_arr = np.stack(recs) if recs else np.zeros((0, rec2), np.float32)
    upp = np.asarray(unit_pptr, dtype=np.int64)
    max_tile_pairs = int((upp[tup[1:]] - upp[tup[:-1]]).max()) if ntile else 0
    return dict(
        rec2=_f32(rec2_arr), bwd_rec2=rec2, unit_pptr=_i32(unit_pptr), npair_items=len(recs),
        bwd_max_tile_pairs=max_tile_pairs,
        bwd_chunk=chunk, bwd_ntile=ntile, bwd_warps=nwarps, bwd_rec=rec, vrec=_f32(vrec),
        unit_bone=_i32(unit_bone), unit_ptr=_i32(unit_ptr), unit_tile=_i32(unit_tile),
        tile_unit_ptr=_i32(tile_unit_ptr), warp_unit_ptr=_i32(warp_unit_ptr), warp_unit=_i32(warp_unit),
        it_v=_i32(it_v), it_w=_f32(it_w), it_first=_f32(it_first), nunits=nunits, nitems=len(it_v),
        sb_ptr=_i32(sb_ptr), unit_slot=_i32(unit_slot), bwd_max_tile_items=max_tile_items,
    )
