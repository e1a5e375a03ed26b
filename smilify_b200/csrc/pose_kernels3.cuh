// Rodrigues + kinematic chain + closed-form joints, LANE = SAMPLE: a CTA owns 32 consecutive samples, its warps split
// the kinematic tree into CHAINS (root-to-leaf paths) and walk them with the running global transform in registers.
//
// Replaces the same reference code as pose_kernels.cuh (batch_lbs.py:31-50, :75-197; smal_torch.py:257-264, :343-351)
// for the headline configuration: axis-angle pose, no per-joint log-scales, no betas_trans, rest joints in closed form
// (JT + beta JS) or static.  Everything else keeps the warp-per-sample kernels of pose_kernels.cuh.
//
// Why: with a warp per sample (pose_kernels.cuh) the 55 joints of the STICK rig fill 55 of 64 lane slots in the
// per-joint phases and 6-12 of 32 lanes in the level walk, every table read is a per-warp L1/L2 round trip, and the
// kernels execute 3150 (forward) / 5420 (backward) warp instructions PER SAMPLE at 47 % issue utilisation
// (profiles/r02_ncu_step_summary.md).  With lane = sample no lane idles, the tables are warp-uniform (staged once per
// CTA in shared memory, read as broadcasts) and there are no shuffles.  A first lane = sample version that kept the
// level-synchronous walk through shared memory needed as many instructions as the old kernel (10.7 M, 32 us:
// gpurun_out/r02_pose_ls_a); walking whole chains in registers removes the per-level load / store / barrier traffic:
// a joint costs one Rodrigues, one 3x4 composition and its output stores.
//
// The chain schedule (heavy-path decomposition, rounds, warp assignment) is built on the host at model creation
// (smil_b200.cu) and staged with the other tables.  Per-joint state that other warps or the coalesced global copies
// need lives in [row][32 samples] tiles with a row stride of 33 floats: conflict-free both for lane = sample accesses
// (immediate offsets from a per-joint base) and for the transposed accesses of the copies.
#pragma once
#include "pose_kernels.cuh"

namespace smil {

constexpr int LS_NW = 16;              // warps per CTA (the STICK tree has 16 chains)
constexpr int LS_THREADS = 32 * LS_NW;
constexpr int LS_RS = 33;              // row stride of the [row][sample] tiles

// float offsets of the shared-memory regions (the same code sizes the launch on the host)
struct PoseLs {
  int ax, nj, jrs, st, rstg, tab, flags, total;
  __host__ void init_fwd(const DevModel& m) {
    const int J = m.J;
    int o = 0;
    ax = o; o += 12 * J * LS_RS;          // before the walk: local rotations R_j; after: A = [GR | Gt - GR Jrest]
    nj = o; o += 3 * J * LS_RS;           // posed joints (new_J = Gt)
    jrs = o; o += 3 * J * LS_RS;          // rest joints
    st = o; o += 3 * J * LS_RS;           // theta staging; then rowsum*trans and the closed-form joints
    rstg = o; o += LS_NW * 9 * LS_RS;     // per-warp staging of one joint's local rotation for the Rs rows
    o = (o + 3) & ~3;
    tab = o; o += m.ls_words;             // the model's table blob (DevModel::ls_blob)
    total = o;
  }
  // backward: nj -> theta rows (later g_theta), st -> g_joints rows, rstg -> per-chain deposits [nchain][15 rows]
  __host__ void init_bwd(const DevModel& m) {
    const int J = m.J;
    int o = 0;
    ax = o; o += 12 * J * LS_RS;          // A of the forward (its 3x3 part is the global rotation); at the end: reduction scratch
    jrs = o; o += 3 * J * LS_RS;          // rest joints
    nj = o; o += 3 * J * LS_RS;           // theta; joint by joint replaced by g_theta
    st = o; o += 3 * J * LS_RS;           // g_joints
    rstg = o; o += (m.ls_nchain > 0 ? m.ls_nchain : 1) * 15 * LS_RS;
    o = (o + 3) & ~3;
    tab = o; o += m.ls_words;
    flags = o; o += ((m.ls_nchain > 0 ? m.ls_nchain : 1) + 3) & ~3;   // one "deposit written" flag per chain
    total = o;
  }
};

// Staging of the CTA's inputs.  Loads and stores are separate calls so that a kernel can issue EVERY global load of its
// prologue before the first shared-memory store: the whole prologue then costs one memory round trip instead of one
// per array (the stores of one array would otherwise sit between the loads of the next).
struct LsBlobRegs { float4 v[4]; };
__device__ __forceinline__ void ls_blob_load(LsBlobRegs& r, const float* src, int words, int i0, int tid) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
  const int n4 = words >> 2;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int i = i0 + u * LS_THREADS + tid;
    r.v[u] = (i < n4) ? s4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
__device__ __forceinline__ void ls_blob_store(const LsBlobRegs& r, float* dst, int words, int i0, int tid) {
  float4* d4 = reinterpret_cast<float4*>(dst);
  const int n4 = words >> 2;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int i = i0 + u * LS_THREADS + tid;
    if (i < n4) d4[i] = r.v[u];
  }
}
// rows [e][sample] of a [B][n] global array for the CTA's 32 samples (2 per warp), 192 elements per call
struct LsRowRegs { float v[2][6]; };
__device__ __forceinline__ void ls_rows_load(LsRowRegs& r, const float* src, long long stride, int n, int e0, int b0, int B, int w, int l) {
  static_assert(LS_NW == 16, "two sample rows per warp");
#pragma unroll
  for (int ss = 0; ss < 2; ++ss) {
    const float* row = src + (long long)min(b0 + w + LS_NW * ss, B - 1) * stride;
#pragma unroll
    for (int t = 0; t < 6; ++t) {
      const int e = e0 + l + 32 * t;
      r.v[ss][t] = (e < n) ? row[e] : 0.f;
    }
  }
}
__device__ __forceinline__ void ls_rows_store(const LsRowRegs& r, float* dst, int n, int e0, int w, int l) {
#pragma unroll
  for (int ss = 0; ss < 2; ++ss) {
#pragma unroll
    for (int t = 0; t < 6; ++t) {
      const int e = e0 + l + 32 * t;
      if (e < n) dst[e * LS_RS + w + LS_NW * ss] = r.v[ss][t];
    }
  }
}
// whatever the first call of each did not cover (J > 64, blobs above 32 KB): plain load-then-store chunks
__device__ __forceinline__ void ls_blob_rest(float* dst, const float* src, int words, int tid) {
  for (int i0 = 4 * LS_THREADS; i0 < (words >> 2); i0 += 4 * LS_THREADS) {
    LsBlobRegs r;
    ls_blob_load(r, src, words, i0, tid);
    ls_blob_store(r, dst, words, i0, tid);
  }
}
__device__ __forceinline__ void ls_rows_rest(float* dst, const float* src, long long stride, int n, int b0, int B, int w, int l) {
  for (int e0 = 192; e0 < n; e0 += 192) {
    LsRowRegs r;
    ls_rows_load(r, src, stride, n, e0, b0, B, w, l);
    ls_rows_store(r, dst, n, e0, w, l);
  }
}

template <int NBMAX>
__global__ void __launch_bounds__(LS_THREADS, 1) pose_fwd_ls_kernel(DevModel m, PoseArgs a, PoseLs L) {
  extern __shared__ __align__(16) float sm_ls[];
  const int tid = threadIdx.x, w = tid >> 5, l = tid & 31;
  const int J = m.J, nb = a.nb, J3 = 3 * J;
  const int b0 = blockIdx.x * 32;
  const int nlive = min(32, a.B - b0);
  const int bs = min(b0 + l, a.B - 1);   // the lane's sample; lanes past the batch recompute the last one and write nothing
  float* AX = sm_ls + L.ax;
  float* NJ = sm_ls + L.nj;
  float* JRS = sm_ls + L.jrs;
  float* ST = sm_ls + L.st;
  float* RSTG = sm_ls + L.rstg + w * 9 * LS_RS;
  float* TAB = sm_ls + L.tab;
  const int* iS = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_SCHED];
  const int* iChainPtr = iS + m.ls_off[0];
  const int* iChainJoint = iS + m.ls_off[1];
  const int* iChainParent = iS + m.ls_off[2];
  const int* iSchedPtr = iS + m.ls_off[3];
  const int* iSchedChain = iS + m.ls_off[4];
  const int* iPairJb = iS + m.ls_off[7];
  const int* iPjp = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_PJP];
  const int* iPjb = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_PJB];
  const float* tRS = TAB + m.ls_o[LS_O_RSUM];
  const float* tJST = TAB + m.ls_o[LS_O_JST];
  const float* tPREC = TAB + m.ls_o[LS_O_PREC];
  const bool pairs = a.joints && a.joints_mode == 1;
  const int jsw = m.ls_jsw, prf = m.ls_prf;

  // ---- tables and inputs -> shared memory (one memory round trip for everything) ------------------------------------
  float be[NBMAX];
#pragma unroll
  for (int k = 0; k < NBMAX; ++k) be[k] = (k < nb) ? a.beta[(long long)bs * a.beta_stride + k] : 0.f;
  float tr[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) tr[c] = a.trans ? a.trans[(long long)bs * a.trans_stride + c] : 0.f;
  {
    const float* blob = m.ls_blob[a.jrest_mode == 0 ? 0 : 1];
    LsRowRegs rt;
    LsBlobRegs rb;
    ls_rows_load(rt, a.theta, a.theta_stride, J3, 0, b0, a.B, w, l);
    ls_blob_load(rb, blob, m.ls_words, 0, tid);
    ls_rows_store(rt, ST, J3, 0, w, l);
    ls_blob_store(rb, TAB, m.ls_words, 0, tid);
    ls_rows_rest(ST, a.theta, a.theta_stride, J3, b0, a.B, w, l);
    ls_blob_rest(TAB, blob, m.ls_words, tid);
  }
  __syncthreads();

  // rest joint component e = 3j + c:  JT[e] + sum_k beta_k JS[k][e]  (row (JT | JS_0 .. | 0) of the re-laid table)
  auto rest_joint = [&](int e) {
    const float4* row = reinterpret_cast<const float4*>(tJST + e * jsw);
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < (NBMAX + 4) / 4; ++q) {
      if (4 * q < jsw) {
        const float4 t = row[q];
        const float tv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int k = 4 * q + u - 1;
          if (k < 0) v += tv[u];
          else if (k < NBMAX) v = fmaf(be[k], tv[u], v);
        }
      }
    }
    return v;
  };

  // ---- phase 1, warp per joint: local rotation and rest joint; Rs rows; the seed of the closed-form joint sum -----
  for (int i = w; i < J; i += LS_NW) {
    const float* tp = ST + 3 * i * LS_RS + l;
    const m3 R = rodrigues(mk3(tp[0], tp[LS_RS], tp[2 * LS_RS]));
    const f3 jr = mk3(rest_joint(3 * i), rest_joint(3 * i + 1), rest_joint(3 * i + 2));
    float* ax = AX + i * 12 * LS_RS + l;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) ax[(4 * r + c) * LS_RS] = R.a[3 * r + c];
    }
    float* jrp = JRS + 3 * i * LS_RS + l;
    jrp[0] = jr.x; jrp[LS_RS] = jr.y; jrp[2 * LS_RS] = jr.z;
    if (i == 0) {   // root: G_0 = (R_0 | J_0), A_0 = [R_0 | J_0 - R_0 J_0]
      const f3 rot = mul(R, jr);
      ax[3 * LS_RS] = jr.x - rot.x; ax[7 * LS_RS] = jr.y - rot.y; ax[11 * LS_RS] = jr.z - rot.z;
      float* njp = NJ + l;
      njp[0] = jr.x; njp[LS_RS] = jr.y; njp[2 * LS_RS] = jr.z;
    }
    if (pairs) {   // theta_i is consumed: its rows now seed the closed-form joint sum with rowsum_i * trans
      const float rs = tRS[i];
      float* sp = ST + 3 * i * LS_RS + l;
      sp[0] = rs * tr[0]; sp[LS_RS] = rs * tr[1]; sp[2 * LS_RS] = rs * tr[2];
    }
    if (a.Rs) {    // [B][J][9]: 36 contiguous bytes per sample; staged so that a store instruction covers ~3.5 samples
#pragma unroll
      for (int q = 0; q < 9; ++q) RSTG[q * LS_RS + l] = R.a[q];
      __syncwarp();
      float* out = a.Rs + (long long)b0 * 9 * J + 9 * i;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const int idx = l + 32 * t;
        const int s = (idx * 57) >> 9;   // idx / 9 for idx < 288
        const int q = idx - 9 * s;
        if (s < nlive) out[(long long)s * 9 * J + q] = RSTG[q * LS_RS + s];
      }
      __syncwarp();
    }
  }
  __syncthreads();

  // ---- phase 2: chains, round by round; the running global transform stays in registers.  (Per-joint "final" flags
  // instead of the CTA barrier between the rounds, as in the backward, measured SLOWER here: 23.2-24.3 vs 21.6-22.2 us --
  // the fence + flag store per joint costs more than the two short barriers.) ------------------------------------------
  for (int r = 0; r < m.ls_nround; ++r) {
    for (int ci = iSchedPtr[r * LS_NW + w]; ci < iSchedPtr[r * LS_NW + w + 1]; ++ci) {
      const int c = iSchedChain[ci];
      const int p = iChainParent[c];
      m3 GR;
      {
        const float* ax = AX + p * 12 * LS_RS + l;   // the parent is final: the root, or a joint of an earlier round
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
          for (int cc = 0; cc < 3; ++cc) GR.a[3 * rr + cc] = ax[(4 * rr + cc) * LS_RS];
        }
      }
      const float* njp = NJ + 3 * p * LS_RS + l;
      f3 Gt = mk3(njp[0], njp[LS_RS], njp[2 * LS_RS]);
      const float* jpp = JRS + 3 * p * LS_RS + l;
      f3 Jp = mk3(jpp[0], jpp[LS_RS], jpp[2 * LS_RS]);
      for (int t = iChainPtr[c]; t < iChainPtr[c + 1]; ++t) {
        const int i = iChainJoint[t];
        float* ax = AX + i * 12 * LS_RS + l;
        m3 R;
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
          for (int cc = 0; cc < 3; ++cc) R.a[3 * rr + cc] = ax[(4 * rr + cc) * LS_RS];
        }
        const float* jrp = JRS + 3 * i * LS_RS + l;
        const f3 jr = mk3(jrp[0], jrp[LS_RS], jrp[2 * LS_RS]);
        const f3 gt = mul(GR, mk3(jr.x - Jp.x, jr.y - Jp.y, jr.z - Jp.z));
        GR = matmul(GR, R);
        Gt = mk3(gt.x + Gt.x, gt.y + Gt.y, gt.z + Gt.z);
        Jp = jr;
        const f3 rot = mul(GR, jr);
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
          for (int cc = 0; cc < 3; ++cc) ax[(4 * rr + cc) * LS_RS] = GR.a[3 * rr + cc];
        }
        ax[3 * LS_RS] = Gt.x - rot.x; ax[7 * LS_RS] = Gt.y - rot.y; ax[11 * LS_RS] = Gt.z - rot.z;
        float* nji = NJ + 3 * i * LS_RS + l;
        nji[0] = Gt.x; nji[LS_RS] = Gt.y; nji[2 * LS_RS] = Gt.z;
      }
    }
    __syncthreads();
  }

  // ---- posed joints in closed form: joints_j = rowsum_j trans + sum_k A_k [c0 + cm beta ; c0.w] --------------------
  if (pairs) {
    for (int j = iPairJb[w]; j < iPairJb[w + 1]; ++j) {
      float* sp = ST + 3 * j * LS_RS + l;
      f3 acc = mk3(sp[0], sp[LS_RS], sp[2 * LS_RS]);
      for (int pi = iPjp[j]; pi < iPjp[j + 1]; ++pi) {
        const float* ax = AX + iPjb[pi] * 12 * LS_RS + l;
        const float4* rec = reinterpret_cast<const float4*>(tPREC + pi * prf);
        const float4 c0 = rec[0];
        float q[3] = {c0.x, c0.y, c0.z};
#pragma unroll
        for (int qi = 1; qi < (4 + 3 * NBMAX + 3) / 4; ++qi) {
          if (4 * qi < prf) {
            const float4 t = rec[qi];
            const float tv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = 4 * qi + u - 4;   // element of cm[nB][3]
              if (e < 3 * NBMAX) q[e % 3] = fmaf(be[e / 3], tv[u], q[e % 3]);
            }
          }
        }
        acc.x += fmaf(ax[0], q[0], fmaf(ax[LS_RS], q[1], fmaf(ax[2 * LS_RS], q[2], ax[3 * LS_RS] * c0.w)));
        acc.y += fmaf(ax[4 * LS_RS], q[0], fmaf(ax[5 * LS_RS], q[1], fmaf(ax[6 * LS_RS], q[2], ax[7 * LS_RS] * c0.w)));
        acc.z += fmaf(ax[8 * LS_RS], q[0], fmaf(ax[9 * LS_RS], q[1], fmaf(ax[10 * LS_RS], q[2], ax[11 * LS_RS] * c0.w)));
      }
      sp[0] = acc.x; sp[LS_RS] = acc.y; sp[2 * LS_RS] = acc.z;
    }
  }
  // ---- pair blocks for the skinning passes: A2 [J][PPB] interleaved over the pair | beta2 | trans2 -----------------
  {
    const int npair = (nlive + 1) >> 1;
    const int nb2 = (a.pp_blk - PPB * J - 8) >> 1;
    for (int pr = w; pr < npair; pr += LS_NW) {
      float* blk = a.pp + ((long long)(b0 >> 1) + pr) * a.pp_blk;
      float4* out = reinterpret_cast<float4*>(blk);
      const float* x0 = AX + 2 * pr;   // slot 2 pr + 1 holds a copy of the last sample when the batch is odd
      int j = (l * 171) >> 10, t = l - 6 * j;   // l / 6 for l < 32
      for (int i = l; i < 6 * J; i += 32) {
        const float* x = x0 + (j * 12 + 2 * t) * LS_RS;
        out[j * (PPB / 4) + t] = make_float4(x[0], x[1], x[LS_RS], x[LS_RS + 1]);
        t += 2; j += 5;
        if (t >= 6) { t -= 6; ++j; }
      }
      for (int i0 = 0; i0 < 2 * nb2; i0 += 32) {   // beta2: (k, h) -> beta_k of sample 2 pr + h, from the lanes' registers
        const int i = i0 + l, k = i >> 1, h = i & 1;
        float v = 0.f;
#pragma unroll
        for (int kk = 0; kk < NBMAX; ++kk) {
          const float bv = __shfl_sync(0xffffffffu, be[kk], 2 * pr + h);
          if (kk == k) v = bv;
        }
        if (i < 2 * nb2) blk[PPB * J + i] = v;
      }
      {
        const int c = l >> 1, h = l & 1;
        float v = 0.f;
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
          const float tv = __shfl_sync(0xffffffffu, tr[cc], 2 * pr + h);
          if (cc == c) v = tv;
        }
        if (l < 8) blk[PPB * J + 2 * nb2 + l] = v;
      }
    }
  }
  __syncthreads();   // closed-form joints complete
  for (int s = w; s < nlive; s += LS_NW) {
    const long long o = (long long)(b0 + s) * J3;
    for (int e = l; e < J3; e += 32) {
      const float nj = NJ[e * LS_RS + s];
      if (a.newJ) a.newJ[o + e] = nj;
      if (a.joints) a.joints[o + e] = pairs ? ST[e * LS_RS + s] : nj;
      if (a.jrest_out) a.jrest_out[o + e] = JRS[e * LS_RS + s];
    }
  }
}

// Backward of the same path (autograd of batch_lbs.py:75-197 + Rodrigues + the closed-form joints), lane = sample.
// Configuration of pose_bwd_lean_kernel (no log-scales) without betas_trans / g_Rs, dense dL/dA from the skinning pass.
// The chains are walked from their tails to their heads with the running (dL/dGR, dL/dGt, dL/dJrest-of-parent)
// contribution in registers; a chain hands its final contribution to the joint it hangs off through a deposit slot
// and a shared-memory flag (no CTA barrier between the rounds: a warp waits only for the chains that hang off the joint
// it is about to process), the root is processed last, by one warp, while the others run the per-joint epilogue.
// (Measured and dropped: a parallel pre-pass that folds the closed-form posed joints into dL/dA before the walk and hands
// the result over through the dense scratch buffer -- 35.5 vs 31.2 us: the walk is bound by the L2 latency of its
// per-joint dL/dA loads, which the inline pair work used to cover.)
// Per joint i with parent p (lean kernel, same algebra):
//   dL/dG_i = (dL/dA_i.R - dL/dA_i.t Jrest_i^T | dL/dA_i.t) + children,   jacc_i = -GR_i^T dL/dA_i.t
//   M_i = GR_p^T GR_i, tl_i = Jrest_i - Jrest_p;   dL/dM_i = GR_p^T dL/dGR_i -> Rodrigues VJP -> g_theta_i
//   to the parent: (dL/dGR_i M_i^T + dL/dGt_i tl_i^T | dL/dGt_i), dL/dJrest_p -= GR_p^T dL/dGt_i, dL/dJrest_i = jacc_i + GR_p^T dL/dGt_i
template <int NBMAX>
__global__ void __launch_bounds__(LS_THREADS, 1) pose_bwd_ls_kernel(DevModel m, PoseArgs a, PoseLs L, int pairs_mode) {
  extern __shared__ __align__(16) float sm_ls[];
  const int tid = threadIdx.x, w = tid >> 5, l = tid & 31;
  const int J = m.J, nb = a.nb, J3 = 3 * J;
  const int b0 = blockIdx.x * 32;
  const int nlive = min(32, a.B - b0);
  const int bs = min(b0 + l, a.B - 1);
  float* AX = sm_ls + L.ax;
  float* JRS = sm_ls + L.jrs;
  float* TH = sm_ls + L.nj;
  float* GJ = sm_ls + L.st;
  float* DEP = sm_ls + L.rstg;
  float* TAB = sm_ls + L.tab;
  const int* iS = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_SCHED];
  const int* iChainPtr = iS + m.ls_off[0];
  const int* iChainJoint = iS + m.ls_off[1];
  const int* iChainParent = iS + m.ls_off[2];
  const int* iSchedPtr = iS + m.ls_off[3];
  const int* iSchedChain = iS + m.ls_off[4];
  const int* iDepPtr = iS + m.ls_off[5];
  const int* iDepChain = iS + m.ls_off[6];
  volatile int* FLAG = reinterpret_cast<volatile int*>(sm_ls + L.flags);
  const int* iPbp = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_PBP];
  const int* iPbq = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_PBQ];
  const int* iPjj = reinterpret_cast<const int*>(TAB) + m.ls_o[LS_O_PJJ];
  const float* tRS = TAB + m.ls_o[LS_O_RSUM];
  const float* tJST = TAB + m.ls_o[LS_O_JST];
  const float* tPREC = TAB + m.ls_o[LS_O_PREC];
  const int jsw = m.ls_jsw, prf = m.ls_prf;
  const bool is_static = a.jrest_mode == 1;
  const bool gj_direct = a.joints_mode == 2 && a.g_joints;   // static rigs: joints = new_J

  // ---- inputs -> shared memory ------------------------------------------------------------------------------------------
  float be[NBMAX];
#pragma unroll
  for (int k = 0; k < NBMAX; ++k) be[k] = (k < nb) ? a.beta[(long long)bs * a.beta_stride + k] : 0.f;
  {
    static_assert(LS_NW == 16, "one sample pair per warp");
    const float* blob = m.ls_blob[is_static ? 1 : 0];
    // the forward's pair blocks: float4 (j, t) = (A_s0[2t], A_s1[2t], A_s0[2t+1], A_s1[2t+1]); one pair per warp
    const int npair_tot = (a.B + 1) >> 1;
    const float4* blk = reinterpret_cast<const float4*>(a.pp_saved + (long long)min((b0 >> 1) + w, npair_tot - 1) * a.pp_blk);
    float* x0 = AX + 2 * w;
    constexpr int PPC = 12;   // float4 per lane and chunk: 384 per warp covers 6 J for J <= 64 in one round trip
    auto pp_chunk = [&](int i0, bool load, float4 (&v)[PPC]) {
#pragma unroll
      for (int u = 0; u < PPC; ++u) {
        const int i = i0 + 32 * u + l;
        const int jj = (int)(((unsigned)i * 43691u) >> 18);   // i / 6 for i < 2^16
        const int tt = i - 6 * jj;
        if (load) {
          v[u] = (i < 6 * J) ? blk[jj * (PPB / 4) + tt] : make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (i < 6 * J) {
          float* x = x0 + (jj * 12 + 2 * tt) * LS_RS;
          x[0] = v[u].x; x[1] = v[u].y; x[LS_RS] = v[u].z; x[LS_RS + 1] = v[u].w;
        }
      }
    };
    for (int c = tid; c < m.ls_nchain; c += LS_THREADS) FLAG[c] = 0;
    LsRowRegs rt, rg, rj;
    LsBlobRegs rb;
    float4 vp[PPC];
    ls_rows_load(rt, a.theta, a.theta_stride, J3, 0, b0, a.B, w, l);
    if (a.g_joints) ls_rows_load(rg, a.g_joints, J3, J3, 0, b0, a.B, w, l);
    if (!is_static) ls_rows_load(rj, a.jrest_in, J3, J3, 0, b0, a.B, w, l);
    ls_blob_load(rb, blob, m.ls_words, 0, tid);
    pp_chunk(0, true, vp);
    ls_rows_store(rt, TH, J3, 0, w, l);
    if (a.g_joints) ls_rows_store(rg, GJ, J3, 0, w, l);
    if (!is_static) ls_rows_store(rj, JRS, J3, 0, w, l);
    ls_blob_store(rb, TAB, m.ls_words, 0, tid);
    pp_chunk(0, false, vp);
    ls_rows_rest(TH, a.theta, a.theta_stride, J3, b0, a.B, w, l);
    if (a.g_joints) ls_rows_rest(GJ, a.g_joints, J3, J3, b0, a.B, w, l);
    if (!is_static) ls_rows_rest(JRS, a.jrest_in, J3, J3, b0, a.B, w, l);
    ls_blob_rest(TAB, blob, m.ls_words, tid);
    for (int i0 = 32 * PPC; i0 < 6 * J; i0 += 32 * PPC) {
      pp_chunk(i0, true, vp);
      pp_chunk(i0, false, vp);
    }
  }
  __syncthreads();
  if (is_static) {   // rest joints of a static rig: the same for every sample
    for (int e = w; e < J3; e += LS_NW) JRS[e * LS_RS + l] = tJST[e * jsw];
    __syncthreads();
  }

  float gbeta[NBMAX];
#pragma unroll
  for (int k = 0; k < NBMAX; ++k) gbeta[k] = 0.f;
  float gtr[3] = {0.f, 0.f, 0.f};

  auto load_gA = [&](int i, float4 (&g)[3]) {
    const float4* p = reinterpret_cast<const float4*>(a.gA_dense) + ((long long)bs * J + i) * 3;
    g[0] = p[0]; g[1] = p[1]; g[2] = p[2];
  };
  // dL/dG_i of joint i from its own dL/dA_i (+ the closed-form posed joints) and the deposits of the chains hanging off
  // it; also folds rowsum_i g_joints_i into g_trans and GR_i^T g_joints . cm into g_beta.  Returns jacc in `ja`.
  auto own_grad = [&](int i, const float4 (&g)[3], const m3& GRi, f3 Jri, m3& gGR, f3& gGt, f3& ja, int dep0, int dep1, bool wait) {
    m3 gAR;
    gAR.a[0] = g[0].x; gAR.a[1] = g[0].y; gAR.a[2] = g[0].z;
    gAR.a[3] = g[1].x; gAR.a[4] = g[1].y; gAR.a[5] = g[1].z;
    gAR.a[6] = g[2].x; gAR.a[7] = g[2].y; gAR.a[8] = g[2].z;
    f3 gAt = mk3(g[0].w, g[1].w, g[2].w);
    if (pairs_mode) {
      const float* gi = GJ + 3 * i * LS_RS + l;
      const float rs = tRS[i];
      gtr[0] = fmaf(rs, gi[0], gtr[0]); gtr[1] = fmaf(rs, gi[LS_RS], gtr[1]); gtr[2] = fmaf(rs, gi[2 * LS_RS], gtr[2]);
      for (int t = iPbp[i]; t < iPbp[i + 1]; ++t) {
        const int pi = iPbq[t];
        const float* gp = GJ + 3 * iPjj[pi] * LS_RS + l;
        const f3 gj = mk3(gp[0], gp[LS_RS], gp[2 * LS_RS]);
        const f3 tg = mulT(GRi, gj);
        const float4* rec = reinterpret_cast<const float4*>(tPREC + pi * prf);
        const float4 c0 = rec[0];
        float q[3] = {c0.x, c0.y, c0.z};
#pragma unroll
        for (int qi = 1; qi < (4 + 3 * NBMAX + 3) / 4; ++qi) {
          if (4 * qi < prf) {
            const float4 tq = rec[qi];
            const float tv[4] = {tq.x, tq.y, tq.z, tq.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = 4 * qi + u - 4;   // element of cm[nB][3]
              if (e < 3 * NBMAX) {
                q[e % 3] = fmaf(be[e / 3], tv[u], q[e % 3]);
                gbeta[e / 3] = fmaf(e % 3 == 0 ? tg.x : (e % 3 == 1 ? tg.y : tg.z), tv[u], gbeta[e / 3]);
              }
            }
          }
        }
        gAR.a[0] = fmaf(gj.x, q[0], gAR.a[0]); gAR.a[1] = fmaf(gj.x, q[1], gAR.a[1]); gAR.a[2] = fmaf(gj.x, q[2], gAR.a[2]);
        gAR.a[3] = fmaf(gj.y, q[0], gAR.a[3]); gAR.a[4] = fmaf(gj.y, q[1], gAR.a[4]); gAR.a[5] = fmaf(gj.y, q[2], gAR.a[5]);
        gAR.a[6] = fmaf(gj.z, q[0], gAR.a[6]); gAR.a[7] = fmaf(gj.z, q[1], gAR.a[7]); gAR.a[8] = fmaf(gj.z, q[2], gAR.a[8]);
        gAt.x = fmaf(gj.x, c0.w, gAt.x); gAt.y = fmaf(gj.y, c0.w, gAt.y); gAt.z = fmaf(gj.z, c0.w, gAt.z);
      }
    }
    gGR.a[0] = gAR.a[0] - gAt.x * Jri.x; gGR.a[1] = gAR.a[1] - gAt.x * Jri.y; gGR.a[2] = gAR.a[2] - gAt.x * Jri.z;
    gGR.a[3] = gAR.a[3] - gAt.y * Jri.x; gGR.a[4] = gAR.a[4] - gAt.y * Jri.y; gGR.a[5] = gAR.a[5] - gAt.y * Jri.z;
    gGR.a[6] = gAR.a[6] - gAt.z * Jri.x; gGR.a[7] = gAR.a[7] - gAt.z * Jri.y; gGR.a[8] = gAR.a[8] - gAt.z * Jri.z;
    gGt = gAt;
    const f3 r = mulT(GRi, gAt);
    ja = mk3(-r.x, -r.y, -r.z);
    if (gj_direct) {
      const float* gi = GJ + 3 * i * LS_RS + l;
      gGt.x += gi[0]; gGt.y += gi[LS_RS]; gGt.z += gi[2 * LS_RS];
    }
    for (int t = dep0; t < dep1; ++t) {   // chains that hang off joint i (finished in earlier rounds)
      const int dc = iDepChain[t];
      if (wait) {
        while (FLAG[dc] == 0) __nanosleep(40);
        __threadfence_block();
      }
      const float* d = DEP + dc * 15 * LS_RS + l;
#pragma unroll
      for (int q = 0; q < 9; ++q) gGR.a[q] += d[q * LS_RS];
      gGt.x += d[9 * LS_RS]; gGt.y += d[10 * LS_RS]; gGt.z += d[11 * LS_RS];
      ja.x += d[12 * LS_RS]; ja.y += d[13 * LS_RS]; ja.z += d[14 * LS_RS];
    }
  };
  auto load_GR = [&](int j, m3& GR) {
    const float* ax = AX + j * 12 * LS_RS + l;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) GR.a[3 * r + c] = ax[(4 * r + c) * LS_RS];
    }
  };
  auto load_J = [&](int j) {
    const float* p = JRS + 3 * j * LS_RS + l;
    return mk3(p[0], p[LS_RS], p[2 * LS_RS]);
  };
  // final values of joint i: g_theta_i (into the theta rows) and dL/dJrest_i -> g_beta through JS
  auto finish = [&](int i, const m3& gM, f3 dJ) {
    float* tp = TH + 3 * i * LS_RS + l;
    if (a.g_theta) {
      const f3 gth = rodrigues_vjp(mk3(tp[0], tp[LS_RS], tp[2 * LS_RS]), gM);
      tp[0] = gth.x; tp[LS_RS] = gth.y; tp[2 * LS_RS] = gth.z;
    }
    if (a.jrest_closed && a.g_beta) {
      const float dj[3] = {dJ.x, dJ.y, dJ.z};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float4* row = reinterpret_cast<const float4*>(tJST + (3 * i + c) * jsw);
#pragma unroll
        for (int q = 0; q < (NBMAX + 4) / 4; ++q) {
          if (4 * q < jsw) {
            const float4 t = row[q];
            const float tv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int k = 4 * q + u - 1;
              if (k >= 0 && k < NBMAX) gbeta[k] = fmaf(tv[u], dj[c], gbeta[k]);
            }
          }
        }
      }
    }
  };

  // joint i is done with the walk: its A slot (dead from here on) keeps (dL/dM_i | dL/dJrest_i) for the per-joint pass
  auto park = [&](int i, const m3& gM, f3 dJ) {
    float* ax = AX + i * 12 * LS_RS + l;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) ax[(4 * r + c) * LS_RS] = gM.a[3 * r + c];
    }
    ax[3 * LS_RS] = dJ.x; ax[7 * LS_RS] = dJ.y; ax[11 * LS_RS] = dJ.z;
  };

  // ---- chains, deepest round first; a chain of a deeper round is never scheduled behind a shallower one on its warp, so
  // the flag waits cannot deadlock -------------------------------------------------------------------------------------------
  for (int r = m.ls_nround - 1; r >= 0; --r) {
    for (int ci = iSchedPtr[r * LS_NW + w]; ci < iSchedPtr[r * LS_NW + w + 1]; ++ci) {
      const int c = iSchedChain[ci];
      const int t0 = iChainPtr[c], t1 = iChainPtr[c + 1];
      m3 cGR;        // running contribution of the joints below, to be added to the next joint's dL/dG
      f3 cGt = mk3(0.f, 0.f, 0.f), cJ = mk3(0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 9; ++q) cGR.a[q] = 0.f;
      float4 g[3];
      load_gA(iChainJoint[t1 - 1], g);
      m3 GRi;
      load_GR(iChainJoint[t1 - 1], GRi);
      for (int t = t1 - 1; t >= t0; --t) {
        const int i = iChainJoint[t];
        const int p = (t > t0) ? iChainJoint[t - 1] : iChainParent[c];
        float4 gn[3];
        if (t > t0) load_gA(p, gn);   // the next joint's dL/dA, in flight under this joint's arithmetic
        m3 GRp;
        load_GR(p, GRp);
        const f3 Ji = load_J(i), Jp = load_J(p);
        m3 gGR; f3 gGt, ja;
        own_grad(i, g, GRi, Ji, gGR, gGt, ja, iDepPtr[i], iDepPtr[i + 1], true);
#pragma unroll
        for (int q = 0; q < 9; ++q) gGR.a[q] += cGR.a[q];
        gGt.x += cGt.x; gGt.y += cGt.y; gGt.z += cGt.z;
        ja.x += cJ.x; ja.y += cJ.y; ja.z += cJ.z;
        const m3 M = matmul_AtB(GRp, GRi);
        const f3 tl = mk3(Ji.x - Jp.x, Ji.y - Jp.y, Ji.z - Jp.z);
        const m3 gM = matmul_AtB(GRp, gGR);
        const f3 gtl = mulT(GRp, gGt);
        cGR = matmul_ABt(gGR, M);
        cGR.a[0] = fmaf(gGt.x, tl.x, cGR.a[0]); cGR.a[1] = fmaf(gGt.x, tl.y, cGR.a[1]); cGR.a[2] = fmaf(gGt.x, tl.z, cGR.a[2]);
        cGR.a[3] = fmaf(gGt.y, tl.x, cGR.a[3]); cGR.a[4] = fmaf(gGt.y, tl.y, cGR.a[4]); cGR.a[5] = fmaf(gGt.y, tl.z, cGR.a[5]);
        cGR.a[6] = fmaf(gGt.z, tl.x, cGR.a[6]); cGR.a[7] = fmaf(gGt.z, tl.y, cGR.a[7]); cGR.a[8] = fmaf(gGt.z, tl.z, cGR.a[8]);
        cGt = gGt;
        cJ = mk3(-gtl.x, -gtl.y, -gtl.z);
        park(i, gM, mk3(ja.x + gtl.x, ja.y + gtl.y, ja.z + gtl.z));
        GRi = GRp;
        if (t > t0) { g[0] = gn[0]; g[1] = gn[1]; g[2] = gn[2]; }
      }
      float* d = DEP + c * 15 * LS_RS + l;
#pragma unroll
      for (int q = 0; q < 9; ++q) d[q * LS_RS] = cGR.a[q];
      d[9 * LS_RS] = cGt.x; d[10 * LS_RS] = cGt.y; d[11 * LS_RS] = cGt.z;
      d[12 * LS_RS] = cJ.x; d[13 * LS_RS] = cJ.y; d[14 * LS_RS] = cJ.z;
      __threadfence_block();
      __syncwarp();
      if (l == 0) FLAG[c] = 1;
    }
  }
  __syncthreads();
  // root: G_0 = (R_0 | J_0).  Its deposits (one per chain that starts below the root: 13 on the STICK rig) are first
  // summed row by row by 15 warps into the first chain's slot, then warp 0 finishes the joint
  const int rd0 = iDepPtr[0], rd1 = iDepPtr[1];
  if (w < 15 && rd1 - rd0 > 1) {
    float* first = DEP + iDepChain[rd0] * 15 * LS_RS + w * LS_RS + l;
    float sum = *first;
    for (int t = rd0 + 1; t < rd1; ++t) sum += DEP[iDepChain[t] * 15 * LS_RS + w * LS_RS + l];
    *first = sum;
  }
  __syncthreads();
  if (w == 0) {   // root, while the other warps finish the joints below it
    float4 g[3];
    load_gA(0, g);
    m3 GR0, gGR; f3 gGt, ja;
    load_GR(0, GR0);
    own_grad(0, g, GR0, load_J(0), gGR, gGt, ja, rd0, min(rd1, rd0 + 1), false);
    finish(0, gGR, mk3(ja.x + gGt.x, ja.y + gGt.y, ja.z + gGt.z));
  } else {
    // ---- per joint: Rodrigues VJP and the rest-joint gradient through JS ----------------------------------------------
    for (int i = w; i < J; i += LS_NW - 1) {
      const float* ax = AX + i * 12 * LS_RS + l;
      m3 gM;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) gM.a[3 * r + c] = ax[(4 * r + c) * LS_RS];
      }
      finish(i, gM, mk3(ax[3 * LS_RS], ax[7 * LS_RS], ax[11 * LS_RS]));
    }
  }
  __syncthreads();

  // ---- g_theta rows out; g_beta / g_trans: warp partials -> sums over the warps (fixed order) -------------------------
  if (a.g_theta)
    for (int s = w; s < nlive; s += LS_NW) {
      float* out = a.g_theta + (long long)(b0 + s) * J3;
      for (int e = l; e < J3; e += 32) out[e] = TH[e * LS_RS + s];
    }
  float* RED = AX;   // [NW][NBMAX + 3] rows
#pragma unroll
  for (int k = 0; k < NBMAX; ++k) RED[(w * (NBMAX + 3) + k) * LS_RS + l] = gbeta[k];
#pragma unroll
  for (int c = 0; c < 3; ++c) RED[(w * (NBMAX + 3) + NBMAX + c) * LS_RS + l] = gtr[c];
  __syncthreads();
  for (int e = w; e < nb + 3; e += LS_NW) {   // warp per output element, lane = sample
    const int row = e < nb ? e : NBMAX + (e - nb);
    float sum = 0.f;
#pragma unroll
    for (int wv = 0; wv < LS_NW; ++wv) sum += RED[(wv * (NBMAX + 3) + row) * LS_RS + l];
    if (a.gbt_part) sum += a.gbt_part[(long long)bs * (nb + 3) + e];   // dense mode: the skinning pass's summed partials
    if (l < nlive) {
      if (e < nb) { if (a.g_beta) a.g_beta[(long long)(b0 + l) * nb + e] = sum; }
      else if (a.g_trans) a.g_trans[(long long)(b0 + l) * 3 + (e - nb)] = sum;
    }
  }
}

}  // namespace smil
