// Rodrigues + kinematic chain, forward and backward: ONE WARP PER BATCH ELEMENT, the parent tree
// walked level by level out of shared memory.
//
// Replaces  batch_rodrigues                         /root/reference/smal_model/batch_lbs.py:31-50
//           batch_global_rigid_transformation       /root/reference/smal_model/batch_lbs.py:75-197
//           rest-joint regression                   /root/reference/smal_model/smal_torch.py:257-264
//           posed-joint regression                  /root/reference/smal_model/smal_torch.py:343-351
// The two regressions are evaluated in closed form from the packed tables (packer.py) when the
// call has no del_v / v_template override / pose blendshapes, so they never touch vertices.
//
// Why a warp per sample: at B=4096 the path has only 4096 independent samples, so the per-sample
// work has to be spread over lanes to fill 148 SMs.  Per-joint work (Rodrigues, local transforms,
// outputs, the Rodrigues VJP) runs lane-per-joint; the level walk runs lane-per-joint-of-the-level
// (<= max level width lanes busy) and is kept as short as possible: the local transform
// M_i = S_p^-1 R_i S_i and its translation are precomputed, the backward child->parent accumulation is
// a float4 gather over records that share one layout, and the backward reads the global transforms
// the forward saved instead of re-walking the tree.  Shared memory per sample: 32 J floats forward,
// 48 J floats backward (J = 55: 7.1 KB / 10.6 KB -> 31 / 21 samples resident per SM).
#pragma once
#include "smil_common.cuh"

namespace smil {

constexpr int POSE_WPB = 2;                 // warps (= samples) per CTA
constexpr int POSE_THREADS = 32 * POSE_WPB;

struct PoseArgs {
  int B, nb;
  uint32_t flags;
  int jrest_mode;   // 0 = closed form (JT + beta JS), 1 = static J, 2 = buffer jrest_in
  int joints_mode;  // 0 = none, 1 = closed-form pairs (+rowsum*trans), 2 = copy of new_J (static)
  const float* beta;  long long beta_stride;
  const float* theta; long long theta_stride;
  const float* trans; long long trans_stride;
  const float* ls;    long long ls_stride;
  const float* bt;    long long bt_stride;
  const float* jrest_in;  // [B][J][3]
  float* jrest_out;       // [B][J][3] forward: the rest joints it used, kept for the backward (or null)
  int jrest_closed;       // backward: the rest joints were JT + beta JS, so dL/dbeta gets JS . dL/dJrest
  // forward outputs
  float* Rs;      // [B][J][9] or null
  float* A;       // [B][J][12] (3x4 row-major) or null
  float* pp;      // [ceil(B/2)][pp_blk] per-pair parameter blocks for the fused skinning passes, or null:
                  //   A2 [J][12][2] | beta2 [NB2][2] | trans2 [4][2], every entry interleaved (sample 2p, sample 2p+1)
  int pp_blk;     // floats per block (fused_kernels.cuh::pair_block_floats)
  float* newJ;    // [B][J][3] or null
  float* joints;  // [B][J][3] or null
  // backward inputs
  const float* A_saved;   // [B][J][12] written by the forward (its 3x3 part is the global rotation) or null
  const float* pp_saved;  // the pair blocks the forward wrote (A2 part is read) or null
  const float* gA_dense;  // [B][J][12] or null
  const float* gA_slots;  // per-(tile, bone) partials of the skin pass or null; the slots of bone k are sb_ptr[k] .. sb_ptr[k+1]
  int nslots;
  const int* sb_ptr;             // [J+1] (device) slot ranges per bone for the table set the skin pass used
  long long slot_bstride;        // float4 units between consecutive samples of one slot ...
  long long slot_sstride;        // ... and between consecutive slots of one sample ([B][nslots][12]: 3 nslots, 3; [nslots][B][12]: 3, 3 B)
  const float* gbt_part;  // [B][ntile][nb+3] partial (g_beta, g_trans) from the fused skin pass, or null
  int ntile;
  int joints_in_skin;     // 1: the fused skin pass already folded Jreg g_joints into g_verts
  const float* g_joints;  // [B][J][3] or null
  const float* g_newJ;    // [B][J][3] or null (extra gradient on J_transformed)
  const float* g_Rs;      // [B][J][9] or null
  const float* g_Rs2;     // second additive [B][J][9] (pose-blendshape path) or null
  // backward outputs
  float* g_theta;   // [B][J][3] or [B][J][9]
  float* g_ls;      // [B][J][3]
  float* g_bt;      // [B][J][3]
  float* g_beta;    // [B][nb]   (written when beta_final != 0)
  float* g_trans;   // [B][3]
  float* g_jrest;   // [B][J][3] (generic mode: consumed by the regressor backward)
  int beta_final;   // 1: this kernel writes the complete g_beta / g_trans (fused mode)
};

// shared-memory layout per sample (float offsets); every block is 16-byte aligned
struct PoseSmem {
  int J, nbp;
  __host__ __device__ int x() const { return 0; }               // [J][12] local (M | tl); bwd later: (gM | g_logscale)
  __host__ __device__ int g() const { return 12 * J; }          // [J][12] global (GR | Gt); before the walk: Rs staging [J][9]
  __host__ __device__ int jr() const { return 24 * J; }         // [J][4]  rest joints; bwd later: gradient on them
  __host__ __device__ int beta() const { return 28 * J; }       // [nbp]
  __host__ __device__ int tail() const { return 28 * J + nbp; }
  __host__ __device__ int scale() const { return tail(); }      // [J][4] exp(log-scales), dead once the M_j are built
  __host__ __device__ int fwd_floats() const { return tail() + 4 * J; }
  // backward: gradient records [J][20] = (gGR | gGt)(12) (g_logscale acc, -)(4) (gJ acc, -)(4); overlays the scales
  __host__ __device__ int rec() const { return tail(); }
  __host__ __device__ int bwd_floats() const { return tail() + 20 * J; }
};

__device__ __forceinline__ void ld_rec(const float* rec, m3& M, f3& w) {
  const float4* p = reinterpret_cast<const float4*>(rec);
  float4 a = p[0], b = p[1], c = p[2];
  M.a[0] = a.x; M.a[1] = a.y; M.a[2] = a.z; w.x = a.w;
  M.a[3] = b.x; M.a[4] = b.y; M.a[5] = b.z; w.y = b.w;
  M.a[6] = c.x; M.a[7] = c.y; M.a[8] = c.z; w.z = c.w;
}
__device__ __forceinline__ void st_rec(float* rec, const m3& M, f3 w) {
  float4* p = reinterpret_cast<float4*>(rec);
  p[0] = make_float4(M.a[0], M.a[1], M.a[2], w.x);
  p[1] = make_float4(M.a[3], M.a[4], M.a[5], w.y);
  p[2] = make_float4(M.a[6], M.a[7], M.a[8], w.z);
}
__device__ __forceinline__ f3 rec_w(const float* rec) { return mk3(rec[3], rec[7], rec[11]); }

// Phase A (forward and backward): beta, rest joints, rotations, scales, local transforms.
//   X[j] = (M_j | tl_j) with M_j = S_p^-1 R_j S_j (batch_lbs.py:163-173), tl_j = J_j - J_p (+ flipped
//   betas_trans, batch_lbs.py:148,159-160);  X[0] = (R_0 | J_0)  (the root is never scaled, :151).
//   jr[j] = rest joint J_j.   `stage` receives Rs packed [J][9]; `sc` the scales [J][4].
__device__ __forceinline__ void pose_local_state(const DevModel& m, const PoseArgs& a, float* X, float* jr, float* sbeta,
                                                 float* stage, float* sc, int b, int l) {
  const int J = m.J;
  const bool rotmat = (a.flags & SMIL_F_THETA_IS_ROTMAT) != 0;
  const int nbp = (max(a.nb, 1) + 3) & ~3;   // == PoseSmem::nbp: the padding reads as zero betas
  for (int k = l; k < nbp; k += 32) sbeta[k] = (k < a.nb) ? a.beta[(long long)b * a.beta_stride + k] : 0.f;
  if (rotmat) {  // [J][9] contiguous per sample -> staging, coalesced
    const float* src = a.theta + (long long)b * a.theta_stride;
    for (int e = l; e < 9 * J; e += 32) stage[e] = src[e];
  }
  __syncwarp();
  for (int e = l; e < 3 * J; e += 32) {
    float v;
    if (a.jrest_mode == 1) {
      v = m.J_static[e];
    } else if (a.jrest_mode == 2) {
      v = a.jrest_in[(long long)b * 3 * J + e];
    } else {
      v = m.JT[e];
      for (int k = 0; k < a.nb; k += 4) {   // four table loads in flight per step (the tables sit in L1/L2)
        const float j0 = m.JS[k * 3 * J + e];
        const float j1 = (k + 1 < a.nb) ? m.JS[(k + 1) * 3 * J + e] : 0.f;
        const float j2 = (k + 2 < a.nb) ? m.JS[(k + 2) * 3 * J + e] : 0.f;
        const float j3 = (k + 3 < a.nb) ? m.JS[(k + 3) * 3 * J + e] : 0.f;
        const float4 bb = reinterpret_cast<const float4*>(sbeta)[k >> 2];
        v = fmaf(bb.x, j0, v); v = fmaf(bb.y, j1, v); v = fmaf(bb.z, j2, v); v = fmaf(bb.w, j3, v);
      }
    }
    const int j = e / 3, c = e - 3 * j;
    jr[j * 4 + c] = v;
  }
  for (int j = l; j < J; j += 32) {
    m3 R;
    if (rotmat) {
#pragma unroll
      for (int q = 0; q < 9; ++q) R.a[q] = stage[j * 9 + q];
    } else {
      R = rodrigues(ld3(a.theta + (long long)b * a.theta_stride + j * 3));
#pragma unroll
      for (int q = 0; q < 9; ++q) stage[j * 9 + q] = R.a[q];
    }
    st_rec(X + j * 12, R, mk3(0.f, 0.f, 0.f));
    f3 s = mk3(1.f, 1.f, 1.f);
    if (a.ls) {
      const float* p = a.ls + (long long)b * a.ls_stride + j * 3;
      s = mk3(expf(p[0]), expf(p[1]), expf(p[2]));
    }
    sc[j * 4 + 0] = s.x; sc[j * 4 + 1] = s.y; sc[j * 4 + 2] = s.z;
  }
  __syncwarp();
  const bool prop = (a.flags & SMIL_F_PROPAGATE_SCALING) != 0;
  for (int j = l; j < J; j += 32) {
    m3 R; f3 dummy;
    ld_rec(X + j * 12, R, dummy);
    const f3 Jj = ld3(jr + j * 4);
    if (j == 0) {
      st_rec(X, R, Jj);
      continue;
    }
    const int p = m.parents[j];
    const f3 Jp = ld3(jr + p * 4);
    const f3 si = ld3(sc + j * 4);
    f3 pin = mk3(1.f, 1.f, 1.f);
    if (!prop) { const f3 sp = ld3(sc + p * 4); pin = mk3(1.0f / sp.x, 1.0f / sp.y, 1.0f / sp.z); }
    m3 M;
    M.a[0] = pin.x * R.a[0] * si.x; M.a[1] = pin.x * R.a[1] * si.y; M.a[2] = pin.x * R.a[2] * si.z;
    M.a[3] = pin.y * R.a[3] * si.x; M.a[4] = pin.y * R.a[4] * si.y; M.a[5] = pin.y * R.a[5] * si.z;
    M.a[6] = pin.z * R.a[6] * si.x; M.a[7] = pin.z * R.a[7] * si.y; M.a[8] = pin.z * R.a[8] * si.z;
    f3 tl = mk3(Jj.x - Jp.x, Jj.y - Jp.y, Jj.z - Jp.z);
    if (a.bt) {
      const float* bp = a.bt + (long long)b * a.bt_stride + j * 3;
      tl.x += bp[0]; tl.y -= bp[1]; tl.z += bp[2];  // y flipped for Unreal data, batch_lbs.py:148
    }
    st_rec(X + j * 12, M, tl);
  }
  __syncwarp();
}

// Level walk: G[i] = (G[p].R M_i | G[p].R tl_i + G[p].t);  G[0] = X[0].
__device__ __forceinline__ void pose_walk(const DevModel& m, const float* X, float* G, int l) {
  if (l == 0) {
    m3 R0; f3 J0;
    ld_rec(X, R0, J0);
    st_rec(G, R0, J0);
  }
  __syncwarp();
  for (int lev = 1; lev < m.num_levels; ++lev) {
    const int beg = m.level_ptr[lev], end = m.level_ptr[lev + 1];
    for (int q = beg + l; q < end; q += 32) {
      const int i = m.order[q], p = m.parents[i];
      m3 M, GRp; f3 tl, Gtp;
      ld_rec(X + i * 12, M, tl);
      ld_rec(G + p * 12, GRp, Gtp);
      const m3 GR = matmul(GRp, M);
      f3 Gt = mul(GRp, tl);
      Gt.x += Gtp.x; Gt.y += Gtp.y; Gt.z += Gtp.z;
      st_rec(G + i * 12, GR, Gt);
    }
    __syncwarp();
  }
}

// coalesced float4 copy of n floats (n % 4 == 0, both sides 16-byte aligned)
__device__ __forceinline__ void warp_copy4(float* dst, const float* src, int n, int l) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  for (int e = l; e < (n >> 2); e += 32) d4[e] = s4[e];
}

__global__ void __launch_bounds__(POSE_THREADS) pose_fwd_kernel(DevModel m, PoseArgs a, PoseSmem L) {
  extern __shared__ __align__(16) float smem_pose[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int b_raw = blockIdx.x * POSE_WPB + w;
  const bool live = b_raw < a.B;           // an odd batch leaves the last CTA's second warp without a sample:
  const int b = live ? b_raw : a.B - 1;    // it recomputes the last one and writes nothing per-sample
  const int J = m.J;
  float* sm = smem_pose + (size_t)w * ((L.fwd_floats() + 3) & ~3);
  float* X = sm + L.x();
  float* G = sm + L.g();
  float* jr = sm + L.jr();
  float* sbeta = sm + L.beta();
  float* sc = sm + L.scale();
  pose_local_state(m, a, X, jr, sbeta, /*stage=*/G, sc, b, l);
  if (a.jrest_out && live) {
    float* o = a.jrest_out + (long long)b * 3 * J;
    for (int e = l; e < 3 * J; e += 32) o[e] = jr[e + e / 3];
  }
  // Rs output from the packed staging copy (a sample's 9J floats are 16-byte aligned only if 9J % 4 == 0)
  if (a.Rs && live) {
    float* out = a.Rs + (long long)b * 9 * J;
    if (((9 * J) & 3) == 0) warp_copy4(out, G, 9 * J, l);
    else for (int e = l; e < 9 * J; e += 32) out[e] = G[e];
  }
  __syncwarp();
  pose_walk(m, X, G, l);
  // A = [GR | Gt - GR Jrest] -> X (free after the walk), new_J = Gt
  for (int j = l; j < J; j += 32) {
    m3 GR; f3 Gt;
    ld_rec(G + j * 12, GR, Gt);
    const f3 rot = mul(GR, ld3(jr + j * 4));
    st_rec(X + j * 12, GR, mk3(Gt.x - rot.x, Gt.y - rot.y, Gt.z - rot.z));
  }
  __syncwarp();
  if (a.A && live) warp_copy4(a.A + (long long)b * 12 * J, X, 12 * J, l);
  if (a.pp) {
    // both samples of the CTA's pair, interleaved element by element: (X0[e], X1[e], X0[e+1], X1[e+1]) per float4
    static_assert(POSE_WPB == 2, "a CTA is one sample pair");
    __syncthreads();
    const int wstride = (L.fwd_floats() + 3) & ~3;
    const float* X0 = smem_pose + L.x();
    const float* X1 = smem_pose + wstride + L.x();
    float* blk = a.pp + (long long)blockIdx.x * a.pp_blk;
    float4* out = reinterpret_cast<float4*>(blk);
    for (int i = threadIdx.x; i < 6 * J; i += POSE_THREADS) {
      const int j = i / 6;
      out[i + j * (PPB / 4 - 6)] = make_float4(X0[2 * i], X1[2 * i], X0[2 * i + 1], X1[2 * i + 1]);
    }
    const int nb2 = (a.pp_blk - PPB * J - 8) >> 1;
    for (int i = threadIdx.x; i < 2 * nb2; i += POSE_THREADS) {
      const int k = i >> 1, h = i & 1;
      blk[PPB * J + i] = (k < a.nb) ? smem_pose[h * wstride + L.beta() + k] : 0.f;
    }
    for (int i = threadIdx.x; i < 8; i += POSE_THREADS) {
      const int c = i >> 1, h = i & 1;
      const int bs = min(blockIdx.x * 2 + h, a.B - 1);
      blk[PPB * J + 2 * nb2 + i] = (a.trans && c < 3) ? a.trans[(long long)bs * a.trans_stride + c] : 0.f;
    }
  }
  if (!live) return;
  if (a.newJ || (a.joints && a.joints_mode == 2)) {
    for (int e = l; e < 3 * J; e += 32) {
      const int j = e / 3, c = e - 3 * j;
      const float v = G[j * 12 + 4 * c + 3];
      if (a.newJ) a.newJ[(long long)b * 3 * J + e] = v;
      if (a.joints && a.joints_mode == 2) a.joints[(long long)b * 3 * J + e] = v;
    }
  }
  if (a.joints && a.joints_mode == 1) {
    // joints_j = sum over coupled bones k of A_k [c0 + sum_m beta_m cm ; c0.w] + rowsum_j * trans, lane per
    // (joint, bone) pair over the joint-sorted records (coalesced float4 table loads); the pairs of one joint
    // are neighbouring lanes, folded by a segmented shuffle sum, run heads accumulate into sc (dead scales).
    f3 tr = mk3(0.f, 0.f, 0.f);
    if (a.trans) tr = ld3(a.trans + (long long)b * a.trans_stride);
    for (int j = l; j < J; j += 32) {
      const float rs = m.jreg_rowsum[j];
      sc[j * 4 + 0] = rs * tr.x; sc[j * 4 + 1] = rs * tr.y; sc[j * 4 + 2] = rs * tr.z;
    }
    __syncwarp();
    const int NQ = m.pr_nq;
    for (int s0 = 0; s0 < m.npairs_pad; s0 += 32) {
      const float4* R4 = m.pjrec4 + (size_t)(s0 >> 5) * NQ * 32 + l;
      const float4 h = R4[0], c0 = R4[32];
      const int bone = __float_as_int(h.x), joint = __float_as_int(h.y);
      f3 q = mk3(c0.x, c0.y, c0.z);
      for (int t4 = 0; 4 * t4 < a.nb; ++t4) {
        const float4 cx = R4[(2 + 3 * t4) * 32], cy = R4[(3 + 3 * t4) * 32], cz = R4[(4 + 3 * t4) * 32];
        const float4 bb = reinterpret_cast<const float4*>(sbeta)[t4];
        q.x = fmaf(bb.x, cx.x, fmaf(bb.y, cx.y, fmaf(bb.z, cx.z, fmaf(bb.w, cx.w, q.x))));
        q.y = fmaf(bb.x, cy.x, fmaf(bb.y, cy.y, fmaf(bb.z, cy.z, fmaf(bb.w, cy.w, q.y))));
        q.z = fmaf(bb.x, cz.x, fmaf(bb.y, cz.y, fmaf(bb.z, cz.z, fmaf(bb.w, cz.w, q.z))));
      }
      m3 AR; f3 At;
      ld_rec(X + max(bone, 0) * 12, AR, At);
      const f3 r = mul(AR, q);
      float vx = fmaf(At.x, h.z, r.x), vy = fmaf(At.y, h.z, r.y), vz = fmaf(At.z, h.z, r.z);
      if (bone < 0) { vx = 0.f; vy = 0.f; vz = 0.f; }
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int oj = __shfl_down_sync(0xffffffffu, joint, d);
        const float ox = __shfl_down_sync(0xffffffffu, vx, d), oy = __shfl_down_sync(0xffffffffu, vy, d),
                    oz = __shfl_down_sync(0xffffffffu, vz, d);
        if ((l + d < 32) && oj == joint) { vx += ox; vy += oy; vz += oz; }
      }
      const int pjn = __shfl_up_sync(0xffffffffu, joint, 1);
      if (bone >= 0 && (l == 0 || pjn != joint)) {   // one run head per joint and pass
        sc[joint * 4 + 0] += vx; sc[joint * 4 + 1] += vy; sc[joint * 4 + 2] += vz;
      }
      __syncwarp();
    }
    for (int e = l; e < 3 * J; e += 32) a.joints[(long long)b * 3 * J + e] = sc[e + e / 3];
  }
}

// Backward.  Gradient records rec[j] (20 floats) share one layout between "what joint j has received"
// and "what joint j hands to its parent", so the child -> parent accumulation is a plain float4 gather
// in a fixed order (deterministic, no atomics):
//   [0..11]  (gGR | gGt)   before j is processed: dL/dG_j;        after: contribution to dL/dG_parent
//   [12..14] log-scale     before: sum of the children's d/ds_j;   after: d/ds_parent from j (non-propagating mode)
//   [16..18] rest joint    before: dL/dJ_j so far;                 after: dL/dJ_parent from j (= -gtl_j)
// When j is processed its final values move to X[j] = (gM_j | g_logscale_j) and jr[j] = dL/dJ_j, from where
// the lane-per-joint epilogue produces g_theta (Rodrigues VJP), g_logscale, g_betas_trans, g_beta.
__global__ void __launch_bounds__(POSE_THREADS) pose_bwd_kernel(DevModel m, PoseArgs a, PoseSmem L, int pairs_mode) {
  extern __shared__ __align__(16) float smem_pose[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int b = blockIdx.x * POSE_WPB + w;
  if (b >= a.B) return;
  const int J = m.J, nb = a.nb;
  const int per = (L.bwd_floats() + 3) & ~3;
  float* sm = smem_pose + (size_t)w * per;
  float* X = sm + L.x();
  float* G = sm + L.g();
  float* jr = sm + L.jr();
  float* sbeta = sm + L.beta();
  float* rec = sm + L.rec();
  float gbeta[16];   // lane-private dL/dbeta partials (the fused path serves nb <= 16)
#pragma unroll
  for (int k = 0; k < 16; ++k) gbeta[k] = 0.f;
  const bool prop = (a.flags & SMIL_F_PROPAGATE_SCALING) != 0;
  const bool rotmat = (a.flags & SMIL_F_THETA_IS_ROTMAT) != 0;

  pose_local_state(m, a, X, jr, sbeta, /*stage=*/G, /*sc=*/rec, b, l);
  if (a.A_saved) {
    warp_copy4(G, a.A_saved + (long long)b * 12 * J, 12 * J, l);  // (GR | At): only GR is used below
    __syncwarp();
  } else if (a.pp_saved) {
    const float* src = a.pp_saved + (long long)(b >> 1) * a.pp_blk + (b & 1);
    for (int e = l; e < 12 * J; e += 32) G[e] = src[2 * e + (e / 12) * (PPB - 24)];
    __syncwarp();
  } else {
    pose_walk(m, X, G, l);
  }

  // ---- dL/dA per bone: dense buffer + slot partials of the skin pass (+ closed-form posed joints) --------
  f3 gtr = mk3(0.f, 0.f, 0.f);
  for (int e = l; e < 3 * J; e += 32) {
    const int j = e / 3, r = e - 3 * j;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.gA_dense) acc = reinterpret_cast<const float4*>(a.gA_dense)[(long long)b * 3 * J + e];
    if (a.gA_slots) {
      const float4* sl = reinterpret_cast<const float4*>(a.gA_slots) + (long long)b * a.slot_bstride + r;
      const long long ss = a.slot_sstride;
      int si = a.sb_ptr[j];
      const int se = a.sb_ptr[j + 1];
      for (; si + 3 < se; si += 4) {   // four independent loads in flight (the slots live in L2 / HBM)
        const float4 t0 = sl[si * ss], t1 = sl[(si + 1) * ss], t2 = sl[(si + 2) * ss], t3 = sl[(si + 3) * ss];
        acc.x += (t0.x + t1.x) + (t2.x + t3.x); acc.y += (t0.y + t1.y) + (t2.y + t3.y);
        acc.z += (t0.z + t1.z) + (t2.z + t3.z); acc.w += (t0.w + t1.w) + (t2.w + t3.w);
      }
      if (si + 1 < se) {
        const float4 t0 = sl[si * ss], t1 = sl[(si + 1) * ss];
        acc.x += t0.x + t1.x; acc.y += t0.y + t1.y; acc.z += t0.z + t1.z; acc.w += t0.w + t1.w;
        si += 2;
      }
      if (si < se) {
        const float4 t = sl[si * ss];
        acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
      }
    }
    if (pairs_mode) {  // joints_j = ... + rowsum_j * trans
      const float rs = m.jreg_rowsum[j];
      const float gjj = a.g_joints[(long long)b * 3 * J + e];
      if (r == 0) gtr.x = fmaf(rs, gjj, gtr.x);
      if (r == 1) gtr.y = fmaf(rs, gjj, gtr.y);
      if (r == 2) gtr.z = fmaf(rs, gjj, gtr.z);
    }
    reinterpret_cast<float4*>(rec)[j * 5 + r] = acc;
  }
  __syncwarp();
  if (pairs_mode) {
    // closed-form posed joints, joints_j = sum_k A_k [q_jk ; c0w_jk] with q_jk = c0_jk + sum_m beta_m cm_jkm:
    // lane per (joint, bone) pair in bone-sorted order, dL/dA_k = sum_j g_joints_j (x) [q_jk ; c0w_jk] by a
    // segmented shuffle reduction over the lanes that share a bone (fixed order, no atomics).
    const int NQ = m.pr_nq;
    for (int s0 = 0; s0 < m.npairs_pad; s0 += 32) {
      const float4* R4 = m.pbrec4 + (size_t)(s0 >> 5) * NQ * 32 + l;   // coalesced float4 table loads
      const float4 h = R4[0], c0 = R4[32];
      const int bone = __float_as_int(h.x);
      const bool real = bone >= 0;
      const int kb = real ? bone : 0;
      f3 gj = mk3(0.f, 0.f, 0.f);
      if (real) gj = ld3(a.g_joints + ((long long)b * J + __float_as_int(h.y)) * 3);
      const float* Gk = G + kb * 12;
      const f3 tg = mk3(fmaf(Gk[0], gj.x, fmaf(Gk[4], gj.y, Gk[8] * gj.z)), fmaf(Gk[1], gj.x, fmaf(Gk[5], gj.y, Gk[9] * gj.z)),
                        fmaf(Gk[2], gj.x, fmaf(Gk[6], gj.y, Gk[10] * gj.z)));   // GR_k^T g_joints_j
      f3 q = mk3(c0.x, c0.y, c0.z);
      const float c0w = h.z;
#pragma unroll
      for (int t4 = 0; t4 < 4; ++t4) {
        if (4 * t4 < nb) {
          const float4 cx = R4[(2 + 3 * t4) * 32], cy = R4[(3 + 3 * t4) * 32], cz = R4[(4 + 3 * t4) * 32];
          const float4 bb = reinterpret_cast<const float4*>(sbeta)[t4];
          q.x = fmaf(bb.x, cx.x, fmaf(bb.y, cx.y, fmaf(bb.z, cx.z, fmaf(bb.w, cx.w, q.x))));
          q.y = fmaf(bb.x, cy.x, fmaf(bb.y, cy.y, fmaf(bb.z, cy.z, fmaf(bb.w, cy.w, q.y))));
          q.z = fmaf(bb.x, cz.x, fmaf(bb.y, cz.y, fmaf(bb.z, cz.z, fmaf(bb.w, cz.w, q.z))));
          gbeta[4 * t4 + 0] += fmaf(tg.x, cx.x, fmaf(tg.y, cy.x, tg.z * cz.x));
          gbeta[4 * t4 + 1] += fmaf(tg.x, cx.y, fmaf(tg.y, cy.y, tg.z * cz.y));
          gbeta[4 * t4 + 2] += fmaf(tg.x, cx.z, fmaf(tg.y, cy.z, tg.z * cz.z));
          gbeta[4 * t4 + 3] += fmaf(tg.x, cx.w, fmaf(tg.y, cy.w, tg.z * cz.w));
        }
      }
      float v[12];
      v[0] = gj.x * q.x; v[1] = gj.x * q.y; v[2] = gj.x * q.z; v[3] = gj.x * c0w;
      v[4] = gj.y * q.x; v[5] = gj.y * q.y; v[6] = gj.y * q.z; v[7] = gj.y * c0w;
      v[8] = gj.z * q.x; v[9] = gj.z * q.y; v[10] = gj.z * q.z; v[11] = gj.z * c0w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int ob = __shfl_down_sync(0xffffffffu, bone, d);
        const bool same = (l + d < 32) && ob == bone;
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          const float o = __shfl_down_sync(0xffffffffu, v[i], d);
          if (same) v[i] += o;
        }
      }
      const int pb_ = __shfl_up_sync(0xffffffffu, bone, 1);
      if (real && (l == 0 || pb_ != bone)) {   // first lane of a bone's run holds the run's sum
        float4* d4 = reinterpret_cast<float4*>(rec + bone * 20);
        float4 t0 = d4[0], t1 = d4[1], t2 = d4[2];
        t0.x += v[0]; t0.y += v[1]; t0.z += v[2]; t0.w += v[3];
        t1.x += v[4]; t1.y += v[5]; t1.z += v[6]; t1.w += v[7];
        t2.x += v[8]; t2.y += v[9]; t2.z += v[10]; t2.w += v[11];
        d4[0] = t0; d4[1] = t1; d4[2] = t2;
      }
      __syncwarp();
    }
  }
  // ---- A = [GR | Gt - GR Jrest], new_J = Gt  ->  initial (gGR | gGt) and dL/dJrest ---------------------------
  for (int j = l; j < J; j += 32) {
    m3 gAR, GR; f3 gAt, dummy;
    ld_rec(rec + j * 20, gAR, gAt);
    ld_rec(G + j * 12, GR, dummy);
    const f3 Jr = ld3(jr + j * 4);
    m3 gGR;
    gGR.a[0] = gAR.a[0] - gAt.x * Jr.x; gGR.a[1] = gAR.a[1] - gAt.x * Jr.y; gGR.a[2] = gAR.a[2] - gAt.x * Jr.z;
    gGR.a[3] = gAR.a[3] - gAt.y * Jr.x; gGR.a[4] = gAR.a[4] - gAt.y * Jr.y; gGR.a[5] = gAR.a[5] - gAt.y * Jr.z;
    gGR.a[6] = gAR.a[6] - gAt.z * Jr.x; gGR.a[7] = gAR.a[7] - gAt.z * Jr.y; gGR.a[8] = gAR.a[8] - gAt.z * Jr.z;
    f3 gGt = gAt;
    if (a.joints_mode == 2 && a.g_joints) {
      const f3 g = ld3(a.g_joints + ((long long)b * J + j) * 3);
      gGt.x += g.x; gGt.y += g.y; gGt.z += g.z;
    }
    if (a.g_newJ) {
      const f3 g = ld3(a.g_newJ + ((long long)b * J + j) * 3);
      gGt.x += g.x; gGt.y += g.y; gGt.z += g.z;
    }
    st_rec(rec + j * 20, gGR, gGt);
    const f3 r = mulT(GR, gAt);
    float4* tailp = reinterpret_cast<float4*>(rec + j * 20 + 12);
    tailp[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    tailp[1] = make_float4(-r.x, -r.y, -r.z, 0.f);
  }
  __syncwarp();
  // ---- leaves -> root -----------------------------------------------------------------------------------
  for (int lev = m.num_levels - 1; lev >= 1; --lev) {
    const int beg = m.level_ptr[lev], end = m.level_ptr[lev + 1];
    for (int q = beg + l; q < end; q += 32) {
      const int i = m.order[q], p = m.parents[i];
      m3 M, GRp, gGR; f3 tl, gGt, dummy;
      ld_rec(X + i * 12, M, tl);
      ld_rec(G + p * 12, GRp, dummy);
      ld_rec(rec + i * 20, gGR, gGt);
      const float4 lsacc = reinterpret_cast<const float4*>(rec + i * 20 + 12)[0];
      const float4 jacc = reinterpret_cast<const float4*>(rec + i * 20 + 12)[1];
      const m3 gM = matmul_AtB(GRp, gGR);  // GRp^T gGR
      const f3 gtl = mulT(GRp, gGt);
      m3 dGR = matmul_ABt(gGR, M);
      dGR.a[0] = fmaf(gGt.x, tl.x, dGR.a[0]); dGR.a[1] = fmaf(gGt.x, tl.y, dGR.a[1]); dGR.a[2] = fmaf(gGt.x, tl.z, dGR.a[2]);
      dGR.a[3] = fmaf(gGt.y, tl.x, dGR.a[3]); dGR.a[4] = fmaf(gGt.y, tl.y, dGR.a[4]); dGR.a[5] = fmaf(gGt.y, tl.z, dGR.a[5]);
      dGR.a[6] = fmaf(gGt.z, tl.x, dGR.a[6]); dGR.a[7] = fmaf(gGt.z, tl.y, dGR.a[7]); dGR.a[8] = fmaf(gGt.z, tl.z, dGR.a[8]);
      // M = diag(1/s_p) R diag(s_i): d/dlog s_i = column sums of gM o M, d/dlog s_p = -(row sums)
      const float e0 = gM.a[0] * M.a[0], e1 = gM.a[1] * M.a[1], e2 = gM.a[2] * M.a[2];
      const float e3 = gM.a[3] * M.a[3], e4 = gM.a[4] * M.a[4], e5 = gM.a[5] * M.a[5];
      const float e6 = gM.a[6] * M.a[6], e7 = gM.a[7] * M.a[7], e8 = gM.a[8] * M.a[8];
      const f3 gls = mk3(lsacc.x + (e0 + e3 + e6), lsacc.y + (e1 + e4 + e7), lsacc.z + (e2 + e5 + e8));
      f3 dls = mk3(0.f, 0.f, 0.f);
      if (!prop) dls = mk3(-(e0 + e1 + e2), -(e3 + e4 + e5), -(e6 + e7 + e8));
      st_rec(X + i * 12, gM, gls);
      jr[i * 4 + 0] = jacc.x + gtl.x; jr[i * 4 + 1] = jacc.y + gtl.y; jr[i * 4 + 2] = jacc.z + gtl.z;
      st_rec(rec + i * 20, dGR, gGt);
      float4* tailp = reinterpret_cast<float4*>(rec + i * 20 + 12);
      tailp[0] = make_float4(dls.x, dls.y, dls.z, 0.f);
      tailp[1] = make_float4(-gtl.x, -gtl.y, -gtl.z, 0.f);
    }
    __syncwarp();
    // parents (level lev-1) gather what their children just wrote; 5 float4 per record
    const int pbeg = m.level_ptr[lev - 1], pend = m.level_ptr[lev];
    const int ntask = (pend - pbeg) * 5;
    for (int t = l; t < ntask; t += 32) {
      const int pi = t / 5, q5 = t - pi * 5;
      const int p = m.order[pbeg + pi];
      const int c0 = m.child_ptr[p], c1 = m.child_ptr[p + 1];
      if (c0 == c1) continue;
      float4* dst = reinterpret_cast<float4*>(rec + p * 20) + q5;
      float4 acc = *dst;
      for (int ci = c0; ci < c1; ++ci) {
        const float4 t4 = reinterpret_cast<const float4*>(rec + m.child_idx[ci] * 20)[q5];
        acc.x += t4.x; acc.y += t4.y; acc.z += t4.z; acc.w += t4.w;
      }
      *dst = acc;
    }
    __syncwarp();
  }
  // root: G_0 = (R_0 | J_0)
  if (l == 0) {
    m3 gGR; f3 gGt;
    ld_rec(rec, gGR, gGt);
    const float4 lsacc = reinterpret_cast<const float4*>(rec + 12)[0];
    const float4 jacc = reinterpret_cast<const float4*>(rec + 12)[1];
    st_rec(X, gGR, mk3(lsacc.x, lsacc.y, lsacc.z));
    jr[0] = jacc.x + gGt.x; jr[1] = jacc.y + gGt.y; jr[2] = jacc.z + gGt.z;
  }
  __syncwarp();
  // ---- lane-per-joint epilogue: g_theta, g_logscale, g_betas_trans -----------------------------------------
  for (int j = l; j < J; j += 32) {
    m3 gR; f3 gls;
    ld_rec(X + j * 12, gR, gls);  // gM_j (root: dL/dR_0)
    if (j > 0 && a.ls) {          // gR = diag(1/s_p) gM diag(s_j); without log-scales both are the identity
      const float* pi_ = a.ls + (long long)b * a.ls_stride + j * 3;
      const f3 si = mk3(expf(pi_[0]), expf(pi_[1]), expf(pi_[2]));
      f3 pin = mk3(1.f, 1.f, 1.f);
      if (!prop) {
        const float* pp = a.ls + (long long)b * a.ls_stride + m.parents[j] * 3;
        pin = mk3(1.0f / expf(pp[0]), 1.0f / expf(pp[1]), 1.0f / expf(pp[2]));
      }
      gR.a[0] *= pin.x * si.x; gR.a[1] *= pin.x * si.y; gR.a[2] *= pin.x * si.z;
      gR.a[3] *= pin.y * si.x; gR.a[4] *= pin.y * si.y; gR.a[5] *= pin.y * si.z;
      gR.a[6] *= pin.z * si.x; gR.a[7] *= pin.z * si.y; gR.a[8] *= pin.z * si.z;
    }
    if (a.g_Rs) {
      const float* gp = a.g_Rs + ((long long)b * J + j) * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t) gR.a[t] += gp[t];
    }
    if (a.g_Rs2) {
      const float* gp = a.g_Rs2 + ((long long)b * J + j) * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t) gR.a[t] += gp[t];
    }
    if (a.g_theta) {
      if (rotmat) {
        float* o = a.g_theta + ((long long)b * J + j) * 9;
#pragma unroll
        for (int t = 0; t < 9; ++t) o[t] = gR.a[t];
      } else {
        st3(a.g_theta + ((long long)b * J + j) * 3, rodrigues_vjp(ld3(a.theta + (long long)b * a.theta_stride + j * 3), gR));
      }
    }
    if (a.g_ls) st3(a.g_ls + ((long long)b * J + j) * 3, gls);
    if (a.g_bt) {
      f3 g = mk3(0.f, 0.f, 0.f);
      if (j > 0) g = mk3(-rec[j * 20 + 16], rec[j * 20 + 17], -rec[j * 20 + 18]);  // gtl with y flipped back
      st3(a.g_bt + ((long long)b * J + j) * 3, g);
    }
  }
  if (a.g_jrest) {
    float* o = a.g_jrest + (long long)b * 3 * J;
    for (int e = l; e < 3 * J; e += 32) { const int j = e / 3; o[e] = jr[j * 4 + (e - 3 * j)]; }
  }
  // ---- g_beta, g_trans (fused mode: this kernel owns the final values) ----------------------------------------
  if (a.beta_final) {
    const int ne = nb + 3;
    float part = 0.f;  // lane k < nb+3: sum over tiles of the skin pass' partial (g_beta_k | g_trans)
    if (a.gbt_part && l < ne)
      for (int t = 0; t < a.ntile; ++t) part += a.gbt_part[((long long)b * a.ntile + t) * ne + l];
    float mine = 0.f;
    if (a.g_beta) {
      if (a.jrest_closed) {   // rest joints JT + beta JS: dL/dbeta_k += JS_k . dL/dJrest, all k in one pass over the joints
        for (int e = l; e < 3 * J; e += 32) {
          const float gje = jr[e + e / 3];
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (k < nb) gbeta[k] = fmaf(m.JS[(long long)k * 3 * J + e], gje, gbeta[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        if (k < nb) {
          float s = gbeta[k];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
          if (l == k) mine = s;
        }
      }
      if (l < nb) a.g_beta[(long long)b * nb + l] = mine + part;
    }
    if (a.g_trans) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gtr.x += __shfl_xor_sync(0xffffffffu, gtr.x, o);
        gtr.y += __shfl_xor_sync(0xffffffffu, gtr.y, o);
        gtr.z += __shfl_xor_sync(0xffffffffu, gtr.z, o);
      }
      if (l >= nb && l < ne) {
        const int c = l - nb;
        a.g_trans[(long long)b * 3 + c] = part + (c == 0 ? gtr.x : (c == 1 ? gtr.y : gtr.z));
      }
    }
  }
}

// Backward without per-joint log-scales (the headline configuration): the local rotation of joint i is then
// M_i = R_i = GR_p^T GR_i and tl_i = J_i - J_p (+ betas_trans), both recomputable from the saved global
// transforms and rest joints, so the local state (Rodrigues, X) is never rebuilt.  Per sample:
//   G   [J][12]  (GR_j | Jrest_j) in the w column; once j is processed: (gM_j | dL/dJrest_j) for the epilogue
//   rec [J][16]  (gGR | gGt)(12) + (dL/dJrest acc / contribution to the parent)(4)
// = 28 J + nbp floats (6.2 KB for J = 55), half of the general kernel's, so a whole 4096-sample batch is
// resident at once (one wave) instead of 1.4 waves.
struct PoseSmemLean {
  int J, nbp;
  __host__ __device__ int g() const { return 0; }
  __host__ __device__ int rec() const { return 12 * J; }
  __host__ __device__ int beta() const { return 28 * J; }
  __host__ __device__ int floats() const { return (28 * J + nbp + 3) & ~3; }
};

__global__ void __launch_bounds__(POSE_THREADS, 14) pose_bwd_lean_kernel(DevModel m, PoseArgs a, PoseSmemLean L, int pairs_mode) {
  extern __shared__ __align__(16) float smem_pose[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int b = blockIdx.x * POSE_WPB + w;
  if (b >= a.B) return;
  const int J = m.J, nb = a.nb;
  float* sm = smem_pose + (size_t)w * L.floats();
  float* G = sm + L.g();
  float* rec = sm + L.rec();
  float* sbeta = sm + L.beta();
  float gbeta[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) gbeta[k] = 0.f;
  const bool rotmat = (a.flags & SMIL_F_THETA_IS_ROTMAT) != 0;

  for (int k = l; k < L.nbp; k += 32) sbeta[k] = (k < nb) ? a.beta[(long long)b * a.beta_stride + k] : 0.f;
  {
    const float* src = a.pp_saved + (long long)(b >> 1) * a.pp_blk + (b & 1);
    const float* jsrc = (a.jrest_mode == 1) ? m.J_static : a.jrest_in + (long long)b * 3 * J;
    for (int e = l; e < 12 * J; e += 32) {
      const int j = e / 12, q = e - 12 * j;
      G[e] = ((q & 3) == 3) ? jsrc[j * 3 + (q >> 2)] : src[2 * e + j * (PPB - 24)];
    }
  }
  __syncwarp();

  // ---- dL/dA per bone: slot partials of the skin pass (+ closed-form posed joints) ---------------------------
  f3 gtr = mk3(0.f, 0.f, 0.f);
  for (int e = l; e < 3 * J; e += 32) {
    const int j = e / 3, r = e - 3 * j;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.gA_dense) acc = reinterpret_cast<const float4*>(a.gA_dense)[(long long)b * 3 * J + e];   // slots already summed
    if (a.gA_slots) {
      const float4* sl = reinterpret_cast<const float4*>(a.gA_slots) + (long long)b * a.slot_bstride + r;
      const long long ss = a.slot_sstride;
      int si = a.sb_ptr[j];
      const int se = a.sb_ptr[j + 1];
      for (; si + 3 < se; si += 4) {
        const float4 t0 = sl[si * ss], t1 = sl[(si + 1) * ss], t2 = sl[(si + 2) * ss], t3 = sl[(si + 3) * ss];
        acc.x += (t0.x + t1.x) + (t2.x + t3.x); acc.y += (t0.y + t1.y) + (t2.y + t3.y);
        acc.z += (t0.z + t1.z) + (t2.z + t3.z); acc.w += (t0.w + t1.w) + (t2.w + t3.w);
      }
      if (si + 1 < se) {
        const float4 t0 = sl[si * ss], t1 = sl[(si + 1) * ss];
        acc.x += t0.x + t1.x; acc.y += t0.y + t1.y; acc.z += t0.z + t1.z; acc.w += t0.w + t1.w;
        si += 2;
      }
      if (si < se) {
        const float4 t = sl[si * ss];
        acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
      }
    }
    if (pairs_mode) {  // joints_j = ... + rowsum_j * trans
      const float rs = m.jreg_rowsum[j];
      const float gjj = a.g_joints[(long long)b * 3 * J + e];
      if (r == 0) gtr.x = fmaf(rs, gjj, gtr.x);
      if (r == 1) gtr.y = fmaf(rs, gjj, gtr.y);
      if (r == 2) gtr.z = fmaf(rs, gjj, gtr.z);
    }
    reinterpret_cast<float4*>(rec)[j * 4 + r] = acc;
  }
  __syncwarp();
  if (pairs_mode) {   // same (joint, bone) pair pass as pose_bwd_kernel
    const int NQ = m.pr_nq;
    for (int s0 = 0; s0 < m.npairs_pad; s0 += 32) {
      const float4* R4 = m.pbrec4 + (size_t)(s0 >> 5) * NQ * 32 + l;
      const float4 h = R4[0], c0 = R4[32];
      const int bone = __float_as_int(h.x);
      const bool real = bone >= 0;
      const int kb = real ? bone : 0;
      f3 gj = mk3(0.f, 0.f, 0.f);
      if (real) gj = ld3(a.g_joints + ((long long)b * J + __float_as_int(h.y)) * 3);
      const float* Gk = G + kb * 12;
      const f3 tg = mk3(fmaf(Gk[0], gj.x, fmaf(Gk[4], gj.y, Gk[8] * gj.z)), fmaf(Gk[1], gj.x, fmaf(Gk[5], gj.y, Gk[9] * gj.z)),
                        fmaf(Gk[2], gj.x, fmaf(Gk[6], gj.y, Gk[10] * gj.z)));   // GR_k^T g_joints_j
      f3 q = mk3(c0.x, c0.y, c0.z);
      const float c0w = h.z;
#pragma unroll
      for (int t4 = 0; t4 < 4; ++t4) {
        if (4 * t4 < nb) {
          const float4 cx = R4[(2 + 3 * t4) * 32], cy = R4[(3 + 3 * t4) * 32], cz = R4[(4 + 3 * t4) * 32];
          const float4 bb = reinterpret_cast<const float4*>(sbeta)[t4];
          q.x = fmaf(bb.x, cx.x, fmaf(bb.y, cx.y, fmaf(bb.z, cx.z, fmaf(bb.w, cx.w, q.x))));
          q.y = fmaf(bb.x, cy.x, fmaf(bb.y, cy.y, fmaf(bb.z, cy.z, fmaf(bb.w, cy.w, q.y))));
          q.z = fmaf(bb.x, cz.x, fmaf(bb.y, cz.y, fmaf(bb.z, cz.z, fmaf(bb.w, cz.w, q.z))));
          gbeta[4 * t4 + 0] += fmaf(tg.x, cx.x, fmaf(tg.y, cy.x, tg.z * cz.x));
          gbeta[4 * t4 + 1] += fmaf(tg.x, cx.y, fmaf(tg.y, cy.y, tg.z * cz.y));
          gbeta[4 * t4 + 2] += fmaf(tg.x, cx.z, fmaf(tg.y, cy.z, tg.z * cz.z));
          gbeta[4 * t4 + 3] += fmaf(tg.x, cx.w, fmaf(tg.y, cy.w, tg.z * cz.w));
        }
      }
      float v[12];
      v[0] = gj.x * q.x; v[1] = gj.x * q.y; v[2] = gj.x * q.z; v[3] = gj.x * c0w;
      v[4] = gj.y * q.x; v[5] = gj.y * q.y; v[6] = gj.y * q.z; v[7] = gj.y * c0w;
      v[8] = gj.z * q.x; v[9] = gj.z * q.y; v[10] = gj.z * q.z; v[11] = gj.z * c0w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int ob = __shfl_down_sync(0xffffffffu, bone, d);
        const bool same = (l + d < 32) && ob == bone;
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          const float o = __shfl_down_sync(0xffffffffu, v[i], d);
          if (same) v[i] += o;
        }
      }
      const int pb_ = __shfl_up_sync(0xffffffffu, bone, 1);
      if (real && (l == 0 || pb_ != bone)) {
        float4* d4 = reinterpret_cast<float4*>(rec + bone * 16);
        float4 t0 = d4[0], t1 = d4[1], t2 = d4[2];
        t0.x += v[0]; t0.y += v[1]; t0.z += v[2]; t0.w += v[3];
        t1.x += v[4]; t1.y += v[5]; t1.z += v[6]; t1.w += v[7];
        t2.x += v[8]; t2.y += v[9]; t2.z += v[10]; t2.w += v[11];
        d4[0] = t0; d4[1] = t1; d4[2] = t2;
      }
      __syncwarp();
    }
  }
  // ---- A = [GR | Gt - GR Jrest], new_J = Gt  ->  initial (gGR | gGt) and dL/dJrest ---------------------------
  for (int j = l; j < J; j += 32) {
    m3 gAR, GR; f3 gAt, Jr;
    ld_rec(rec + j * 16, gAR, gAt);
    ld_rec(G + j * 12, GR, Jr);
    m3 gGR;
    gGR.a[0] = gAR.a[0] - gAt.x * Jr.x; gGR.a[1] = gAR.a[1] - gAt.x * Jr.y; gGR.a[2] = gAR.a[2] - gAt.x * Jr.z;
    gGR.a[3] = gAR.a[3] - gAt.y * Jr.x; gGR.a[4] = gAR.a[4] - gAt.y * Jr.y; gGR.a[5] = gAR.a[5] - gAt.y * Jr.z;
    gGR.a[6] = gAR.a[6] - gAt.z * Jr.x; gGR.a[7] = gAR.a[7] - gAt.z * Jr.y; gGR.a[8] = gAR.a[8] - gAt.z * Jr.z;
    f3 gGt = gAt;
    if (a.joints_mode == 2 && a.g_joints) {
      const f3 g = ld3(a.g_joints + ((long long)b * J + j) * 3);
      gGt.x += g.x; gGt.y += g.y; gGt.z += g.z;
    }
    st_rec(rec + j * 16, gGR, gGt);
    const f3 r = mulT(GR, gAt);
    reinterpret_cast<float4*>(rec + j * 16)[3] = make_float4(-r.x, -r.y, -r.z, 0.f);
  }
  __syncwarp();
  // ---- leaves -> root -----------------------------------------------------------------------------------
  for (int lev = m.num_levels - 1; lev >= 1; --lev) {
    const int beg = m.level_ptr[lev], end = m.level_ptr[lev + 1];
    for (int q = beg + l; q < end; q += 32) {
      const int i = m.order[q], p = m.parents[i];
      m3 GRi, GRp, gGR; f3 Ji, Jp, gGt;
      ld_rec(G + i * 12, GRi, Ji);
      ld_rec(G + p * 12, GRp, Jp);
      ld_rec(rec + i * 16, gGR, gGt);
      const float4 jacc = reinterpret_cast<const float4*>(rec + i * 16)[3];
      const m3 M = matmul_AtB(GRp, GRi);   // R_i
      f3 tl = mk3(Ji.x - Jp.x, Ji.y - Jp.y, Ji.z - Jp.z);
      if (a.bt) {
        const float* bp = a.bt + (long long)b * a.bt_stride + i * 3;
        tl.x += bp[0]; tl.y -= bp[1]; tl.z += bp[2];
      }
      const m3 gM = matmul_AtB(GRp, gGR);
      const f3 gtl = mulT(GRp, gGt);
      m3 dGR = matmul_ABt(gGR, M);
      dGR.a[0] = fmaf(gGt.x, tl.x, dGR.a[0]); dGR.a[1] = fmaf(gGt.x, tl.y, dGR.a[1]); dGR.a[2] = fmaf(gGt.x, tl.z, dGR.a[2]);
      dGR.a[3] = fmaf(gGt.y, tl.x, dGR.a[3]); dGR.a[4] = fmaf(gGt.y, tl.y, dGR.a[4]); dGR.a[5] = fmaf(gGt.y, tl.z, dGR.a[5]);
      dGR.a[6] = fmaf(gGt.z, tl.x, dGR.a[6]); dGR.a[7] = fmaf(gGt.z, tl.y, dGR.a[7]); dGR.a[8] = fmaf(gGt.z, tl.z, dGR.a[8]);
      st_rec(G + i * 12, gM, mk3(jacc.x + gtl.x, jacc.y + gtl.y, jacc.z + gtl.z));   // children of i are done: G[i] is free
      st_rec(rec + i * 16, dGR, gGt);
      reinterpret_cast<float4*>(rec + i * 16)[3] = make_float4(-gtl.x, -gtl.y, -gtl.z, 0.f);
      if (a.g_bt) st3(a.g_bt + ((long long)b * J + i) * 3, mk3(gtl.x, -gtl.y, gtl.z));
    }
    __syncwarp();
    const int pbeg = m.level_ptr[lev - 1], pend = m.level_ptr[lev];
    const int ntask = (pend - pbeg) * 4;
    for (int t = l; t < ntask; t += 32) {
      const int pi = t >> 2, q4 = t & 3;
      const int p = m.order[pbeg + pi];
      const int c0 = m.child_ptr[p], c1 = m.child_ptr[p + 1];
      if (c0 == c1) continue;
      float4* dst = reinterpret_cast<float4*>(rec + p * 16) + q4;
      float4 acc = *dst;
      for (int ci = c0; ci < c1; ++ci) {
        const float4 t4 = reinterpret_cast<const float4*>(rec + m.child_idx[ci] * 16)[q4];
        acc.x += t4.x; acc.y += t4.y; acc.z += t4.z; acc.w += t4.w;
      }
      *dst = acc;
    }
    __syncwarp();
  }
  if (l == 0) {   // root: G_0 = (R_0 | J_0)
    m3 gGR; f3 gGt;
    ld_rec(rec, gGR, gGt);
    const float4 jacc = reinterpret_cast<const float4*>(rec)[3];
    st_rec(G, gGR, mk3(jacc.x + gGt.x, jacc.y + gGt.y, jacc.z + gGt.z));
    if (a.g_bt) st3(a.g_bt + (long long)b * J * 3, mk3(0.f, 0.f, 0.f));
  }
  __syncwarp();
  // ---- lane-per-joint epilogue: g_theta -------------------------------------------------------------------
  if (a.g_theta) {
    for (int j = l; j < J; j += 32) {
      m3 gR; f3 dJ;
      ld_rec(G + j * 12, gR, dJ);
      if (a.g_Rs) {
        const float* gp = a.g_Rs + ((long long)b * J + j) * 9;
#pragma unroll
        for (int t = 0; t < 9; ++t) gR.a[t] += gp[t];
      }
      if (rotmat) {
        float* o = a.g_theta + ((long long)b * J + j) * 9;
#pragma unroll
        for (int t = 0; t < 9; ++t) o[t] = gR.a[t];
      } else {
        st3(a.g_theta + ((long long)b * J + j) * 3, rodrigues_vjp(ld3(a.theta + (long long)b * a.theta_stride + j * 3), gR));
      }
    }
  }
  // ---- g_beta, g_trans ---------------------------------------------------------------------------------------
  const int ne = nb + 3;
  float part = 0.f;
  if (a.gbt_part && l < ne)
    for (int t = 0; t < a.ntile; ++t) part += a.gbt_part[((long long)b * a.ntile + t) * ne + l];
  float mine = 0.f;
  if (a.g_beta) {
    if (a.jrest_closed) {
      for (int e = l; e < 3 * J; e += 32) {
        const int j = e / 3;
        const float gje = G[j * 12 + 3 + 4 * (e - 3 * j)];
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (k < nb) gbeta[k] = fmaf(m.JS[(long long)k * 3 * J + e], gje, gbeta[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (k < nb) {
        float s = gbeta[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (l == k) mine = s;
      }
    }
    if (l < nb) a.g_beta[(long long)b * nb + l] = mine + part;
  }
  if (a.g_trans) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      gtr.x += __shfl_xor_sync(0xffffffffu, gtr.x, o);
      gtr.y += __shfl_xor_sync(0xffffffffu, gtr.y, o);
      gtr.z += __shfl_xor_sync(0xffffffffu, gtr.z, o);
    }
    if (l >= nb && l < ne) {
      const int c = l - nb;
      a.g_trans[(long long)b * 3 + c] = part + (c == 0 ? gtr.x : (c == 1 ? gtr.y : gtr.z));
    }
  }
}

// ---- stand-alone Rodrigues (drop-in for batch_rodrigues) --------------------------------------------
__global__ void rodrigues_fwd_kernel(const float* __restrict__ theta, float* __restrict__ R, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  m3 M = rodrigues(ld3(theta + 3 * i));
#pragma unroll
  for (int q = 0; q < 9; ++q) R[9 * i + q] = M.a[q];
}
__global__ void rodrigues_bwd_kernel(const float* __restrict__ theta, const float* __restrict__ gR,
                                     float* __restrict__ gth, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  m3 G;
#pragma unroll
  for (int q = 0; q < 9; ++q) G.a[q] = gR[9 * i + q];
  st3(gth + 3 * i, rodrigues_vjp(ld3(theta + 3 * i), G));
}

}  // namespace smil
