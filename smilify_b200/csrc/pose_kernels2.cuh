// Rodrigues + kinematic chain, forward and (log-scale-free) backward: SEVERAL SAMPLES PER WARP.
//
// Replaces  batch_rodrigues                         /root/reference/smal_model/batch_lbs.py:31-50
//           batch_global_rigid_transformation       /root/reference/smal_model/batch_lbs.py:75-197
//           rest / posed joint regression           /root/reference/smal_model/smal_torch.py:257-264, 343-351
//
// The one-warp-per-sample kernels of pose_kernels.cuh leave most lanes idle: a tree level of the STICK rig is at
// most 12 joints wide, J = 55 joints fill 1.7 warp trips, and the (joint, bone) couplings of the closed-form joint
// regression were folded by 5-level segmented shuffle sums.  Here a warp owns SPW samples (2 for trees up to 16
// joints wide, 4 up to 8) and every stage runs over the flattened (sample, joint) -- or (sample, joint of the
// level) -- task list, so a level of 12 joints keeps 24 lanes busy, and the couplings are handled without
// shuffles: a (pair)-parallel stage evaluates q = c0 + cm beta for all samples of the warp into shared memory, a
// (sample, joint)-parallel stage (forward) / (sample, bone, row)-parallel stage (backward) sums them.  The walk is
// done IN PLACE (X_i = (M_i | tl_i) becomes (GR_i | Gt_i), then (GR_i | At_i)), which halves the shared memory per
// sample.  Only __syncwarp() separates the stages: a warp never waits for another one.
#pragma once
#include "pose_kernels.cuh"

namespace smil {

constexpr int POSE2_WPB = 4;   // warps per CTA (independent of each other)

// shared-memory floats per sample of the forward pass
__host__ __device__ inline int pose2_fwd_floats(int J, int nbp, int npairs) { return 16 * J + nbp + 4 * npairs; }

// task t of a flattened (sample, item) list with `n` items per sample -> (sample, item)
template <int SPW>
__device__ __forceinline__ void split_task(int t, int n, int& s, int& i) {
  s = 0; i = t;
#pragma unroll
  for (int q = 1; q < SPW; ++q)
    if (i >= n) { i -= n; ++s; }
}

template <int SPW>
__global__ void __launch_bounds__(32 * POSE2_WPB) pose_fwd2_kernel(DevModel m, PoseArgs a, int nbp) {
  extern __shared__ __align__(16) float smem_pose[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int J = m.J, nb = a.nb;
  const int b0 = (blockIdx.x * POSE2_WPB + w) * SPW;   // first sample of this warp (even: a warp holds whole pairs)
  if (b0 >= a.B) return;
  const int np = (a.joints && a.joints_mode == 1) ? m.npairs : 0;
  const int PS = pose2_fwd_floats(J, nbp, (a.joints && a.joints_mode == 1) ? m.npairs_pad : 0);
  float* sm = smem_pose + (size_t)w * SPW * PS;
  // a sample beyond the batch (odd tail) is computed as a duplicate of the last one and never written per-sample
  auto bs_of = [&](int s) { return min(b0 + s, a.B - 1); };
  auto live = [&](int s) { return b0 + s < a.B; };
#define P2_X(s) (sm + (s) * PS)
#define P2_JR(s) (sm + (s) * PS + 12 * J)
#define P2_BE(s) (sm + (s) * PS + 16 * J)
#define P2_Q(s) (sm + (s) * PS + 16 * J + nbp)
  const bool rotmat = (a.flags & SMIL_F_THETA_IS_ROTMAT) != 0;
  const bool prop = (a.flags & SMIL_F_PROPAGATE_SCALING) != 0;

  // ---- betas ---------------------------------------------------------------------------------------------
#pragma unroll
  for (int s = 0; s < SPW; ++s)
    for (int k = l; k < nbp; k += 32) P2_BE(s)[k] = (k < nb) ? a.beta[(long long)bs_of(s) * a.beta_stride + k] : 0.f;
  __syncwarp();
  // ---- rest joints: J = JT + sum_m beta_m JS_m, element-major so that the table loads are coalesced and shared by
  //      the warp's samples (smal_torch.py:257-264 in closed form; static rigs: the J table; generic path: a buffer)
  for (int e = l; e < 3 * J; e += 32) {
    const int j = e / 3, c = e - 3 * j;
    float v[SPW];
    if (a.jrest_mode == 1) {
      const float t = m.J_static[e];
#pragma unroll
      for (int s = 0; s < SPW; ++s) v[s] = t;
    } else if (a.jrest_mode == 2) {
#pragma unroll
      for (int s = 0; s < SPW; ++s) v[s] = a.jrest_in[(long long)bs_of(s) * 3 * J + e];
    } else {
      const float t = m.JT[e];
#pragma unroll
      for (int s = 0; s < SPW; ++s) v[s] = t;
      for (int k = 0; k < nb; ++k) {
        const float js = m.JS[k * 3 * J + e];
#pragma unroll
        for (int s = 0; s < SPW; ++s) v[s] = fmaf(P2_BE(s)[k], js, v[s]);
      }
    }
#pragma unroll
    for (int s = 0; s < SPW; ++s) {
      P2_JR(s)[j * 4 + c] = v[s];
      if (a.jrest_out && live(s)) a.jrest_out[(long long)(b0 + s) * 3 * J + e] = v[s];
    }
  }
  __syncwarp();
  // ---- rotations and local transforms X_j = (M_j | tl_j), M_j = S_p^-1 R_j S_j (batch_lbs.py:163-173),
  //      tl_j = J_j - J_p (+ flipped betas_trans, :148,159-160); X_0 = (R_0 | J_0) (the root is never scaled, :151)
  for (int t = l; t < SPW * J; t += 32) {
    int s, j;
    split_task<SPW>(t, J, s, j);
    const int bs = bs_of(s);
    m3 R;
    if (rotmat) {
      const float* src = a.theta + (long long)bs * a.theta_stride + j * 9;
#pragma unroll
      for (int q = 0; q < 9; ++q) R.a[q] = src[q];
    } else {
      R = rodrigues(ld3(a.theta + (long long)bs * a.theta_stride + j * 3));
      if (a.Rs && live(s)) {
        float* o = a.Rs + ((long long)(b0 + s) * J + j) * 9;
#pragma unroll
        for (int q = 0; q < 9; ++q) o[q] = R.a[q];
      }
    }
    const f3 Jj = ld3(P2_JR(s) + j * 4);
    float* Xj = P2_X(s) + j * 12;
    if (j == 0) {
      st_rec(Xj, R, Jj);
      continue;
    }
    const int p = m.parents[j];
    const f3 Jp = ld3(P2_JR(s) + p * 4);
    if (a.ls) {
      const float* pl = a.ls + (long long)bs * a.ls_stride;
      const f3 si = mk3(expf(pl[j * 3]), expf(pl[j * 3 + 1]), expf(pl[j * 3 + 2]));
      f3 pin = mk3(1.f, 1.f, 1.f);
      if (!prop) pin = mk3(1.0f / expf(pl[p * 3]), 1.0f / expf(pl[p * 3 + 1]), 1.0f / expf(pl[p * 3 + 2]));
      R.a[0] = pin.x * R.a[0] * si.x; R.a[1] = pin.x * R.a[1] * si.y; R.a[2] = pin.x * R.a[2] * si.z;
      R.a[3] = pin.y * R.a[3] * si.x; R.a[4] = pin.y * R.a[4] * si.y; R.a[5] = pin.y * R.a[5] * si.z;
      R.a[6] = pin.z * R.a[6] * si.x; R.a[7] = pin.z * R.a[7] * si.y; R.a[8] = pin.z * R.a[8] * si.z;
    }
    f3 tl = mk3(Jj.x - Jp.x, Jj.y - Jp.y, Jj.z - Jp.z);
    if (a.bt) {
      const float* bp = a.bt + (long long)bs * a.bt_stride + j * 3;
      tl.x += bp[0]; tl.y -= bp[1]; tl.z += bp[2];   // y flipped for Unreal data, batch_lbs.py:148
    }
    st_rec(Xj, R, tl);
  }
  __syncwarp();
  // ---- level walk, in place: X_i <- (GR_p M_i | GR_p tl_i + Gt_p) ----------------------------------------------
  for (int lev = 1; lev < m.num_levels; ++lev) {
    const int beg = m.level_ptr[lev], wid = m.level_ptr[lev + 1] - beg;
    for (int t = l; t < SPW * wid; t += 32) {
      int s, q;
      split_task<SPW>(t, wid, s, q);
      const int i = m.order[beg + q], p = m.parents[i];
      m3 M, GRp; f3 tl, Gtp;
      ld_rec(P2_X(s) + i * 12, M, tl);
      ld_rec(P2_X(s) + p * 12, GRp, Gtp);
      const m3 GR = matmul(GRp, M);
      f3 Gt = mul(GRp, tl);
      Gt.x += Gtp.x; Gt.y += Gtp.y; Gt.z += Gtp.z;
      st_rec(P2_X(s) + i * 12, GR, Gt);
    }
    __syncwarp();
  }
  // ---- outputs per joint: new_J = Gt; A = [GR | Gt - GR Jrest] (in place) -------------------------------------------
  for (int t = l; t < SPW * J; t += 32) {
    int s, j;
    split_task<SPW>(t, J, s, j);
    m3 GR; f3 Gt;
    ld_rec(P2_X(s) + j * 12, GR, Gt);
    const f3 rot = mul(GR, ld3(P2_JR(s) + j * 4));
    const f3 At = mk3(Gt.x - rot.x, Gt.y - rot.y, Gt.z - rot.z);
    st_rec(P2_X(s) + j * 12, GR, At);
    if (live(s)) {
      const long long o = ((long long)(b0 + s) * J + j) * 3;
      if (a.newJ) st3(a.newJ + o, Gt);
      if (a.joints && a.joints_mode == 2) st3(a.joints + o, Gt);
      if (a.A) {
        float4* d = reinterpret_cast<float4*>(a.A + ((long long)(b0 + s) * J + j) * 12);
        d[0] = make_float4(GR.a[0], GR.a[1], GR.a[2], At.x);
        d[1] = make_float4(GR.a[3], GR.a[4], GR.a[5], At.y);
        d[2] = make_float4(GR.a[6], GR.a[7], GR.a[8], At.z);
      }
    }
  }
  __syncwarp();
  // ---- per-pair parameter blocks of the fused skin passes: A2 [J][PPB] | beta2 | trans2, pair-interleaved ------------
  if (a.pp) {
#pragma unroll
    for (int pr = 0; pr < SPW / 2; ++pr) {
      if (b0 + 2 * pr >= a.B) break;
      const float* X0 = P2_X(2 * pr);
      const float* X1 = P2_X(2 * pr + 1);
      float* blk = a.pp + (long long)((b0 >> 1) + pr) * a.pp_blk;
      float4* out = reinterpret_cast<float4*>(blk);
      for (int i = l; i < 6 * J; i += 32) {
        const int j = i / 6, q = i - 6 * j;
        const float2 u0 = *reinterpret_cast<const float2*>(X0 + 12 * j + 2 * q);
        const float2 u1 = *reinterpret_cast<const float2*>(X1 + 12 * j + 2 * q);
        out[j * (PPB / 4) + q] = make_float4(u0.x, u1.x, u0.y, u1.y);
      }
      const int nb2 = (a.pp_blk - PPB * J - 8) >> 1;
      for (int i = l; i < 2 * nb2; i += 32) {
        const int k = i >> 1, h = i & 1;
        blk[PPB * J + i] = (k < nb) ? P2_BE(2 * pr + h)[k] : 0.f;
      }
      if (l < 8) {
        const int c = l >> 1, h = l & 1;
        blk[PPB * J + 2 * nb2 + l] = (a.trans && c < 3) ? a.trans[(long long)bs_of(2 * pr + h) * a.trans_stride + c] : 0.f;
      }
    }
  }
  // ---- posed joints in closed form: joints_j = sum_k A_k [q_jk ; c0w_jk] + rowsum_j trans, q_jk = c0_jk + cm_jk beta
  if (a.joints && a.joints_mode == 1) {   // (a rig without regressor entries still gets joints = rowsum * trans = 0)
    const int NQ = m.pr_nq;
    for (int s0 = 0; s0 < (np > 0 ? m.npairs_pad : 0); s0 += 32) {   // lane = pair: coalesced float4 table loads shared by the samples
      const float4* R4 = m.pjrec4 + (size_t)(s0 >> 5) * NQ * 32 + l;
      const float4 h = R4[0], c0 = R4[32];
      float qx[SPW], qy[SPW], qz[SPW];
#pragma unroll
      for (int s = 0; s < SPW; ++s) { qx[s] = c0.x; qy[s] = c0.y; qz[s] = c0.z; }
      for (int t4 = 0; 4 * t4 < nb; ++t4) {
        const float4 cx = R4[(2 + 3 * t4) * 32], cy = R4[(3 + 3 * t4) * 32], cz = R4[(4 + 3 * t4) * 32];
#pragma unroll
        for (int s = 0; s < SPW; ++s) {
          const float4 bb = reinterpret_cast<const float4*>(P2_BE(s))[t4];
          qx[s] = fmaf(bb.x, cx.x, fmaf(bb.y, cx.y, fmaf(bb.z, cx.z, fmaf(bb.w, cx.w, qx[s]))));
          qy[s] = fmaf(bb.x, cy.x, fmaf(bb.y, cy.y, fmaf(bb.z, cy.z, fmaf(bb.w, cy.w, qy[s]))));
          qz[s] = fmaf(bb.x, cz.x, fmaf(bb.y, cz.y, fmaf(bb.z, cz.z, fmaf(bb.w, cz.w, qz[s]))));
        }
      }
#pragma unroll
      for (int s = 0; s < SPW; ++s) reinterpret_cast<float4*>(P2_Q(s))[s0 + l] = make_float4(qx[s], qy[s], qz[s], h.z);
    }
    __syncwarp();
    for (int t = l; t < SPW * J; t += 32) {
      int s, j;
      split_task<SPW>(t, J, s, j);
      if (!live(s)) continue;
      const float rs = m.jreg_rowsum[j];
      f3 acc = mk3(0.f, 0.f, 0.f);
      if (a.trans) {
        const f3 tr = ld3(a.trans + (long long)(b0 + s) * a.trans_stride);
        acc = mk3(rs * tr.x, rs * tr.y, rs * tr.z);
      }
      for (int pi = m.pj_ptr[j]; pi < m.pj_ptr[j + 1]; ++pi) {
        const int k = m.pj_bone[pi];
        const float4 q = reinterpret_cast<const float4*>(P2_Q(s))[pi];
        m3 AR; f3 At;
        ld_rec(P2_X(s) + k * 12, AR, At);
        const f3 r = mul(AR, mk3(q.x, q.y, q.z));
        acc.x += fmaf(At.x, q.w, r.x); acc.y += fmaf(At.y, q.w, r.y); acc.z += fmaf(At.z, q.w, r.z);
      }
      st3(a.joints + ((long long)(b0 + s) * J + j) * 3, acc);
    }
  }
#undef P2_X
#undef P2_JR
#undef P2_BE
#undef P2_Q
}

}  // namespace smil

namespace smil {

// shared-memory floats per sample of the backward pass without per-joint log-scales:
//   G [J][12] (GR_j | Jrest_j in the w column; once j is processed: (gM_j | dL/dJrest_j)) | rec [J][16] ((gGR | gGt)(12) +
//   (dL/dJrest accumulator / contribution to the parent)(4)) | beta [nbp] | Q [npairs_pad][4] (q_jk ; c0w_jk), bone-sorted
__host__ __device__ inline int pose2_bwd_floats(int J, int nbp, int npairs_pad) { return 28 * J + nbp + 4 * npairs_pad; }

// Backward of the pose pass for calls without per-joint log-scales (the same contract as pose_bwd_lean_kernel: the local
// rotation R_i = GR_p^T GR_i and tl_i = J_i - J_p (+ betas_trans) are recomputed from the forward's saved global transforms
// and rest joints), SPW samples per warp, every stage over the flattened (sample, item) task list.
template <int SPW, int NBMAX>
__global__ void __launch_bounds__(32 * POSE2_WPB) pose_bwd2_kernel(DevModel m, PoseArgs a, int nbp, int pairs_mode) {
  extern __shared__ __align__(16) float smem_pose[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int J = m.J, nb = a.nb;
  const int b0 = (blockIdx.x * POSE2_WPB + w) * SPW;
  if (b0 >= a.B) return;
  const int npp = pairs_mode ? m.npairs_pad : 0;
  const int PS = pose2_bwd_floats(J, nbp, npp);
  float* sm = smem_pose + (size_t)w * SPW * PS;
  auto bs_of = [&](int s) { return min(b0 + s, a.B - 1); };
  auto live = [&](int s) { return b0 + s < a.B; };
#define P2_G(s) (sm + (s) * PS)
#define P2_REC(s) (sm + (s) * PS + 12 * J)
#define P2_BE(s) (sm + (s) * PS + 28 * J)
#define P2_Q(s) (sm + (s) * PS + 28 * J + nbp)
  const bool rotmat = (a.flags & SMIL_F_THETA_IS_ROTMAT) != 0;
  float gbeta[SPW][NBMAX];   // lane-private dL/dbeta partials, folded across the warp at the end
  float gtr[SPW][3];         // lane-private dL/dtrans partials from the posed-joint regression
#pragma unroll
  for (int s = 0; s < SPW; ++s) {
#pragma unroll
    for (int k = 0; k < NBMAX; ++k) gbeta[s][k] = 0.f;
    gtr[s][0] = gtr[s][1] = gtr[s][2] = 0.f;
  }

  // ---- betas; saved global transforms (pair-interleaved blocks: one float4 feeds both samples of a pair) ---------------
#pragma unroll
  for (int s = 0; s < SPW; ++s)
    for (int k = l; k < nbp; k += 32) P2_BE(s)[k] = (k < nb) ? a.beta[(long long)bs_of(s) * a.beta_stride + k] : 0.f;
#pragma unroll
  for (int pr = 0; pr < SPW / 2; ++pr) {
    const float4* src = reinterpret_cast<const float4*>(a.pp_saved + (long long)((bs_of(2 * pr)) >> 1) * a.pp_blk);
    float* G0 = P2_G(2 * pr);
    float* G1 = P2_G(2 * pr + 1);
    for (int i = l; i < 6 * J; i += 32) {
      const int j = i / 6, q = i - 6 * j;
      const float4 v = src[j * (PPB / 4) + q];
      *reinterpret_cast<float2*>(G0 + 12 * j + 2 * q) = make_float2(v.x, v.z);
      *reinterpret_cast<float2*>(G1 + 12 * j + 2 * q) = make_float2(v.y, v.w);
    }
  }
  __syncwarp();
  // rest joints into the w column (the saved block holds A's translation there, which the backward never reads)
  for (int e = l; e < 3 * J; e += 32) {
    const int j = e / 3, c = e - 3 * j;
#pragma unroll
    for (int s = 0; s < SPW; ++s)
      P2_G(s)[12 * j + 4 * c + 3] = (a.jrest_mode == 1) ? m.J_static[e] : a.jrest_in[(long long)bs_of(s) * 3 * J + e];
  }
  // ---- dL/dA per bone: the skin pass' partial slots -> rec rows ----------------------------------------------------------
  for (int t = l; t < SPW * 3 * J; t += 32) {
    int s, e;
    split_task<SPW>(t, 3 * J, s, e);
    const int j = e / 3, r = e - 3 * j;
    const int bs = bs_of(s);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.gA_dense) acc = reinterpret_cast<const float4*>(a.gA_dense)[(long long)bs * 3 * J + e];   // slots already summed
    if (a.gA_slots) {
      const float4* sl = reinterpret_cast<const float4*>(a.gA_slots) + (long long)bs * a.slot_bstride + r;
      const long long ss = a.slot_sstride;
      int si = a.sb_ptr[j];
      const int se = a.sb_ptr[j + 1];
      for (; si + 3 < se; si += 4) {   // four independent loads in flight (the slots live in L2 / HBM)
        const float4 t0 = sl[si * ss], t1 = sl[(si + 1) * ss], t2 = sl[(si + 2) * ss], t3 = sl[(si + 3) * ss];
        acc.x += (t0.x + t1.x) + (t2.x + t3.x); acc.y += (t0.y + t1.y) + (t2.y + t3.y);
        acc.z += (t0.z + t1.z) + (t2.z + t3.z); acc.w += (t0.w + t1.w) + (t2.w + t3.w);
      }
      for (; si < se; ++si) {
        const float4 t0 = sl[si * ss];
        acc.x += t0.x; acc.y += t0.y; acc.z += t0.z; acc.w += t0.w;
      }
    }
    if (pairs_mode) {  // joints_j = ... + rowsum_j * trans
      const float v = m.jreg_rowsum[j] * a.g_joints[(long long)bs * 3 * J + e];
#pragma unroll
      for (int q = 0; q < SPW; ++q)
        if (q == s) { if (r == 0) gtr[q][0] += v; else if (r == 1) gtr[q][1] += v; else gtr[q][2] += v; }
    }
    reinterpret_cast<float4*>(P2_REC(s))[j * 4 + r] = acc;
  }
  __syncwarp();
  if (pairs_mode) {
    // closed-form posed joints, joints_j = sum_k A_k [q_jk ; c0w_jk], q_jk = c0_jk + cm_jk beta.  Lane = pair (bone-sorted,
    // coalesced float4 table loads shared by the warp's samples): q into shared memory, and the dL/dbeta term
    // (GR_k^T g_joints_j) . cm_jk into lane-private partials.
    const int NQ = m.pr_nq;
    for (int s0 = 0; s0 < m.npairs_pad; s0 += 32) {
      const float4* R4 = m.pbrec4 + (size_t)(s0 >> 5) * NQ * 32 + l;
      const float4 h = R4[0], c0 = R4[32];
      const int bone = __float_as_int(h.x);
      const bool real = bone >= 0;
      const int kb = real ? bone : 0, jn = real ? __float_as_int(h.y) : 0;
      float qx[SPW], qy[SPW], qz[SPW], tx[SPW], ty[SPW], tz[SPW];
#pragma unroll
      for (int s = 0; s < SPW; ++s) {
        qx[s] = c0.x; qy[s] = c0.y; qz[s] = c0.z;
        f3 gj = mk3(0.f, 0.f, 0.f);
        if (real) gj = ld3(a.g_joints + ((long long)bs_of(s) * J + jn) * 3);
        const float* Gk = P2_G(s) + kb * 12;
        tx[s] = fmaf(Gk[0], gj.x, fmaf(Gk[4], gj.y, Gk[8] * gj.z));    // GR_k^T g_joints_j
        ty[s] = fmaf(Gk[1], gj.x, fmaf(Gk[5], gj.y, Gk[9] * gj.z));
        tz[s] = fmaf(Gk[2], gj.x, fmaf(Gk[6], gj.y, Gk[10] * gj.z));
      }
#pragma unroll
      for (int t4 = 0; t4 < NBMAX / 4; ++t4) {
        if (4 * t4 < nb) {
          const float4 cx = R4[(2 + 3 * t4) * 32], cy = R4[(3 + 3 * t4) * 32], cz = R4[(4 + 3 * t4) * 32];
#pragma unroll
          for (int s = 0; s < SPW; ++s) {
            const float4 bb = reinterpret_cast<const float4*>(P2_BE(s))[t4];
            qx[s] = fmaf(bb.x, cx.x, fmaf(bb.y, cx.y, fmaf(bb.z, cx.z, fmaf(bb.w, cx.w, qx[s]))));
            qy[s] = fmaf(bb.x, cy.x, fmaf(bb.y, cy.y, fmaf(bb.z, cy.z, fmaf(bb.w, cy.w, qy[s]))));
            qz[s] = fmaf(bb.x, cz.x, fmaf(bb.y, cz.y, fmaf(bb.z, cz.z, fmaf(bb.w, cz.w, qz[s]))));
            gbeta[s][4 * t4 + 0] += fmaf(tx[s], cx.x, fmaf(ty[s], cy.x, tz[s] * cz.x));
            gbeta[s][4 * t4 + 1] += fmaf(tx[s], cx.y, fmaf(ty[s], cy.y, tz[s] * cz.y));
            gbeta[s][4 * t4 + 2] += fmaf(tx[s], cx.z, fmaf(ty[s], cy.z, tz[s] * cz.z));
            gbeta[s][4 * t4 + 3] += fmaf(tx[s], cx.w, fmaf(ty[s], cy.w, tz[s] * cz.w));
          }
        }
      }
#pragma unroll
      for (int s = 0; s < SPW; ++s) reinterpret_cast<float4*>(P2_Q(s))[s0 + l] = make_float4(qx[s], qy[s], qz[s], h.z);
    }
    __syncwarp();
    // dL/dA_k += sum over the bone's couplings of g_joints_j (x) [q_jk ; c0w_jk]: one (sample, bone, row) task per lane, the
    // bone's couplings are the contiguous range pb_ptr[k] .. pb_ptr[k+1] of the bone-sorted table (fixed order, no atomics)
    for (int t = l; t < SPW * 3 * J; t += 32) {
      int s, e;
      split_task<SPW>(t, 3 * J, s, e);
      const int k = e / 3, r = e - 3 * k;
      const int p0 = m.pb_ptr[k], p1 = m.pb_ptr[k + 1];
      if (p0 == p1) continue;
      float4 acc = reinterpret_cast<const float4*>(P2_REC(s))[k * 4 + r];
      for (int pos = p0; pos < p1; ++pos) {
        const int jn = __float_as_int(reinterpret_cast<const float*>(m.pbrec4 + (size_t)(pos >> 5) * NQ * 32 + (pos & 31))[1]);
        const float gjr = a.g_joints[((long long)bs_of(s) * J + jn) * 3 + r];
        const float4 q = reinterpret_cast<const float4*>(P2_Q(s))[pos];
        acc.x = fmaf(gjr, q.x, acc.x); acc.y = fmaf(gjr, q.y, acc.y); acc.z = fmaf(gjr, q.z, acc.z); acc.w = fmaf(gjr, q.w, acc.w);
      }
      reinterpret_cast<float4*>(P2_REC(s))[k * 4 + r] = acc;
    }
    __syncwarp();
  }
  // ---- A = [GR | Gt - GR Jrest], new_J = Gt  ->  initial (gGR | gGt) and dL/dJrest -------------------------------------------
  for (int t = l; t < SPW * J; t += 32) {
    int s, j;
    split_task<SPW>(t, J, s, j);
    m3 gAR, GR; f3 gAt, Jr;
    ld_rec(P2_REC(s) + j * 16, gAR, gAt);
    ld_rec(P2_G(s) + j * 12, GR, Jr);
    m3 gGR;
    gGR.a[0] = gAR.a[0] - gAt.x * Jr.x; gGR.a[1] = gAR.a[1] - gAt.x * Jr.y; gGR.a[2] = gAR.a[2] - gAt.x * Jr.z;
    gGR.a[3] = gAR.a[3] - gAt.y * Jr.x; gGR.a[4] = gAR.a[4] - gAt.y * Jr.y; gGR.a[5] = gAR.a[5] - gAt.y * Jr.z;
    gGR.a[6] = gAR.a[6] - gAt.z * Jr.x; gGR.a[7] = gAR.a[7] - gAt.z * Jr.y; gGR.a[8] = gAR.a[8] - gAt.z * Jr.z;
    f3 gGt = gAt;
    if (a.joints_mode == 2 && a.g_joints) {
      const f3 g = ld3(a.g_joints + ((long long)bs_of(s) * J + j) * 3);
      gGt.x += g.x; gGt.y += g.y; gGt.z += g.z;
    }
    st_rec(P2_REC(s) + j * 16, gGR, gGt);
    const f3 r = mulT(GR, gAt);
    reinterpret_cast<float4*>(P2_REC(s) + j * 16)[3] = make_float4(-r.x, -r.y, -r.z, 0.f);
  }
  __syncwarp();
  // ---- leaves -> root -------------------------------------------------------------------------------------------------------
  for (int lev = m.num_levels - 1; lev >= 1; --lev) {
    const int beg = m.level_ptr[lev], wid = m.level_ptr[lev + 1] - beg;
    for (int t = l; t < SPW * wid; t += 32) {
      int s, q;
      split_task<SPW>(t, wid, s, q);
      const int i = m.order[beg + q], p = m.parents[i];
      float* G = P2_G(s);
      float* rec = P2_REC(s);
      m3 GRi, GRp, gGR; f3 Ji, Jp, gGt;
      ld_rec(G + i * 12, GRi, Ji);
      ld_rec(G + p * 12, GRp, Jp);
      ld_rec(rec + i * 16, gGR, gGt);
      const float4 jacc = reinterpret_cast<const float4*>(rec + i * 16)[3];
      const m3 M = matmul_AtB(GRp, GRi);   // R_i
      f3 tl = mk3(Ji.x - Jp.x, Ji.y - Jp.y, Ji.z - Jp.z);
      if (a.bt) {
        const float* bp = a.bt + (long long)bs_of(s) * a.bt_stride + i * 3;
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
      if (a.g_bt && live(s)) st3(a.g_bt + ((long long)(b0 + s) * J + i) * 3, mk3(gtl.x, -gtl.y, gtl.z));
    }
    __syncwarp();
    // parents (level lev-1) gather what their children just wrote; 4 float4 per record
    const int pbeg = m.level_ptr[lev - 1], pwid = m.level_ptr[lev] - pbeg;
    for (int t = l; t < SPW * pwid * 4; t += 32) {
      int s, u;
      split_task<SPW>(t, pwid * 4, s, u);
      const int pi = u >> 2, q4 = u & 3;
      const int p = m.order[pbeg + pi];
      const int c0 = m.child_ptr[p], c1 = m.child_ptr[p + 1];
      if (c0 == c1) continue;
      float* rec = P2_REC(s);
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
  if (l < SPW) {   // root: G_0 = (R_0 | J_0)
    const int s = l;
    m3 gGR; f3 gGt;
    ld_rec(P2_REC(s), gGR, gGt);
    const float4 jacc = reinterpret_cast<const float4*>(P2_REC(s))[3];
    st_rec(P2_G(s), gGR, mk3(jacc.x + gGt.x, jacc.y + gGt.y, jacc.z + gGt.z));
    if (a.g_bt && live(s)) st3(a.g_bt + (long long)(b0 + s) * J * 3, mk3(0.f, 0.f, 0.f));
  }
  __syncwarp();
  // ---- lane-per-(sample, joint) epilogue: g_theta --------------------------------------------------------------------------
  if (a.g_theta) {
    for (int t = l; t < SPW * J; t += 32) {
      int s, j;
      split_task<SPW>(t, J, s, j);
      if (!live(s)) continue;
      const int b = b0 + s;
      m3 gR; f3 dJ;
      ld_rec(P2_G(s) + j * 12, gR, dJ);
      if (a.g_Rs) {
        const float* gp = a.g_Rs + ((long long)b * J + j) * 9;
#pragma unroll
        for (int q = 0; q < 9; ++q) gR.a[q] += gp[q];
      }
      if (rotmat) {
        float* o = a.g_theta + ((long long)b * J + j) * 9;
#pragma unroll
        for (int q = 0; q < 9; ++q) o[q] = gR.a[q];
      } else {
        st3(a.g_theta + ((long long)b * J + j) * 3, rodrigues_vjp(ld3(a.theta + (long long)b * a.theta_stride + j * 3), gR));
      }
    }
  }
  // ---- g_beta, g_trans ---------------------------------------------------------------------------------------------------------
  const int ne = nb + 3;
  if (a.g_beta && a.jrest_closed) {   // rest joints were JT + beta JS: dL/dbeta_k += JS_k . dL/dJrest
    for (int e = l; e < 3 * J; e += 32) {
      const int j = e / 3, c = e - 3 * j;
      float gje[SPW];
#pragma unroll
      for (int s = 0; s < SPW; ++s) gje[s] = P2_G(s)[j * 12 + 3 + 4 * c];
#pragma unroll
      for (int k = 0; k < NBMAX; ++k)
        if (k < nb) {
          const float js = m.JS[(long long)k * 3 * J + e];
#pragma unroll
          for (int s = 0; s < SPW; ++s) gbeta[s][k] = fmaf(js, gje[s], gbeta[s][k]);
        }
    }
  }
#pragma unroll
  for (int s = 0; s < SPW; ++s) {
    float part = 0.f;   // lane e < nb+3: sum over the tiles of the skin pass' partial (g_beta_e | g_trans)
    if (a.gbt_part && l < ne)
      for (int t = 0; t < a.ntile; ++t) part += a.gbt_part[((long long)bs_of(s) * a.ntile + t) * ne + l];
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < NBMAX; ++k) {
      if (k < nb) {
        float v = gbeta[s][k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (l == k) mine = v;
      }
    }
    float g0 = gtr[s][0], g1 = gtr[s][1], g2 = gtr[s][2];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      g0 += __shfl_xor_sync(0xffffffffu, g0, o);
      g1 += __shfl_xor_sync(0xffffffffu, g1, o);
      g2 += __shfl_xor_sync(0xffffffffu, g2, o);
    }
    if (!live(s)) continue;
    if (a.g_beta && l < nb) a.g_beta[(long long)(b0 + s) * nb + l] = mine + part;
    if (a.g_trans && l >= nb && l < ne) {
      const int c = l - nb;
      a.g_trans[(long long)(b0 + s) * 3 + c] = part + (c == 0 ? g0 : (c == 1 ? g1 : g2));
    }
  }
#undef P2_G
#undef P2_REC
#undef P2_BE
#undef P2_Q
}

}  // namespace smil
