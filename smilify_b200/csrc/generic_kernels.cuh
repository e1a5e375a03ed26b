// Unfused kernels: every variant of SMAL.__call__ the fused path does not cover
// (del_v, v_template override, non-zero posedirs, K > 4 weights per vertex, nB > 16, upstream
// gradient on v_shaped).  One thread per (sample, vertex); simple, correct, not the hot path.
//
// Reference:  /root/reference/smal_model/smal_torch.py:223-253 (shape blend), 257-264 / 343-351
// (joint regression), 294-301 (pose blend), 318-340 (LBS).
#pragma once
#include "smil_common.cuh"

namespace smil {

// v_shaped[b,v,c] = vt[v,c] + sum_k beta[b,k] S[k][3v+c] (+ del_v)
__global__ void shape_fwd_kernel(DevModel m, int B, int nb, const float* __restrict__ beta, long long beta_stride,
                                 const float* __restrict__ vt, const float* __restrict__ del_v, long long dv_stride,
                                 float* __restrict__ v_shaped) {
  const long long n = (long long)B * m.V * 3;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int e = (int)(i % (3LL * m.V));
  const long long b = i / (3LL * m.V);
  float acc = vt[e];
  if (del_v) acc += del_v[b * dv_stride + e];
  const float* bp = beta + b * beta_stride;
  for (int k = 0; k < nb; ++k) acc = fmaf(bp[k], m.shapedirs[(long long)k * 3 * m.V + e], acc);
  v_shaped[i] = acc;
}

// out[b,j,:] = sum_e jc_val[e] * src[b, jc_vid[e], :]     (CSC by joint; one warp per (b,j))
__global__ void regress_fwd_kernel(DevModel m, int B, const float* __restrict__ src, float* __restrict__ out) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * m.J) return;
  const int j = (int)(w % m.J);
  const long long b = w / m.J;
  float ax = 0.f, ay = 0.f, az = 0.f;
  for (int e = m.jc_ptr[j] + lane; e < m.jc_ptr[j + 1]; e += 32) {
    const float* p = src + (b * m.V + m.jc_vid[e]) * 3;
    const float v = m.jc_val[e];
    ax = fmaf(v, p[0], ax); ay = fmaf(v, p[1], ay); az = fmaf(v, p[2], az);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ax += __shfl_xor_sync(0xffffffffu, ax, o);
    ay += __shfl_xor_sync(0xffffffffu, ay, o);
    az += __shfl_xor_sync(0xffffffffu, az, o);
  }
  if (lane == 0) { out[w * 3] = ax; out[w * 3 + 1] = ay; out[w * 3 + 2] = az; }
}

// v_posed[b,v,c] = v_shaped[b,v,c] + sum_p (Rs[b,1+p/9,p%9] - I) * posedirs[p][3v+c]
__global__ void posedirs_fwd_kernel(DevModel m, int B, const float* __restrict__ Rs, const float* __restrict__ v_shaped,
                                    float* __restrict__ v_posed) {
  extern __shared__ float feat[];  // [P]
  const long long b = blockIdx.y;
  for (int p = threadIdx.x; p < m.P; p += blockDim.x) {
    const int q = p % 9;
    feat[p] = Rs[(b * m.J + 1 + p / 9) * 9 + q] - ((q == 0 || q == 4 || q == 8) ? 1.f : 0.f);
  }
  __syncthreads();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 3 * m.V) return;
  float acc = v_shaped[b * 3 * m.V + e];
  for (int p = 0; p < m.P; ++p) acc = fmaf(feat[p], m.posedirs[(long long)p * 3 * m.V + e], acc);
  v_posed[b * 3 * m.V + e] = acc;
}

// verts[b,v] = sum_k w_vk (A_k.R v_posed + A_k.t) + trans
__global__ void lbs_fwd_kernel(DevModel m, int B, const float* __restrict__ A, const float* __restrict__ v_posed,
                               const float* __restrict__ trans, long long trans_stride, float* __restrict__ verts) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * m.V) return;
  const int v = (int)(i % m.V);
  const long long b = i / m.V;
  const f3 p = ld3(v_posed + i * 3);
  float ox = 0.f, oy = 0.f, oz = 0.f;
  for (int k = 0; k < m.K; ++k) {
    const float w = m.skin_w[v * m.K + k];
    if (w == 0.f) continue;
    const float* Ak = A + (b * m.J + m.skin_idx[v * m.K + k]) * 12;
    ox = fmaf(w, fmaf(Ak[0], p.x, fmaf(Ak[1], p.y, fmaf(Ak[2], p.z, Ak[3]))), ox);
    oy = fmaf(w, fmaf(Ak[4], p.x, fmaf(Ak[5], p.y, fmaf(Ak[6], p.z, Ak[7]))), oy);
    oz = fmaf(w, fmaf(Ak[8], p.x, fmaf(Ak[9], p.y, fmaf(Ak[10], p.z, Ak[11]))), oz);
  }
  if (trans) { const float* t = trans + b * trans_stride; ox += t[0]; oy += t[1]; oz += t[2]; }
  verts[i * 3] = ox; verts[i * 3 + 1] = oy; verts[i * 3 + 2] = oz;
}

// A[b][j][e] = pp[b/2][(j*12+e)*2 + b%2]: the fused forward saves the bone transforms pair-interleaved; the unfused
// backward (taken when the call carries an upstream gradient on v_shaped) wants them plain.
__global__ void deinterleave_A_kernel(int B, int J, int blk, const float* __restrict__ pp, float* __restrict__ A) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = 12LL * J;
  if (i >= (long long)B * per) return;
  const long long b = i / per, e = i - b * per;
  A[i] = pp[(b >> 1) * blk + 2 * e + (e / 12) * (PPB - 24) + (b & 1)];
}

// ---- backward -------------------------------------------------------------------------------------
// g = g_verts[b,v] (+ sum_e jr_val * g_joints[b, jr_jid]);  g_vp = sum_k w A_k.R^T g;
// gA[b,k] += w g (x) [v_posed;1]  (atomics);  g_trans[b] += g (atomics, block-reduced first)
__global__ void lbs_bwd_kernel(DevModel m, int B, const float* __restrict__ A, const float* __restrict__ v_posed,
                               const float* __restrict__ g_verts, const float* __restrict__ g_joints_regress,
                               float* __restrict__ g_vp, float* __restrict__ gA, float* __restrict__ g_trans) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * m.V) return;
  const int v = (int)(i % m.V);
  const long long b = i / m.V;
  f3 g = mk3(0.f, 0.f, 0.f);
  if (g_verts) g = ld3(g_verts + i * 3);
  if (g_joints_regress) {
    for (int e = m.jr_ptr[v]; e < m.jr_ptr[v + 1]; ++e) {
      const float* gj = g_joints_regress + (b * m.J + m.jr_jid[e]) * 3;
      const float w = m.jr_val[e];
      g.x = fmaf(w, gj[0], g.x); g.y = fmaf(w, gj[1], g.y); g.z = fmaf(w, gj[2], g.z);
    }
  }
  const f3 p = ld3(v_posed + i * 3);
  f3 o = mk3(0.f, 0.f, 0.f);
  for (int k = 0; k < m.K; ++k) {
    const float w = m.skin_w[v * m.K + k];
    if (w == 0.f) continue;
    const int bone = m.skin_idx[v * m.K + k];
    const float* Ak = A + (b * m.J + bone) * 12;
    o.x = fmaf(w, fmaf(Ak[0], g.x, fmaf(Ak[4], g.y, Ak[8] * g.z)), o.x);
    o.y = fmaf(w, fmaf(Ak[1], g.x, fmaf(Ak[5], g.y, Ak[9] * g.z)), o.y);
    o.z = fmaf(w, fmaf(Ak[2], g.x, fmaf(Ak[6], g.y, Ak[10] * g.z)), o.z);
    float* d = gA + (b * m.J + bone) * 12;
    const float wx = w * g.x, wy = w * g.y, wz = w * g.z;
    atomicAdd(d + 0, wx * p.x); atomicAdd(d + 1, wx * p.y); atomicAdd(d + 2, wx * p.z); atomicAdd(d + 3, wx);
    atomicAdd(d + 4, wy * p.x); atomicAdd(d + 5, wy * p.y); atomicAdd(d + 6, wy * p.z); atomicAdd(d + 7, wy);
    atomicAdd(d + 8, wz * p.x); atomicAdd(d + 9, wz * p.y); atomicAdd(d + 10, wz * p.z); atomicAdd(d + 11, wz);
  }
  st3(g_vp + i * 3, o);
  if (g_trans) {
    atomicAdd(g_trans + b * 3 + 0, g.x); atomicAdd(g_trans + b * 3 + 1, g.y); atomicAdd(g_trans + b * 3 + 2, g.z);
  }
}

// g_feat[b,p] = sum_e posedirs[p][e] g_vp[b,e]  -> g_Rs2[b, 1+p/9, p%9]   (one warp per (b,p))
__global__ void posedirs_bwd_kernel(DevModel m, int B, const float* __restrict__ g_vp, float* __restrict__ g_Rs2) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * m.P) return;
  const int p = (int)(w % m.P);
  const long long b = w / m.P;
  const float* row = m.posedirs + (long long)p * 3 * m.V;
  const float* g = g_vp + b * 3 * m.V;
  float acc = 0.f;
  for (int e = lane; e < 3 * m.V; e += 32) acc = fmaf(row[e], g[e], acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) g_Rs2[(b * m.J + 1 + p / 9) * 9 + p % 9] = acc;
}

// g_vs[b,v] = g_vp + g_v_shaped_upstream + sum_e jr_val * g_jrest[b, jr_jid]   (in place on g_vp)
__global__ void vshaped_grad_kernel(DevModel m, int B, float* __restrict__ g_vp, const float* __restrict__ g_vs_up,
                                    const float* __restrict__ g_jrest, float* __restrict__ g_del_v) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)B * m.V) return;
  const int v = (int)(i % m.V);
  const long long b = i / m.V;
  f3 g = ld3(g_vp + i * 3);
  if (g_vs_up) { const f3 u = ld3(g_vs_up + i * 3); g.x += u.x; g.y += u.y; g.z += u.z; }
  if (g_jrest) {
    for (int e = m.jr_ptr[v]; e < m.jr_ptr[v + 1]; ++e) {
      const float* gj = g_jrest + (b * m.J + m.jr_jid[e]) * 3;
      const float w = m.jr_val[e];
      g.x = fmaf(w, gj[0], g.x); g.y = fmaf(w, gj[1], g.y); g.z = fmaf(w, gj[2], g.z);
    }
  }
  st3(g_vp + i * 3, g);
  if (g_del_v) st3(g_del_v + i * 3, g);
}

// g_beta[b,k] = sum_e S[k][e] g_vs[b,e]   (one warp per (b,k))
__global__ void shape_bwd_kernel(DevModel m, int B, int nb, const float* __restrict__ g_vs, float* __restrict__ g_beta) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)B * nb) return;
  const int k = (int)(w % nb);
  const long long b = w / nb;
  const float* row = m.shapedirs + (long long)k * 3 * m.V;
  const float* g = g_vs + b * 3 * m.V;
  float acc = 0.f;
  for (int e = lane; e < 3 * m.V; e += 32) acc = fmaf(row[e], g[e], acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) g_beta[w] = acc;
}

}  // namespace smil
