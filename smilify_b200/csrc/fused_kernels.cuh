// Fused blendshape + linear-blend-skinning passes (the HBM-bound part of the path).
//
// Forward  (skin_fwd_kernel):  thread per vertex, per-vertex constants (template, shapedirs,
//   skinning pairs) live in REGISTERS and are reused across a group of samples; the group's bone
//   transforms arrive in shared memory through one TMA bulk copy; results are staged in shared
//   memory and leave through TMA bulk stores, so every byte that crosses HBM moves in 16-byte
//   aligned bulk transactions.  Replaces /root/reference/smal_model/smal_torch.py:223-253, 318-340.
//
// Backward (skin_bwd_kernel):  item-major.  An item is one non-zero (vertex, bone, weight) of
//   the skinning matrix; items are grouped per vertex tile and bone ("unit").  A lane owns one
//   SAMPLE (8 samples x 4 item streams per warp), so every reduction over vertices -- g_A per
//   bone, g_beta, g_trans -- is a private register accumulation: no atomics, no cross-lane
//   reduction per item, deterministic.  g_verts tiles arrive through double-buffered TMA bulk
//   copies.  Autograd of smal_torch.py:223-253, 318-340.
#pragma once
#include "smil_common.cuh"

namespace smil {

// ---- PTX: mbarrier + 1-D bulk copies (TMA) --------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared, completion signalled on an mbarrier (bytes multiple of 16, both sides 16B aligned)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global, bulk-group completion
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// =====================================================================================================
// forward
// =====================================================================================================
constexpr int SKIN_FWD_THREADS = 256;  // == packer.FWD_CHUNK
constexpr int SKIN_FWD_GMAX = 16;

struct SkinFwdArgs {
  int B, nb, G;               // G samples per CTA
  const float* beta;  long long beta_stride;
  const float* trans; long long trans_stride;   // or null
  const float* A;             // [B][J][12]
  float* verts;               // [B][V][3]
  float* v_shaped;            // [B][V][3]
};

__host__ __device__ inline int skin_fwd_stage_floats(int chunk) { return ((3 * chunk + 4) + 3) & ~3; }
// beta block padded to a 16-byte multiple: the staging tiles behind it are bulk-store sources
__host__ __device__ inline int skin_fwd_beta_floats(int G, int NBT) { return (G * NBT + 3) & ~3; }
__host__ inline size_t skin_fwd_smem_bytes(int G, int J, int NBT, int chunk) {
  return 16 + (size_t)G * J * 48 + (size_t)skin_fwd_beta_floats(G, NBT) * 4 + (size_t)G * 16 +
         4ull * skin_fwd_stage_floats(chunk) * 4;
}

// Writes one staged tile (n floats starting `phase` floats into `stage`) to global `dst` (+off floats):
// 16B-aligned body through one bulk store issued by thread 0, <=3-float head and tail with scalar stores.
__device__ __forceinline__ void store_tile(float* __restrict__ dst, long long off, const float* stage, int phase, int n,
                                           int tid) {
  const int head = min(n, (4 - phase) & 3);
  const int body = ((n - head) >> 2) << 2;
  if (tid == 0 && body > 0) bulk_s2g(dst + off + head, stage + phase + head, (uint32_t)body * 4u);
  if (tid >= 32 && tid < 32 + head) dst[off + (tid - 32)] = stage[phase + (tid - 32)];
  const int tail = n - head - body;
  if (tid >= 64 && tid < 64 + tail) dst[off + head + body + (tid - 64)] = stage[phase + head + body + (tid - 64)];
}

template <int NBT, int K>
__global__ void __launch_bounds__(SKIN_FWD_THREADS) skin_fwd_kernel(DevModel m, SkinFwdArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x;
  const int chunk = blockIdx.x;
  const int b0 = blockIdx.y * a.G;
  const int Gn = min(a.G, a.B - b0);
  const int J = m.J, V = m.V, C = SKIN_FWD_THREADS;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  float* sA = reinterpret_cast<float*>(smem_raw + 16);
  float* sBeta = sA + (size_t)a.G * J * 12;
  float* sTrans = sBeta + skin_fwd_beta_floats(a.G, NBT);
  float* sStage = sTrans + (size_t)a.G * 4;
  const int SF = skin_fwd_stage_floats(C);

  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t bytes = (uint32_t)Gn * J * 48u;
    mbar_expect_tx(bar, bytes);
    bulk_g2s(sA, a.A + (long long)b0 * J * 12, bytes, bar);
  }
  // per-sample scalars of the group
  for (int i = tid; i < Gn * NBT; i += C) {
    const int g = i / NBT, k = i - g * NBT;
    sBeta[i] = (k < a.nb) ? a.beta[(long long)(b0 + g) * a.beta_stride + k] : 0.f;
  }
  for (int i = tid; i < Gn * 4; i += C) {
    const int g = i >> 2, c = i & 3;
    sTrans[i] = (a.trans && c < 3) ? a.trans[(long long)(b0 + g) * a.trans_stride + c] : 0.f;
  }
  // per-vertex constants -> registers (coalesced SoA loads; padded vertices read zeros)
  const int v = chunk * C + tid;
  float T[3], S[NBT][3], w[K];
  int bo[K];
#pragma unroll
  for (int c = 0; c < 3; ++c) T[c] = m.vconst[(long long)c * m.v_pad + v];
#pragma unroll
  for (int k = 0; k < NBT; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) S[k][c] = (k < m.nB) ? m.vconst[(long long)(3 + 3 * k + c) * m.v_pad + v] : 0.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    w[k] = (k < m.K) ? m.ell_w[(long long)k * m.v_pad + v] : 0.f;
    bo[k] = (k < m.K) ? m.ell_idx[(long long)k * m.v_pad + v] * 12 : 0;
  }
  const int nv = min(C, V - chunk * C);
  const int n = 3 * nv;
  __syncthreads();
  mbar_wait(bar, 0);

  for (int g = 0; g < Gn; ++g) {
    float* stV = sStage + (size_t)((g & 1) * 2 + 0) * SF;
    float* stS = sStage + (size_t)((g & 1) * 2 + 1) * SF;
    const long long off = ((long long)(b0 + g) * V + (long long)chunk * C) * 3;
    const int phase = (int)(off & 3);
    // shape blend
    float vs[3] = {T[0], T[1], T[2]};
    const float* bp = sBeta + g * NBT;
#pragma unroll
    for (int k = 0; k < NBT; ++k) {
      const float bk = bp[k];
      vs[0] = fmaf(bk, S[k][0], vs[0]); vs[1] = fmaf(bk, S[k][1], vs[1]); vs[2] = fmaf(bk, S[k][2], vs[2]);
    }
    // skinning: sum_k w_k (A_k [vs;1])
    float o[3] = {sTrans[g * 4 + 0], sTrans[g * 4 + 1], sTrans[g * 4 + 2]};
    const float* Ag = sA + (size_t)g * J * 12;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float4* Ak = reinterpret_cast<const float4*>(Ag + bo[k]);
      const float4 r0 = Ak[0], r1 = Ak[1], r2 = Ak[2];
      o[0] = fmaf(w[k], fmaf(r0.x, vs[0], fmaf(r0.y, vs[1], fmaf(r0.z, vs[2], r0.w))), o[0]);
      o[1] = fmaf(w[k], fmaf(r1.x, vs[0], fmaf(r1.y, vs[1], fmaf(r1.z, vs[2], r1.w))), o[1]);
      o[2] = fmaf(w[k], fmaf(r2.x, vs[0], fmaf(r2.y, vs[1], fmaf(r2.z, vs[2], r2.w))), o[2]);
    }
    if (tid < nv) {
      stV[phase + 3 * tid + 0] = o[0]; stV[phase + 3 * tid + 1] = o[1]; stV[phase + 3 * tid + 2] = o[2];
      stS[phase + 3 * tid + 0] = vs[0]; stS[phase + 3 * tid + 1] = vs[1]; stS[phase + 3 * tid + 2] = vs[2];
    }
    fence_proxy_async();
    if (tid == 0) bulk_wait_read0();  // stores of sample g-1 have drained their staging buffer
    __syncthreads();
    store_tile(a.verts, off, stV, phase, n, tid);
    store_tile(a.v_shaped, off, stS, phase, n, tid);
    if (tid == 0) bulk_commit();
  }
  if (tid == 0) bulk_wait_read0();
}

// =====================================================================================================
// backward
// =====================================================================================================
constexpr int SKIN_BWD_G = 8;       // samples per warp pass (lane & 7)
constexpr int SKIN_BWD_STREAMS = 4; // item streams per warp (lane >> 3)

struct SkinBwdArgs {
  int B, nb;
  int groups_per_cta;        // consecutive sample groups processed by one CTA
  int max_tile_items;        // smem sizing
  const float* beta;  long long beta_stride;
  const float* A;            // [B][J][12]
  const float* g_verts;      // [B][V][3]
  long long total_floats;    // B*V*3
  float* gA_slots;           // [B][num_units][12]
  float* gbt_part;           // [B][ntile][nb+3]
};

__host__ __device__ inline int skin_bwd_row_floats(int chunk) {
  int r = 3 * chunk + 4;             // +4: alignment slack for the 16-byte superset load
  return r + ((36 - (r & 31)) & 31);  // == 4 (mod 32): the 8 sample rows of a stream hit distinct banks
}
__host__ inline size_t skin_bwd_smem_bytes(int chunk, int rec, int max_items, int NBT) {
  size_t fl = (size_t)chunk * rec + 2ull * SKIN_BWD_G * skin_bwd_row_floats(chunk) + 3ull * ((max_items + 3) & ~3) +
              4ull * SKIN_BWD_G * (NBT + 4);
  return 32 + fl * 4;
}

template <int NBT>
__global__ void __launch_bounds__(128) skin_bwd_kernel(DevModel m, SkinBwdArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane & (SKIN_BWD_G - 1), st = lane >> 3;
  const int ti = blockIdx.x;
  const int C = m.bwd_chunk, V = m.V, J = m.J, rec = m.bwd_rec;
  const int RF = skin_bwd_row_floats(C);
  const int MI = (a.max_tile_items + 3) & ~3;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [2]
  float* sRec = reinterpret_cast<float*>(smem_raw + 32);   // [C][rec]
  float* sRows = sRec + (size_t)C * rec;                   // [2][G][RF]
  float* sItW = sRows + 2ull * SKIN_BWD_G * RF;            // [MI]
  float* sItF = sItW + MI;                                 // [MI]
  int* sItV = reinterpret_cast<int*>(sItF + MI);           // [MI]
  float* sRed = reinterpret_cast<float*>(sItV + MI);       // [4 warps][G][NBT+4]

  const int u_beg = m.tile_unit_ptr[ti], u_end = m.tile_unit_ptr[ti + 1];
  const int it_base = m.unit_ptr[u_beg], it_cnt = m.unit_ptr[u_end] - it_base;
  const int nv = min(C, V - ti * C);
  const int n = 3 * nv;
  const int ngroups = (a.B + SKIN_BWD_G - 1) / SKIN_BWD_G;
  const int grp0 = blockIdx.y * a.groups_per_cta;
  const int grp1 = min(ngroups, grp0 + a.groups_per_cta);

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  // tile constants -> shared memory (once per CTA, reused by every sample group)
  {
    const float4* src = reinterpret_cast<const float4*>(m.vrec + (long long)ti * C * rec);
    float4* dst = reinterpret_cast<float4*>(sRec);
    for (int i = tid; i < C * rec / 4; i += 128) dst[i] = src[i];
    for (int i = tid; i < it_cnt; i += 128) {
      sItV[i] = m.it_v[it_base + i];
      sItW[i] = m.it_w[it_base + i];
      sItF[i] = m.it_first[it_base + i];
    }
  }
  __syncthreads();

  // producer: one bulk copy per sample row (16-byte aligned superset of the tile's float range)
  auto issue = [&](int grp, int stage) {
    const int b0 = grp * SKIN_BWD_G;
    const int Gn = min(SKIN_BWD_G, a.B - b0);
    uint32_t total = 0;
    for (int r = 0; r < Gn; ++r) {
      const long long off = ((long long)(b0 + r) * V + (long long)ti * C) * 3;
      const int ph = (int)(off & 3);
      long long len = (ph + n + 3) & ~3;
      if (off - ph + len > a.total_floats) len -= 4;  // never read past the buffer; tail fixed up below
      total += (uint32_t)len * 4u;
    }
    mbar_expect_tx(&bars[stage], total);
    for (int r = 0; r < Gn; ++r) {
      const long long off = ((long long)(b0 + r) * V + (long long)ti * C) * 3;
      const int ph = (int)(off & 3);
      long long len = (ph + n + 3) & ~3;
      if (off - ph + len > a.total_floats) len -= 4;
      if (len > 0)
        bulk_g2s(sRows + ((size_t)stage * SKIN_BWD_G + r) * RF, a.g_verts + (off - ph), (uint32_t)len * 4u, &bars[stage]);
    }
  };
  if (tid == 0) {
    if (grp0 < grp1) issue(grp0, 0);
    if (grp0 + 1 < grp1) issue(grp0 + 1, 1);
  }

  const int wu_beg = m.warp_unit_ptr[ti * m.bwd_warps + warp], wu_end = m.warp_unit_ptr[ti * m.bwd_warps + warp + 1];

  for (int grp = grp0; grp < grp1; ++grp) {
    const int stage = (grp - grp0) & 1;
    const uint32_t parity = ((grp - grp0) >> 1) & 1;
    const int b0 = grp * SKIN_BWD_G;
    const int b = b0 + g;
    const bool valid = b < a.B;
    const int bb = valid ? b : a.B - 1;
    mbar_wait(&bars[stage], parity);
    // the (at most 3) floats a clipped superset load left out: plain loads by the owning lanes
    {
      const long long off = ((long long)bb * V + (long long)ti * C) * 3;
      const int ph = (int)(off & 3);
      const long long len = (ph + n + 3) & ~3;
      if (valid && off - ph + len > a.total_floats && warp == 0 && st == 0) {
        float* row = sRows + ((size_t)stage * SKIN_BWD_G + g) * RF;
        for (long long e = len - 4; e < ph + n; ++e) row[e] = a.g_verts[off - ph + e];
      }
    }
    __syncthreads();

    float beta[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) beta[k] = (k < a.nb) ? a.beta[(long long)bb * a.beta_stride + k] : 0.f;
    float gb[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) gb[k] = 0.f;
    float gtx = 0.f, gty = 0.f, gtz = 0.f;
    const long long off = ((long long)bb * V + (long long)ti * C) * 3;
    const float* row = sRows + ((size_t)stage * SKIN_BWD_G + g) * RF + (int)(off & 3);

    for (int wu = wu_beg; wu < wu_end; ++wu) {
      const int u = m.warp_unit[wu];
      const int bone = m.unit_bone[u];
      const int i0 = m.unit_ptr[u] - it_base, i1 = m.unit_ptr[u + 1] - it_base;
      const float4* Ak = reinterpret_cast<const float4*>(a.A + ((long long)bb * J + bone) * 12);
      const float4 r0 = Ak[0], r1 = Ak[1], r2 = Ak[2];
      float acc[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) acc[q] = 0.f;
      for (int it = i0 + st; it < i1; it += SKIN_BWD_STREAMS) {
        const int lv = sItV[it];
        const float w = sItW[it], fw = sItF[it];
        const float gx = row[3 * lv], gy = row[3 * lv + 1], gz = row[3 * lv + 2];
        // vertex record {T, S_0.. S_NBT-1} as 16-byte broadcast loads (address is uniform per stream)
        constexpr int RECV = (3 + 3 * NBT + 3) / 4;
        float rc[4 * RECV];
        {
          const float4* rc4 = reinterpret_cast<const float4*>(sRec + (size_t)lv * rec);
#pragma unroll
          for (int q = 0; q < RECV; ++q) {
            const float4 t4 = rc4[q];
            rc[4 * q] = t4.x; rc[4 * q + 1] = t4.y; rc[4 * q + 2] = t4.z; rc[4 * q + 3] = t4.w;
          }
        }
        float vx = rc[0], vy = rc[1], vz = rc[2];
        // u = A_k.R^T g
        const float ux = fmaf(r0.x, gx, fmaf(r1.x, gy, r2.x * gz));
        const float uy = fmaf(r0.y, gx, fmaf(r1.y, gy, r2.y * gz));
        const float uz = fmaf(r0.z, gx, fmaf(r1.z, gy, r2.z * gz));
        const float wux = w * ux, wuy = w * uy, wuz = w * uz;
#pragma unroll
        for (int k = 0; k < NBT; ++k) {
          const float sx = rc[3 + 3 * k], sy = rc[4 + 3 * k], sz = rc[5 + 3 * k];
          vx = fmaf(beta[k], sx, vx); vy = fmaf(beta[k], sy, vy); vz = fmaf(beta[k], sz, vz);
          gb[k] = fmaf(wux, sx, fmaf(wuy, sy, fmaf(wuz, sz, gb[k])));
        }
        const float wx = w * gx, wy = w * gy, wz = w * gz;
        acc[0] = fmaf(wx, vx, acc[0]); acc[1] = fmaf(wx, vy, acc[1]); acc[2] = fmaf(wx, vz, acc[2]); acc[3] += wx;
        acc[4] = fmaf(wy, vx, acc[4]); acc[5] = fmaf(wy, vy, acc[5]); acc[6] = fmaf(wy, vz, acc[6]); acc[7] += wy;
        acc[8] = fmaf(wz, vx, acc[8]); acc[9] = fmaf(wz, vy, acc[9]); acc[10] = fmaf(wz, vz, acc[10]); acc[11] += wz;
        gtx = fmaf(fw, gx, gtx); gty = fmaf(fw, gy, gty); gtz = fmaf(fw, gz, gtz);
      }
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], 8);
        acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], 16);
      }
      if (st == 0 && valid) {
        float4* d = reinterpret_cast<float4*>(a.gA_slots + ((long long)b * m.num_units + m.unit_slot[u]) * 12);
        d[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        d[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        d[2] = make_float4(acc[8], acc[9], acc[10], acc[11]);
      }
    }
    // g_beta / g_trans partials of this tile: streams -> warp -> CTA, fixed order
#pragma unroll
    for (int k = 0; k < NBT; ++k) {
      gb[k] += __shfl_xor_sync(0xffffffffu, gb[k], 8);
      gb[k] += __shfl_xor_sync(0xffffffffu, gb[k], 16);
    }
    gtx += __shfl_xor_sync(0xffffffffu, gtx, 8);  gtx += __shfl_xor_sync(0xffffffffu, gtx, 16);
    gty += __shfl_xor_sync(0xffffffffu, gty, 8);  gty += __shfl_xor_sync(0xffffffffu, gty, 16);
    gtz += __shfl_xor_sync(0xffffffffu, gtz, 8);  gtz += __shfl_xor_sync(0xffffffffu, gtz, 16);
    if (st == 0) {
      float* r = sRed + ((size_t)warp * SKIN_BWD_G + g) * (NBT + 4);
#pragma unroll
      for (int k = 0; k < NBT; ++k) r[k] = gb[k];
      r[NBT] = gtx; r[NBT + 1] = gty; r[NBT + 2] = gtz;
    }
    __syncthreads();  // all warps done with this stage's rows and sRed complete
    if (tid == 0 && grp + 2 < grp1) issue(grp + 2, stage);
    {
      const int ne = a.nb + 3;
      for (int i = tid; i < SKIN_BWD_G * ne; i += 128) {
        const int gg = i / ne, e = i - gg * ne;
        const int src = (e < a.nb) ? e : NBT + (e - a.nb);
        if (b0 + gg < a.B) {
          float s = 0.f;
#pragma unroll
          for (int wv = 0; wv < 4; ++wv) s += sRed[((size_t)wv * SKIN_BWD_G + gg) * (NBT + 4) + src];
          a.gbt_part[((long long)(b0 + gg) * m.bwd_ntile + ti) * ne + e] = s;
        }
      }
    }
    // sRed is rewritten only after the next group's first __syncthreads
  }
}

}  // namespace smil
