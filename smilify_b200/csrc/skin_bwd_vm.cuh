// Fused backward of shape blend + linear-blend skinning, VERTEX-MAJOR (rigs with <= 2 skinning weights per vertex).
//
// Autograd of /root/reference/smal_model/smal_torch.py:223-253 (shape blend) and :318-340 (LBS) with respect to the
// bone transforms A (-> per-(tile, bone) partial slots), beta and trans (-> per-tile partials).
//
// Against the item-major pass (fused_kernels.cuh::skin_bwd_kernel, kept for K > 2):
//   * per VERTEX, not per (vertex, bone) item, the pass evaluates v_shaped = T + S beta, u = sum_k w_k R_k^T g and the
//     dL/dbeta contraction S^T u -- 78 FMA for a two-bone vertex instead of 114, 54 for a single-bone one;
//   * a warp keeps TWO bones resident (slots X and Y: rotation + 3x4 accumulator each), the vertices of a tile are
//     ordered by the packer so that consecutive runs share a bone (packer.py::_vm_order_runs): a run switch replaces
//     one slot, the other keeps accumulating;
//   * the two slots are the two halves of f32x2 registers: acc, R and the weights are (X, Y) pairs, the per-sample
//     gradient g and v_shaped enter the packed FMAs as broadcast scalars (FFMA2 takes a 32-bit operand for both
//     halves), so a two-bone vertex costs 24 packed + 33 scalar FMAs; runs that touch one slot only take a scalar path;
//   * records are per vertex (24 floats for nB = 5 instead of 44 per item pair): less shared memory per CTA, 4 CTAs/SM.
// Unchanged: lane = sample (16 samples x 2 vertex streams per warp), g_verts tiles arrive by TMA bulk copies (one per
// sample row, 16-byte aligned superset), every reduction over vertices is a private register accumulation, slots and
// per-tile partials are written without atomics in a fixed order (deterministic).
#pragma once
#include "fused_kernels.cuh"

namespace smil {

constexpr int VM_WARPS = 4, VM_PHASES = 1;   // what the kernel is compiled for (the packer can emit other splits)
constexpr int VM_MODE_PACKED = 0, VM_MODE_X = 1;   // == packer.VM_MODE_* (Y-only stretches are packed with w_X = 0)

struct SkinBwdVmArgs {
  int B, nb;
  int max_tile_recs, max_tile_segs;   // smem sizing
  const float* beta;  long long beta_stride;
  const float* pp; int blk;           // per-pair parameter blocks of the forward (A2 [J][PPB] first), blk floats each
  const float* g_verts;               // [B][V][3]
  long long total_floats;             // B*V*3
  float* gA_slots;                    // [nslots][B][12]   (slot-major: a flush writes 16 samples x 48 contiguous bytes)
  float* gbt_part;                    // [B][ntile][nb+3]
  // atomic mode (gA_dense != null): a flush ADDS its 3x4 partial to the dense dL/dA [B][J][12] with three 16-byte
  // vector reductions (red.global.add.v4.f32), the per-tile (g_beta, g_trans) partials are added to gbt_sum [B][nb+3];
  // both buffers are zeroed by the caller.  No slot traffic, no slot-summing pass; the summation order over the tiles
  // is then the hardware's (results vary in the last bits from run to run, like any atomically reduced torch op)
  float* gA_dense;
  float* gbt_sum;
  int J;
  int stagger_ns;                     // start delay per resident-CTA slot (see the kernel)
};

__host__ __device__ inline int vm_rec_floats(int NBT) { return (6 + 3 * NBT + 3) & ~3; }
__device__ __forceinline__ void red_add_v4(float* p, float x, float y, float z, float w) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__host__ inline size_t skin_bwd_vm_smem_bytes(int chunk, int nphase, int nwarps, int recf, int max_recs, int max_segs, int NBT) {
  size_t fl = (size_t)max_recs * recf + (size_t)nphase * SKIN_BWD_G * skin_bwd_row_floats(chunk / nphase) +
              (size_t)nwarps * SKIN_BWD_G * (NBT + 4) + 8ull * max_segs;
  return 32 + fl * 4;
}

template <int NBT>
__global__ void __launch_bounds__(128) skin_bwd_vm_kernel(DevModel m, SkinBwdVmArgs a) {
  // Warps per CTA that share a tile, and phases (sub-ranges of a tile with their own row buffer) are COMPILE-TIME here:
  // 1 / 2 warps per tile and 2 / 3 phases were measured slower on B200 (107-135 us and 116-126 us
  // against 95-97 us), and as run-time values they cost the default configuration 5 us.
  constexpr int NW = VM_WARPS;
  static_assert(SKIN_BWD_STREAMS == 2, "the vertex-major pass is written for 16 samples x 2 vertex streams per warp");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int RECF = (6 + 3 * NBT + 3) & ~3;
  constexpr int NQ = RECF / 4;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane & (SKIN_BWD_G - 1), st = lane / SKIN_BWD_G;
  const int ti = blockIdx.x;
  constexpr int NPH = VM_PHASES;
  const int C = m.vm_chunk, V = m.V;
  const int PW = C / NPH;                       // vertices per phase (a multiple of 4)
  const int RF = skin_bwd_row_floats(PW);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [NPH <= 4]
  float* sRec = reinterpret_cast<float*>(smem_raw + 32);   // [max_recs][RECF]
  float* sRows = sRec + (size_t)a.max_tile_recs * RECF;    // [NPH][G][RF]: one buffer per phase
  float* sRed = sRows + (size_t)NPH * SKIN_BWD_G * RF;     // [4 warps][G][NBT+4]
  int4* sSeg = reinterpret_cast<int4*>(sRed + (size_t)NW * SKIN_BWD_G * (NBT + 4));   // [max_segs][2]

  const int rec_base = m.vm_tile_rec_ptr[ti], rec_cnt = m.vm_tile_rec_ptr[ti + 1] - rec_base;
  const int* wsp = m.vm_warp_seg_ptr + (long long)ti * NPH * NW;     // [(phase * NW + warp)], tile-major
  const int seg_base = wsp[0], seg_cnt = wsp[NPH * NW] - seg_base;
  const int ngroups = (a.B + SKIN_BWD_G - 1) / SKIN_BWD_G;
  const int gstep = gridDim.y;   // persistent over sample groups: blockIdx.y, blockIdx.y + gridDim.y, ...

  if (tid == 0) {
    for (int p = 0; p < NPH; ++p) mbar_init(&bars[p], 1);
    fence_mbar_init();
  }
  {   // tile constants -> shared memory (once per CTA, reused by every sample group)
    const float4* src = reinterpret_cast<const float4*>(m.vm_rec + (long long)rec_base * RECF);
    float4* dst = reinterpret_cast<float4*>(sRec);
    for (int i = tid; i < rec_cnt * NQ; i += 32 * NW) dst[i] = src[i];
    const int4* ssrc = m.vm_seg + 2ll * seg_base;
    for (int i = tid; i < 2 * seg_cnt; i += 32 * NW) {
      int4 d = ssrc[i];
      if ((i & 1) == 0) {   // bones -> float offsets into a pair block
        d.x = d.x >= 0 ? d.x * PPB : -1;
        d.y = d.y >= 0 ? d.y * PPB : -1;
      } else if (a.gA_dense) {   // atomic mode: a flush targets its BONE's dense entry instead of a private slot
        d.x = d.x >= 0 ? m.vm_slot_bone[d.x] : -1;
        d.y = d.y >= 0 ? m.vm_slot_bone[d.y] : -1;
      }
      sSeg[i] = d;
    }
  }
  __syncthreads();
  // the warp's segments: phase p owns [wsp[NW p + warp], wsp[NW p + warp + 1]); the warp walks them phase by phase
  const int first_seg = [&]() { for (int p = 0; p < NPH; ++p) if (wsp[NW * p + warp] < wsp[NW * p + warp + 1]) return wsp[NW * p + warp] - seg_base; return -1; }();
  // producer (warp 0, all lanes): lane r issues the bulk copy of sample row r of one phase (16-byte aligned superset of
  // the phase's float range); the byte total is folded with one warp reduction before lane 0 arms the barrier
  // a tile is the vertex range [rec_base, rec_base + rec_cnt): one record per vertex, tiles consecutive; the packer
  // sizes the tiles for equal estimated COST (not equal vertex counts), C = vm_chunk is only their upper bound
  auto phase_n = [&](int p) { return 3 * max(0, min(PW, rec_cnt - p * PW)); };
  auto issue = [&](int grp, int p) {
    const int b0 = grp * SKIN_BWD_G;
    const int Gn = min(SKIN_BWD_G, a.B - b0);
    const int n = phase_n(p);
    const bool act = lane < Gn && n > 0;
    const long long off = ((long long)(b0 + lane) * V + (long long)rec_base + (long long)p * PW) * 3;
    const int ph = (int)(off & 3);
    long long len = (ph + n + 3) & ~3;
    if (off - ph + len > a.total_floats) len -= 4;  // never read past the buffer; tail fixed up below
    const uint32_t bytes = (act && len > 0) ? (uint32_t)len * 4u : 0u;
    const uint32_t total = __reduce_add_sync(0xffffffffu, bytes);
    if (lane == 0) mbar_expect_tx(&bars[p], total);
    __syncwarp();
    if (bytes) bulk_g2s(sRows + ((size_t)p * SKIN_BWD_G + lane) * RF, a.g_verts + (off - ph), bytes, &bars[p]);
  };
  if (a.stagger_ns > 0) {
    const unsigned slot = ((blockIdx.y * gridDim.x + blockIdx.x) / (unsigned)m.sm_count) & 7u;
    if (slot) __nanosleep(slot * (unsigned)a.stagger_ns);
  }
  if (warp == 0 && (int)blockIdx.y < ngroups)
    for (int p = 0; p < NPH; ++p) issue(blockIdx.y, p);

  float beta_next[NBT];   // betas of the group's samples, fetched one group ahead
  auto beta_fetch = [&](int grp) {
    const int bs = min(grp * SKIN_BWD_G + g, a.B - 1);
#pragma unroll
    for (int k = 0; k < NBT; ++k) beta_next[k] = (k < a.nb) ? a.beta[(long long)bs * a.beta_stride + k] : 0.f;
  };
  if ((int)blockIdx.y < ngroups) beta_fetch(blockIdx.y);

  int itn = 0;
  for (int grp = blockIdx.y; grp < ngroups; grp += gstep, ++itn) {
    const uint32_t parity = itn & 1;     // every phase buffer is filled once per group
    const int b0 = grp * SKIN_BWD_G;
    const int b = b0 + g;
    const bool valid = b < a.B;
    const int bb = valid ? b : a.B - 1;
    // rotations of the warp's first segment: issued BEFORE the wait for the g_verts rows (they depend on the
    // forward's parameter blocks only), so their L2 latency hides under the TMA's
    const float* Ab = a.pp + (long long)(bb >> 1) * a.blk + (bb & 1);
    float2 R[9];      // (X, Y) halves of the two resident bones' rotations, row-major
    auto load_R = [&](float (&dst)[9], int boff) {
      const float* Ak = Ab + boff;
      dst[0] = Ak[0]; dst[1] = Ak[2]; dst[2] = Ak[4]; dst[3] = Ak[8]; dst[4] = Ak[10]; dst[5] = Ak[12];
      dst[6] = Ak[16]; dst[7] = Ak[18]; dst[8] = Ak[20];
    };
    float Rn[9];      // one slot's rotation for the NEXT segment, fetched under the current segment's vertices
    int pf = -1;      // which slot Rn is for (0 = X, 1 = Y, -1 = nothing pending)
#pragma unroll
    for (int q = 0; q < 9; ++q) { R[q] = make_float2(0.f, 0.f); Rn[q] = 0.f; }
    if (first_seg >= 0) {
      const int4 f0 = sSeg[2 * first_seg];
      float t[9];
      if (f0.x >= 0) {
        load_R(t, f0.x);
#pragma unroll
        for (int q = 0; q < 9; ++q) R[q].x = t[q];
      }
      if (f0.y >= 0) {
        load_R(t, f0.y);
#pragma unroll
        for (int q = 0; q < 9; ++q) R[q].y = t[q];
      }
    }
    float beta[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) beta[k] = beta_next[k];
    if (grp + gstep < ngroups) beta_fetch(grp + gstep);
    float gb[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) gb[k] = 0.f;
    float gt0 = 0.f, gt1 = 0.f, gt2 = 0.f;
    float2 acc[12];   // (slot X, slot Y) halves of the two resident bones' dL/dA accumulators
#pragma unroll
    for (int q = 0; q < 12; ++q) acc[q] = make_float2(0.f, 0.f);

    for (int p = 0; p < NPH; ++p) {
      const int n = phase_n(p);
      const long long off = ((long long)bb * V + (long long)rec_base + (long long)p * PW) * 3;
      mbar_wait(&bars[p], parity);
      {   // the (at most 3) floats a clipped superset load left out (last sample, last tile, last phase only): plain
          // loads by the owning lanes; the condition is the same for the whole CTA
        const long long offl = ((long long)min(b0 + SKIN_BWD_G - 1, a.B - 1) * V + (long long)rec_base + (long long)p * PW) * 3;
        const int phl = (int)(offl & 3);
        if (n > 0 && offl - phl + ((phl + n + 3) & ~3) > a.total_floats) {
          const int ph = (int)(off & 3);
          const long long len = (ph + n + 3) & ~3;
          if (valid && off - ph + len > a.total_floats && warp == 0 && st == 0) {
            float* row = sRows + ((size_t)p * SKIN_BWD_G + g) * RF;
            for (long long e = len - 4; e < ph + n; ++e) row[e] = a.g_verts[off - ph + e];
          }
          __syncthreads();
        }
      }
      const float* row = sRows + ((size_t)p * SKIN_BWD_G + g) * RF + (int)(off & 3);
      const int sg_beg = wsp[NW * p + warp] - seg_base, sg_end = wsp[NW * p + warp + 1] - seg_base;
      for (int sg = sg_beg; sg < sg_end; ++sg) {
      const int4 d0 = sSeg[2 * sg], d1 = sSeg[2 * sg + 1];
      if (sg != first_seg) {   // a new bone for slot X / Y (its previous content was flushed by the segment before)
        if (d0.x >= 0) {
          if (pf == 0) {
#pragma unroll
            for (int q = 0; q < 9; ++q) R[q].x = Rn[q];
          } else {
            float t[9];
            load_R(t, d0.x);
#pragma unroll
            for (int q = 0; q < 9; ++q) R[q].x = t[q];
          }
        }
        if (d0.y >= 0) {
          if (pf == 1) {
#pragma unroll
            for (int q = 0; q < 9; ++q) R[q].y = Rn[q];
          } else {
            float t[9];
            load_R(t, d0.y);
#pragma unroll
            for (int q = 0; q < 9; ++q) R[q].y = t[q];
          }
        }
      }
      pf = -1;
      if (d1.w >= 0) {   // prefetch one slot of the warp's next segment, possibly of the next phase (a switch replaces one slot almost always)
        const int4 nx = sSeg[2 * d1.w];
        if (nx.x >= 0) { load_R(Rn, nx.x); pf = 0; }
        else if (nx.y >= 0) { load_R(Rn, nx.y); pf = 1; }
      }
      const int rend = d0.z + d0.w;
      if (d1.z == VM_MODE_PACKED) {
        for (int it = d0.z + st; it < rend; it += 2) {
          const float4* rc = reinterpret_cast<const float4*>(sRec + (size_t)it * RECF);
          float c[RECF];
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            const float4 t4 = rc[q];
            c[4 * q] = t4.x; c[4 * q + 1] = t4.y; c[4 * q + 2] = t4.z; c[4 * q + 3] = t4.w;
          }
          const float2 w2 = make_float2(c[0], c[1]);
          const int lv = __float_as_int(c[2]);
          const float gx = row[lv], gy = row[lv + 1], gz = row[lv + 2];
          gt0 += gx; gt1 += gy; gt2 += gz;
          float vx = c[3], vy = c[4], vz = c[5];
#pragma unroll
          for (int k = 0; k < NBT; ++k) {
            vx = fmaf(beta[k], c[6 + 3 * k], vx); vy = fmaf(beta[k], c[7 + 3 * k], vy); vz = fmaf(beta[k], c[8 + 3 * k], vz);
          }
          const float2 wx = __fmul2_rn(w2, dup2(gx)), wy = __fmul2_rn(w2, dup2(gy)), wz = __fmul2_rn(w2, dup2(gz));
          const float2 vx2 = dup2(vx), vy2 = dup2(vy), vz2 = dup2(vz);
          acc[0] = __ffma2_rn(wx, vx2, acc[0]); acc[1] = __ffma2_rn(wx, vy2, acc[1]); acc[2] = __ffma2_rn(wx, vz2, acc[2]);
          acc[3] = __fadd2_rn(acc[3], wx);
          acc[4] = __ffma2_rn(wy, vx2, acc[4]); acc[5] = __ffma2_rn(wy, vy2, acc[5]); acc[6] = __ffma2_rn(wy, vz2, acc[6]);
          acc[7] = __fadd2_rn(acc[7], wy);
          acc[8] = __ffma2_rn(wz, vx2, acc[8]); acc[9] = __ffma2_rn(wz, vy2, acc[9]); acc[10] = __ffma2_rn(wz, vz2, acc[10]);
          acc[11] = __fadd2_rn(acc[11], wz);
          // u = R_X^T (w_X g) + R_Y^T (w_Y g): both bones in one packed chain, halves folded at the end
          const float2 t0 = __ffma2_rn(R[0], wx, __ffma2_rn(R[3], wy, __fmul2_rn(R[6], wz)));
          const float2 t1 = __ffma2_rn(R[1], wx, __ffma2_rn(R[4], wy, __fmul2_rn(R[7], wz)));
          const float2 t2 = __ffma2_rn(R[2], wx, __ffma2_rn(R[5], wy, __fmul2_rn(R[8], wz)));
          const float ux = t0.x + t0.y, uy = t1.x + t1.y, uz = t2.x + t2.y;
#pragma unroll
          for (int k = 0; k < NBT; ++k)
            gb[k] = fmaf(ux, c[6 + 3 * k], fmaf(uy, c[7 + 3 * k], fmaf(uz, c[8 + 3 * k], gb[k])));
        }
      } else {
        // slot X only (the packer sends the rare Y-only stretches through the packed path with w_X = 0): scalar path
        float r[9], sa[12];
#pragma unroll
        for (int q = 0; q < 9; ++q) r[q] = R[q].x;
#pragma unroll
        for (int q = 0; q < 12; ++q) sa[q] = acc[q].x;
        for (int it = d0.z + st; it < rend; it += 2) {
          const float4* rc = reinterpret_cast<const float4*>(sRec + (size_t)it * RECF);
          float c[RECF];
#pragma unroll
          for (int q = 0; q < NQ; ++q) {
            const float4 t4 = rc[q];
            c[4 * q] = t4.x; c[4 * q + 1] = t4.y; c[4 * q + 2] = t4.z; c[4 * q + 3] = t4.w;
          }
          const float w = c[0];
          const int lv = __float_as_int(c[2]);
          const float gx = row[lv], gy = row[lv + 1], gz = row[lv + 2];
          gt0 += gx; gt1 += gy; gt2 += gz;
          float vx = c[3], vy = c[4], vz = c[5];
#pragma unroll
          for (int k = 0; k < NBT; ++k) {
            vx = fmaf(beta[k], c[6 + 3 * k], vx); vy = fmaf(beta[k], c[7 + 3 * k], vy); vz = fmaf(beta[k], c[8 + 3 * k], vz);
          }
          const float wx = w * gx, wy = w * gy, wz = w * gz;
          sa[0] = fmaf(wx, vx, sa[0]); sa[1] = fmaf(wx, vy, sa[1]); sa[2] = fmaf(wx, vz, sa[2]); sa[3] += wx;
          sa[4] = fmaf(wy, vx, sa[4]); sa[5] = fmaf(wy, vy, sa[5]); sa[6] = fmaf(wy, vz, sa[6]); sa[7] += wy;
          sa[8] = fmaf(wz, vx, sa[8]); sa[9] = fmaf(wz, vy, sa[9]); sa[10] = fmaf(wz, vz, sa[10]); sa[11] += wz;
          const float ux = fmaf(r[0], wx, fmaf(r[3], wy, r[6] * wz));
          const float uy = fmaf(r[1], wx, fmaf(r[4], wy, r[7] * wz));
          const float uz = fmaf(r[2], wx, fmaf(r[5], wy, r[8] * wz));
#pragma unroll
          for (int k = 0; k < NBT; ++k)
            gb[k] = fmaf(ux, c[6 + 3 * k], fmaf(uy, c[7 + 3 * k], fmaf(uz, c[8 + 3 * k], gb[k])));
        }
#pragma unroll
        for (int q = 0; q < 12; ++q) acc[q].x = sa[q];
      }
      // flushes: fold the two vertex streams, stream 0 writes the slot (16 samples x 48 contiguous bytes)
      if (d1.x >= 0) {
        float s[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) { s[q] = stream_sum(acc[q].x); acc[q].x = 0.f; }
        if (st == 0 && valid) {
          if (a.gA_dense) {
            float* d = a.gA_dense + ((long long)b * a.J + d1.x) * 12;
            red_add_v4(d, s[0], s[1], s[2], s[3]);
            red_add_v4(d + 4, s[4], s[5], s[6], s[7]);
            red_add_v4(d + 8, s[8], s[9], s[10], s[11]);
          } else {
            float4* d = reinterpret_cast<float4*>(a.gA_slots + ((long long)d1.x * a.B + b) * 12);
            d[0] = make_float4(s[0], s[1], s[2], s[3]);
            d[1] = make_float4(s[4], s[5], s[6], s[7]);
            d[2] = make_float4(s[8], s[9], s[10], s[11]);
          }
        }
      }
      if (d1.y >= 0) {
        float s[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) { s[q] = stream_sum(acc[q].y); acc[q].y = 0.f; }
        if (st == 0 && valid) {
          if (a.gA_dense) {
            float* d = a.gA_dense + ((long long)b * a.J + d1.y) * 12;
            red_add_v4(d, s[0], s[1], s[2], s[3]);
            red_add_v4(d + 4, s[4], s[5], s[6], s[7]);
            red_add_v4(d + 8, s[8], s[9], s[10], s[11]);
          } else {
            float4* d = reinterpret_cast<float4*>(a.gA_slots + ((long long)d1.y * a.B + b) * 12);
            d[0] = make_float4(s[0], s[1], s[2], s[3]);
            d[1] = make_float4(s[4], s[5], s[6], s[7]);
            d[2] = make_float4(s[8], s[9], s[10], s[11]);
          }
        }
      }
    }
      if (p == NPH - 1) {
        // g_beta / g_trans partials of this tile: streams -> warp -> CTA, fixed order (folded below, after the barrier)
#pragma unroll
        for (int k = 0; k < NBT; ++k) gb[k] = stream_sum(gb[k]);
        gt0 = stream_sum(gt0); gt1 = stream_sum(gt1); gt2 = stream_sum(gt2);
        if (st == 0) {
          float* r = sRed + ((size_t)warp * SKIN_BWD_G + g) * (NBT + 4);
#pragma unroll
          for (int k = 0; k < NBT; ++k) r[k] = gb[k];
          r[NBT] = gt0; r[NBT + 1] = gt1; r[NBT + 2] = gt2;
        }
      }
      __syncthreads();  // all warps done with this phase's rows (and, after the last phase, sRed complete)
      if (warp == 0 && grp + gstep < ngroups) issue(grp + gstep, p);
    }
    {   // the warps fold the group's samples four at a time: lane -> (sample, element), no divisions
      const int ne = a.nb + 3;
      for (int gg = warp * 4 + (lane >> 3); gg < SKIN_BWD_G; gg += 4 * NW)
        if (b0 + gg < a.B)
          for (int e = lane & 7; e < ne; e += 8) {
            const int src = (e < a.nb) ? e : NBT + (e - a.nb);
            float sum = 0.f;
#pragma unroll
            for (int wv = 0; wv < NW; ++wv) sum += sRed[((size_t)wv * SKIN_BWD_G + gg) * (NBT + 4) + src];
            if (a.gbt_sum) atomicAdd(a.gbt_sum + (long long)(b0 + gg) * ne + e, sum);
            else a.gbt_part[((long long)(b0 + gg) * m.vm_ntile + ti) * ne + e] = sum;
          }
    }
    // sRed is rewritten only after the next group's phase barriers
  }
}

// Sum of the per-(tile, bone) partial slots -> dense dL/dA [B][J][12], and of the per-tile (g_beta, g_trans) partials ->
// [B][nb+3].  The pose backward is a latency-bound warp-per-sample kernel; gathering up to ~40 slots per bone there is a
// chain of dependent L2 round trips on ~3 lanes.  Here the same sums are a streaming pass: blockIdx.y = bone, threads run
// over the float4 index (sample, row) so that every slot read is a coalesced 512-byte warp access and the loop over a
// bone's slots has the same trip count for the whole block; blockIdx.y == J handles the per-tile partials.
struct SlotReduceArgs {
  int B, J, ne, ntile;
  const float* gA_slots; long long slot_bstride, slot_sstride;   // float4 units: per sample, per slot
  const int* sb_ptr;         // [J+1]
  const float* gbt_part;     // [B][ntile][ne]
  float* gA_dense;           // [B][J][12]
  float* gbt_sum;            // [B][ne]
};
__global__ void __launch_bounds__(256) slot_reduce_kernel(SlotReduceArgs a) {
  const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (j == a.J) {   // per-tile partials: thread per (sample, element)
    if (f >= (long long)a.B * a.ne) return;
    const long long b = f / a.ne;
    const int e = (int)(f - b * a.ne);
    const float* p = a.gbt_part + b * a.ntile * a.ne + e;
    float s = 0.f;
    for (int t = 0; t < a.ntile; ++t) s += p[(long long)t * a.ne];
    a.gbt_sum[f] = s;
    return;
  }
  if (f >= 3LL * a.B) return;
  const long long b = f / 3;
  const int r = (int)(f - 3 * b);
  const float4* sl = reinterpret_cast<const float4*>(a.gA_slots) + b * a.slot_bstride + r;
  int si = a.sb_ptr[j];
  const int se = a.sb_ptr[j + 1];
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (; si + 3 < se; si += 4) {
    const float4 t0 = sl[si * a.slot_sstride], t1 = sl[(si + 1) * a.slot_sstride], t2 = sl[(si + 2) * a.slot_sstride],
                 t3 = sl[(si + 3) * a.slot_sstride];
    acc.x += (t0.x + t1.x) + (t2.x + t3.x); acc.y += (t0.y + t1.y) + (t2.y + t3.y);
    acc.z += (t0.z + t1.z) + (t2.z + t3.z); acc.w += (t0.w + t1.w) + (t2.w + t3.w);
  }
  for (; si < se; ++si) {
    const float4 t0 = sl[si * a.slot_sstride];
    acc.x += t0.x; acc.y += t0.y; acc.z += t0.z; acc.w += t0.w;
  }
  reinterpret_cast<float4*>(a.gA_dense)[(b * a.J + j) * 3 + r] = acc;
}

}  // namespace smil
