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
//   SAMPLE (16 samples x 2 item streams per warp), so every reduction over vertices -- g_A per
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
// Thread = vertex, two samples per pass packed into f32x2 lanes (FFMA2: one issue slot, two samples).
// The per-vertex constants are loop invariant and live in registers as duplicated (c,c) pairs; what
// changes per pass -- beta, trans and the bone transforms of the sample pair -- sits in shared memory
// pair-interleaved, so every LDS delivers (sample b, sample b+1) operands ready for FFMA2.
constexpr int SKIN_FWD_THREADS = 256;  // == packer.FWD_CHUNK
#ifndef SMIL_FWD_PPS
#define SMIL_FWD_PPS 2
#endif
constexpr int SKIN_FWD_PPS = SMIL_FWD_PPS;        // sample pairs computed per CTA barrier / shipping step (1 or 2)
constexpr int SKIN_FWD_STAGES = SKIN_FWD_PPS == 1 ? 3 : 4;   // output staging ring (each stage: 4 rows = verts/v_shaped of 2 samples)
constexpr int SKIN_FWD_GPMAX = 16;     // sample pairs per CTA

struct SkinFwdArgs {
  int B, nb, GP;              // GP sample pairs per group (one TMA load, GP passes)
  int top_sync;               // 1: CTA barrier at the top of every group as well (tuning switch, see the kernel)
  int blk;                    // floats per pair block of `pp`
  const float* pp;            // [ceil(B/2)][blk] per-pair parameter blocks written by the pose pass:
                              //   A2 [J][12][2] | beta2 [NB2max][2] | trans2 [4][2]   (all pair-interleaved)
  float* verts;               // [B][V][3]
  float* v_shaped;            // [B][V][3]
};

__host__ __device__ inline int skin_fwd_stage_floats(int chunk) { return ((3 * chunk + 4) + 3) & ~3; }
__host__ __device__ inline int pair_block_floats(int J, int nbt_model) { return PPB * J + 2 * ((nbt_model + 1) & ~1) + 8; }
__host__ inline size_t skin_fwd_smem_bytes(int GP, int blk, int chunk) {
  return 32 + 2ull * GP * blk * 4 + (size_t)SKIN_FWD_STAGES * 4 * skin_fwd_stage_floats(chunk) * 4;
}

template <int K_>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(K_) : "memory");
}

// One staged row (n floats starting `phase` floats into `stage`) -> global dst[off ..]: the 16-byte
// aligned body leaves through one bulk store issued by lane 0, the <=3-float head / tail through scalar
// stores by lanes 1-3 / 4-6 of the same warp.
__device__ __forceinline__ void store_row(float* __restrict__ dst, long long off, const float* stage, int phase, int n,
                                          int lane) {
  const int head = min(n, (4 - phase) & 3);
  const int body = ((n - head) >> 2) << 2;
  if (lane == 0 && body > 0) bulk_s2g(dst + off + head, stage + phase + head, (uint32_t)body * 4u);
  if (lane >= 1 && lane < 1 + head) dst[off + (lane - 1)] = stage[phase + (lane - 1)];
  const int tail = n - head - body;
  if (lane >= 4 && lane < 4 + tail) dst[off + head + body + (lane - 4)] = stage[phase + head + body + (lane - 4)];
}

template <int NBT, int K, bool ALIGNED>
__global__ void __launch_bounds__(SKIN_FWD_THREADS) skin_fwd_kernel(DevModel m, SkinFwdArgs a) {
  // ALIGNED: 3V % 4 == 0, i.e. every staged row starts and ends on a 16-byte boundary of the output tensors
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int C = SKIN_FWD_THREADS;
  constexpr int SF = ((3 * C + 4) + 3) & ~3;
  constexpr int NB2 = (NBT + 1) & ~1;                // betas per pair block, padded to whole float4s
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chunk = blockIdx.x;
  const int J = m.J, V = m.V;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);        // [2]
  float* sBlk = reinterpret_cast<float*>(smem_raw + 32);          // [2][GP][blk]
  float* sStage = sBlk + 2ull * a.GP * a.blk;                     // [STAGES][4][SF]
  const int npairs = (a.B + 1) >> 1;
  const int ngroups = (npairs + a.GP - 1) / a.GP;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  auto issue = [&](int grp, int buf) {
    const int p0 = grp * a.GP;
    const uint32_t bytes = (uint32_t)min(a.GP, npairs - p0) * (uint32_t)a.blk * 4u;
    mbar_expect_tx(&bars[buf], bytes);
    bulk_g2s(sBlk + (size_t)buf * a.GP * a.blk, a.pp + (long long)p0 * a.blk, bytes, &bars[buf]);
  };
  if (tid == 0 && (int)blockIdx.y < ngroups) issue(blockIdx.y, 0);
  // per-vertex constants -> registers as duplicated pairs (coalesced SoA loads; padded vertices read zeros)
  const int v = chunk * C + tid;
  float2 T2[3], S2[NBT][3], w2[K];
  int bo[K];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float t = m.vconst[(long long)c * m.v_pad + v];
    T2[c] = make_float2(t, t);
  }
#pragma unroll
  for (int k = 0; k < NBT; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float s = (k < m.nB) ? m.vconst[(long long)(3 + 3 * k + c) * m.v_pad + v] : 0.f;
      S2[k][c] = make_float2(s, s);
    }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const float w = (k < m.K) ? m.ell_w[(long long)k * m.v_pad + v] : 0.f;
    w2[k] = make_float2(w, w);
    bo[k] = (k < m.K) ? m.ell_idx[(long long)k * m.v_pad + v] * PPB : 0;
  }
  // warps whose vertices all have a single bone skip the higher skinning slots (the packer groups such vertices)
  bool upper = false;
#pragma unroll
  for (int k = 1; k < K; ++k) upper = upper || (w2[k].x != 0.f);
  const bool multi = __any_sync(0xffffffffu, upper);
  const int nv = min(C, V - chunk * C);
  const int n = 3 * nv;
  const int pos3 = 3 * m.fwd_pos[v];   // this thread's vertex, local to the chunk (packer: forward_thread_assignment)
  const bool writer = pos3 >= 0;
  // row shipping: warp r < 4 owns row r of a stage (0: verts b, 1: verts b+1, 2: v_shaped b, 3: v_shaped b+1)
  const bool shipper = warp < 4;
  const bool second = (warp & 1) != 0;
  float* const ship_dst = (warp < 2) ? a.verts : a.v_shaped;
  const long long off_step = 6LL * V;
  const int NB2M = (a.blk - PPB * J - 8) >> 1;   // beta pairs per block (model-level padding)

  int sidx = 0, it = 0;
  for (int grp = blockIdx.y; grp < ngroups; grp += gridDim.y, ++it) {
    const int buf = it & 1;
    // every thread must be done with the other buffer (group it-1) before it is refilled: on the staged path the
    // barrier of that group's last shipping step, which follows its last read of the buffer, already guarantees it
    if (a.top_sync) __syncthreads();
    if (tid == 0 && grp + (int)gridDim.y < ngroups) issue(grp + gridDim.y, buf ^ 1);
    mbar_wait(&bars[buf], (uint32_t)((it >> 1) & 1));
    const int p0 = grp * a.GP;
    const int GPn = min(a.GP, npairs - p0);
    long long off = ((long long)(2 * p0 + (second ? 1 : 0)) * V + (long long)chunk * C) * 3;   // this warp's row
    long long offw = ((long long)(2 * p0) * V + (long long)chunk * C) * 3;                      // writer phases
    const float* blkp = sBlk + (size_t)buf * a.GP * a.blk;
    for (int g = 0; g < GPn; g += SKIN_FWD_PPS) {
#pragma unroll
     for (int u = 0; u < SKIN_FWD_PPS; ++u) {
      if (g + u >= GPn) break;
      float* st = sStage + (sidx + u) * (4 * SF);
      const float* Ag = blkp + (size_t)u * a.blk;
      const long long offw_u = offw + u * off_step;
      const float4* bp4 = reinterpret_cast<const float4*>(Ag + PPB * J);
      const float4* tp4 = reinterpret_cast<const float4*>(Ag + PPB * J + 2 * NB2M);
      // shape blend for both samples
      float2 vs0 = T2[0], vs1 = T2[1], vs2 = T2[2];
#pragma unroll
      for (int k = 0; k < NB2; k += 2) {
        const float4 bb = bp4[k >> 1];
        const float2 bk = make_float2(bb.x, bb.y);
        vs0 = __ffma2_rn(bk, S2[k][0], vs0); vs1 = __ffma2_rn(bk, S2[k][1], vs1); vs2 = __ffma2_rn(bk, S2[k][2], vs2);
        if (k + 1 < NBT) {
          const float2 bl = make_float2(bb.z, bb.w);
          vs0 = __ffma2_rn(bl, S2[k + 1][0], vs0); vs1 = __ffma2_rn(bl, S2[k + 1][1], vs1);
          vs2 = __ffma2_rn(bl, S2[k + 1][2], vs2);
        }
      }
      // skinning: sum_k w_k (A_k [vs;1]) + trans
      const float4 t01 = tp4[0];
      const float2 t2_ = *reinterpret_cast<const float2*>(tp4 + 1);
      float2 o0 = make_float2(t01.x, t01.y), o1 = make_float2(t01.z, t01.w), o2 = t2_;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        if (k > 0 && !multi) break;
        const float4* Ak = reinterpret_cast<const float4*>(Ag + bo[k]);
        float4 q;
        float2 t;
        q = Ak[1]; t = __ffma2_rn(make_float2(q.x, q.y), vs2, make_float2(q.z, q.w));
        q = Ak[0]; t = __ffma2_rn(make_float2(q.z, q.w), vs1, t); t = __ffma2_rn(make_float2(q.x, q.y), vs0, t);
        o0 = __ffma2_rn(w2[k], t, o0);
        q = Ak[3]; t = __ffma2_rn(make_float2(q.x, q.y), vs2, make_float2(q.z, q.w));
        q = Ak[2]; t = __ffma2_rn(make_float2(q.z, q.w), vs1, t); t = __ffma2_rn(make_float2(q.x, q.y), vs0, t);
        o1 = __ffma2_rn(w2[k], t, o1);
        q = Ak[5]; t = __ffma2_rn(make_float2(q.x, q.y), vs2, make_float2(q.z, q.w));
        q = Ak[4]; t = __ffma2_rn(make_float2(q.z, q.w), vs1, t); t = __ffma2_rn(make_float2(q.x, q.y), vs0, t);
        o2 = __ffma2_rn(w2[k], t, o2);
      }
      if (writer) {
        float* r = st + pos3;
        int ph0 = 0, ph1 = 0;
        if (!ALIGNED) { ph0 = (int)(offw_u & 3); ph1 = (int)((offw_u + 3LL * V) & 3); }
        r[ph0 + 0] = o0.x;  r[ph0 + 1] = o1.x;  r[ph0 + 2] = o2.x;                         // verts, sample b
        r[SF + ph1 + 0] = o0.y;  r[SF + ph1 + 1] = o1.y;  r[SF + ph1 + 2] = o2.y;          // verts, sample b+1
        r[2 * SF + ph0 + 0] = vs0.x; r[2 * SF + ph0 + 1] = vs1.x; r[2 * SF + ph0 + 2] = vs2.x;  // v_shaped, b
        r[3 * SF + ph1 + 0] = vs0.y; r[3 * SF + ph1 + 1] = vs1.y; r[3 * SF + ph1 + 2] = vs2.y;  // v_shaped, b+1
      }
     }
      {
      fence_proxy_async();
      // the stages the NEXT step writes were last read by the stores of step (now)+1-STAGES/PPS
      if (shipper && lane == 0) bulk_wait_read<SKIN_FWD_STAGES / SKIN_FWD_PPS - 2>();
      __syncthreads();
      if (shipper) {
#pragma unroll
        for (int u = 0; u < SKIN_FWD_PPS; ++u) {
          if (g + u < GPn && (!second || 2 * (p0 + g + u) + 1 < a.B)) {
            const float* src = sStage + (sidx + u) * (4 * SF) + warp * SF;
            const long long off_u = off + u * off_step;
            if (ALIGNED) {
              if (lane == 0) bulk_s2g(ship_dst + off_u, src, (uint32_t)n * 4u);
            } else {
              store_row(ship_dst, off_u, src, (int)(off_u & 3), n, lane);
            }
          }
        }
        if (lane == 0) bulk_commit();
      }
      }
      off += SKIN_FWD_PPS * off_step;
      offw += SKIN_FWD_PPS * off_step;
      blkp += (size_t)SKIN_FWD_PPS * a.blk;
      sidx = (sidx + SKIN_FWD_PPS >= SKIN_FWD_STAGES) ? 0 : sidx + SKIN_FWD_PPS;
    }
  }
  if (shipper && lane == 0) bulk_wait_read<0>();
}

// =====================================================================================================
// backward
// =====================================================================================================
// Item-major, lane = sample (SKIN_BWD_G samples x 32/SKIN_BWD_G item streams per warp), two items per step
// packed into f32x2 lanes: the item-pair records hold every constant as an (item i, item i+1) pair, the
// per-sample operands (bone rotation, betas) enter the packed FMAs as broadcast scalars.  All reductions
// over vertices are private register accumulations; the two halves of a pair and the streams are folded
// once per unit.  Persistent: a CTA keeps its tile's records in shared memory and walks its share of the
// sample groups, whose g_verts rows arrive through TMA bulk copies.  The posed-joint regression
// (joints = Jreg^T verts, smal_torch.py:348-351) is NOT folded in here: its backward is the closed-form
// (joint, bone) pair pass of the pose kernel (folding it cost more than it saved).
#ifndef SMIL_SKIN_BWD_G
#define SMIL_SKIN_BWD_G 16
#endif
constexpr int SKIN_BWD_G = SMIL_SKIN_BWD_G;        // samples per warp pass (lane % G): 8, 16 or 32
constexpr int SKIN_BWD_STREAMS = 32 / SKIN_BWD_G;  // item-pair streams per warp (lane / G)
// fold the SKIN_BWD_STREAMS stream partials of a value onto every lane of its sample
__device__ __forceinline__ float stream_sum(float v) {
  if (SKIN_BWD_STREAMS >= 4) v += __shfl_xor_sync(0xffffffffu, v, 8);
  if (SKIN_BWD_STREAMS >= 2) v += __shfl_xor_sync(0xffffffffu, v, 16);
  return v;
}

struct SkinBwdArgs {
  int B, nb;
  int max_tile_pairs;        // smem sizing
  const float* beta;  long long beta_stride;
  const float* pp; int blk;  // per-pair parameter blocks of the forward (A2 [J][12][2] first), blk floats each
  const float* g_verts;      // [B][V][3]
  long long total_floats;    // B*V*3
  float* gA_slots;           // [B][num_units][12]
  float* gbt_part;           // [B][ntile][nb+3]
};

__host__ __device__ inline int skin_bwd_rec_floats(int NBT) { return (6 + 6 * NBT + 8 + 3) & ~3; }
__host__ __device__ inline int skin_bwd_row_floats(int chunk) {
  int r = 3 * chunk + 4;             // +4: alignment slack for the 16-byte superset load
  return r + ((36 - (r & 31)) & 31);  // == 4 (mod 32): the 8 sample rows of a stream hit distinct banks
}
constexpr int SKIN_BWD_MAXU = 64;   // units per tile the shared descriptor table holds (packer keeps tiles below it)
__host__ inline size_t skin_bwd_smem_bytes(int chunk, int rec, int max_pairs, int NBT, int stages) {
  size_t fl = (size_t)max_pairs * rec + (size_t)stages * SKIN_BWD_G * skin_bwd_row_floats(chunk) +
              4ull * SKIN_BWD_G * (NBT + 4) + 4ull * SKIN_BWD_MAXU;
  return 32 + fl * 4;
}

__device__ __forceinline__ float2 dup2(float x) { return make_float2(x, x); }

template <int NBT, int STAGES>
__global__ void __launch_bounds__(128) skin_bwd_kernel(DevModel m, SkinBwdArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int NQ = (6 + 6 * NBT + 3) / 4;   // float4 loads that cover T and the NBT used shape directions
  const int REC = m.bwd_rec;                  // record stride and tail offset follow the MODEL's beta count
  const int TAIL = m.bwd_tail;                // float offset of (w | first | 3*lv) inside a record
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane & (SKIN_BWD_G - 1), st = lane / SKIN_BWD_G;
  const int ti = blockIdx.x;
  const int C = m.bwd_chunk, V = m.V, J = m.J;
  const int RF = skin_bwd_row_floats(C);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [2]
  float* sRec = reinterpret_cast<float*>(smem_raw + 32);   // [max_pairs][REC]
  float* sRows = sRec + (size_t)a.max_tile_pairs * REC;    // [2][G][RF]
  float* sRed = sRows + (size_t)STAGES * SKIN_BWD_G * RF;  // [4 warps][G][NBT+4]
  int4* sUnit = reinterpret_cast<int4*>(sRed + 4ull * SKIN_BWD_G * (NBT + 4));   // [MAXU] (bone*PPB, first pair, end pair, slot)

  const int u_beg = m.tile_unit_ptr[ti], u_end = m.tile_unit_ptr[ti + 1];
  const int pr_base = m.unit_ptr[u_beg], pr_cnt = m.unit_ptr[u_end] - pr_base;
  const int nv = min(C, V - ti * C);
  const int n = 3 * nv;
  const int ngroups = (a.B + SKIN_BWD_G - 1) / SKIN_BWD_G;
  const int gstep = gridDim.y;   // persistent over sample groups: blockIdx.y, blockIdx.y + gridDim.y, ...
  const int wu_base = m.warp_unit_ptr[ti * m.bwd_warps];
  const int wu_beg = m.warp_unit_ptr[ti * m.bwd_warps + warp] - wu_base;
  const int wu_end = m.warp_unit_ptr[ti * m.bwd_warps + warp + 1] - wu_base;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  // tile constants -> shared memory (once per CTA, reused by every sample group)
  {
    const float4* src = reinterpret_cast<const float4*>(m.rec2 + (long long)pr_base * REC);
    float4* dst = reinterpret_cast<float4*>(sRec);
    for (int i = tid; i < pr_cnt * (REC >> 2); i += 128) dst[i] = src[i];
    for (int i = tid; i < u_end - u_beg; i += 128) {   // unit descriptors in the order the warps walk them
      const int u = m.warp_unit[wu_base + i];
      sUnit[i] = make_int4(m.unit_bone[u] * PPB, m.unit_ptr[u] - pr_base, m.unit_ptr[u + 1] - pr_base, m.unit_slot[u]);
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
    if ((int)blockIdx.y < ngroups) issue(blockIdx.y, 0);
    if (STAGES > 1 && (int)blockIdx.y + gstep < ngroups) issue(blockIdx.y + gstep, 1);
  }

  // betas of the group's samples, fetched one group ahead
  float beta_next[NBT];
  auto beta_fetch = [&](int grp) {
    const int bs = min(grp * SKIN_BWD_G + g, a.B - 1);
#pragma unroll
    for (int k = 0; k < NBT; ++k) beta_next[k] = (k < a.nb) ? a.beta[(long long)bs * a.beta_stride + k] : 0.f;
  };
  if ((int)blockIdx.y < ngroups) beta_fetch(blockIdx.y);

  int itn = 0;
  for (int grp = blockIdx.y; grp < ngroups; grp += gstep, ++itn) {
    const int stage = (STAGES > 1) ? (itn & 1) : 0;
    const uint32_t parity = (STAGES > 1) ? ((itn >> 1) & 1) : (itn & 1);
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

    float2 beta2[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) beta2[k] = dup2(beta_next[k]);
    if (grp + gstep < ngroups) beta_fetch(grp + gstep);
    float2 gb2[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) gb2[k] = make_float2(0.f, 0.f);
    float2 gt0 = make_float2(0.f, 0.f), gt1 = gt0, gt2 = gt0;
    const long long off = ((long long)bb * V + (long long)ti * C) * 3;
    const float* row = sRows + ((size_t)stage * SKIN_BWD_G + g) * RF + (int)(off & 3);
    const float* Ab = a.pp + (long long)(bb >> 1) * a.blk + (bb & 1);

    float rn[9];   // rotation of the NEXT unit's bone for this lane's sample (prefetched under the current unit's items)
    auto r_fetch = [&](int wu) {
      const float* Ak = Ab + sUnit[wu].x;
      rn[0] = Ak[0]; rn[1] = Ak[2]; rn[2] = Ak[4]; rn[3] = Ak[8]; rn[4] = Ak[10]; rn[5] = Ak[12];
      rn[6] = Ak[16]; rn[7] = Ak[18]; rn[8] = Ak[20];
    };
    if (wu_beg < wu_end) r_fetch(wu_beg);
    for (int wu = wu_beg; wu < wu_end; ++wu) {
      const int4 ud = sUnit[wu];
      const int i0 = ud.y, i1 = ud.z;
      // rotation part of this sample's bone transform, duplicated into (x,x) pairs
      const float2 r00 = dup2(rn[0]), r01 = dup2(rn[1]), r02 = dup2(rn[2]);
      const float2 r10 = dup2(rn[3]), r11 = dup2(rn[4]), r12 = dup2(rn[5]);
      const float2 r20 = dup2(rn[6]), r21 = dup2(rn[7]), r22 = dup2(rn[8]);
      if (wu + 1 < wu_end) r_fetch(wu + 1);
      float2 acc[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) acc[q] = make_float2(0.f, 0.f);
      for (int it = i0 + st; it < i1; it += SKIN_BWD_STREAMS) {
        const float4* rc4 = reinterpret_cast<const float4*>(sRec + (size_t)it * REC);
        float2 P[2 * NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
          const float4 t4 = rc4[q];
          P[2 * q] = make_float2(t4.x, t4.y);
          P[2 * q + 1] = make_float2(t4.z, t4.w);
        }
        const float2* tl = reinterpret_cast<const float2*>(sRec + (size_t)it * REC + TAIL);
        const float2 w2 = tl[0], fw2 = tl[1], lvp = tl[2];
        const int lv0 = __float_as_int(lvp.x), lv1 = __float_as_int(lvp.y);
        const float2 gx = make_float2(row[lv0], row[lv1]);
        const float2 gy = make_float2(row[lv0 + 1], row[lv1 + 1]);
        const float2 gz = make_float2(row[lv0 + 2], row[lv1 + 2]);
        const float2 wx = __fmul2_rn(w2, gx), wy = __fmul2_rn(w2, gy), wz = __fmul2_rn(w2, gz);
        // u = A_k.R^T (w g)
        const float2 ux = __ffma2_rn(r00, wx, __ffma2_rn(r10, wy, __fmul2_rn(r20, wz)));
        const float2 uy = __ffma2_rn(r01, wx, __ffma2_rn(r11, wy, __fmul2_rn(r21, wz)));
        const float2 uz = __ffma2_rn(r02, wx, __ffma2_rn(r12, wy, __fmul2_rn(r22, wz)));
        float2 vx = P[0], vy = P[1], vz = P[2];
#pragma unroll
        for (int k = 0; k < NBT; ++k) {
          const float2 sx = P[3 + 3 * k], sy = P[4 + 3 * k], sz = P[5 + 3 * k];
          vx = __ffma2_rn(beta2[k], sx, vx); vy = __ffma2_rn(beta2[k], sy, vy); vz = __ffma2_rn(beta2[k], sz, vz);
          gb2[k] = __ffma2_rn(ux, sx, __ffma2_rn(uy, sy, __ffma2_rn(uz, sz, gb2[k])));
        }
        acc[0] = __ffma2_rn(wx, vx, acc[0]); acc[1] = __ffma2_rn(wx, vy, acc[1]); acc[2] = __ffma2_rn(wx, vz, acc[2]);
        acc[3] = __fadd2_rn(acc[3], wx);
        acc[4] = __ffma2_rn(wy, vx, acc[4]); acc[5] = __ffma2_rn(wy, vy, acc[5]); acc[6] = __ffma2_rn(wy, vz, acc[6]);
        acc[7] = __fadd2_rn(acc[7], wy);
        acc[8] = __ffma2_rn(wz, vx, acc[8]); acc[9] = __ffma2_rn(wz, vy, acc[9]); acc[10] = __ffma2_rn(wz, vz, acc[10]);
        acc[11] = __fadd2_rn(acc[11], wz);
        gt0 = __ffma2_rn(fw2, gx, gt0); gt1 = __ffma2_rn(fw2, gy, gt1); gt2 = __ffma2_rn(fw2, gz, gt2);
      }
      float s[12];
#pragma unroll
      for (int q = 0; q < 12; ++q) {
        s[q] = acc[q].x + acc[q].y;
        s[q] = stream_sum(s[q]);
      }
      if (st == 0 && valid) {
        float4* d = reinterpret_cast<float4*>(a.gA_slots + ((long long)b * m.num_units + ud.w) * 12);
        d[0] = make_float4(s[0], s[1], s[2], s[3]);
        d[1] = make_float4(s[4], s[5], s[6], s[7]);
        d[2] = make_float4(s[8], s[9], s[10], s[11]);
      }
    }
    // g_beta / g_trans partials of this tile: pair halves -> streams -> warp -> CTA, fixed order
    float gb[NBT];
#pragma unroll
    for (int k = 0; k < NBT; ++k) {
      gb[k] = gb2[k].x + gb2[k].y;
      gb[k] = stream_sum(gb[k]);
    }
    float gtx = gt0.x + gt0.y, gty = gt1.x + gt1.y, gtz = gt2.x + gt2.y;
    gtx = stream_sum(gtx); gty = stream_sum(gty); gtz = stream_sum(gtz);
    if (st == 0) {
      float* r = sRed + ((size_t)warp * SKIN_BWD_G + g) * (NBT + 4);
#pragma unroll
      for (int k = 0; k < NBT; ++k) r[k] = gb[k];
      r[NBT] = gtx; r[NBT + 1] = gty; r[NBT + 2] = gtz;
    }
    __syncthreads();  // all warps done with this stage's rows and sRed complete
    if (tid == 0 && grp + STAGES * gstep < ngroups) issue(grp + STAGES * gstep, stage);
    {
      const int ne = a.nb + 3;
      for (int i = tid; i < SKIN_BWD_G * ne; i += 128) {
        const int gg = i / ne, e = i - gg * ne;
        const int src = (e < a.nb) ? e : NBT + (e - a.nb);
        if (b0 + gg < a.B) {
          float sum = 0.f;
#pragma unroll
          for (int wv = 0; wv < 4; ++wv) sum += sRed[((size_t)wv * SKIN_BWD_G + gg) * (NBT + 4) + src];
          a.gbt_part[((long long)(b0 + gg) * m.bwd_ntile + ti) * ne + e] = sum;
        }
      }
    }
    // sRed is rewritten only after the next group's first __syncthreads
  }
}

}  // namespace smil
