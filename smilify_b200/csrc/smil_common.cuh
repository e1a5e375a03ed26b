// Shared helpers for the SMAL/SMIL hot-path kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/smil_b200.h"

namespace smil {

// ---- thread-local error message -----------------------------------------------------------
inline char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}
#define SMIL_CUDA(call)                                                                        \
  do {                                                                                         \
    cudaError_t e__ = (call);                                                                  \
    if (e__ != cudaSuccess)                                                                    \
      return smil::fail(SMIL_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),  \
                        __FILE__, __LINE__);                                                   \
  } while (0)
#define SMIL_LAUNCH_CHECK(name)                                                                \
  do {                                                                                         \
    smil::prof_end();                                                                          \
    cudaError_t e__ = cudaGetLastError();                                                      \
    if (e__ != cudaSuccess)                                                                    \
      return smil::fail(SMIL_ECUDA, "launch of %s failed: %s", name, cudaGetErrorString(e__)); \
  } while (0)

// ---- launch counter + optional per-kernel device timing (smil_profile_* in the C ABI) ------------
// prof_begin() goes right before a kernel launch, SMIL_LAUNCH_CHECK() right after it.  The counter
// is always on; the cudaEvent pair around the launch only while profiling is enabled.
void prof_begin(const char* name, cudaStream_t st);
void prof_end();

// ---- device-resident model ------------------------------------------------------------------
struct DevModel {
  int V, J, nB, P, K;
  int jreg_nnz, num_levels, max_width, npairs, has_static;
  int fwd_chunk, fwd_nchunk, v_pad;
  int bwd_chunk, bwd_ntile, bwd_warps, bwd_rec, bwd_tail, num_units, num_item_pairs;
  int fused_ok;
  int device;
  int sm_count;
  const float* v_template;
  const float* shapedirs;
  const float* posedirs;
  const float* J_static;
  const int* skin_idx;
  const float* skin_w;
  const int *jc_ptr, *jc_vid;
  const float* jc_val;
  const int *jr_ptr, *jr_jid;
  const float* jr_val;
  const int *parents, *order, *level_ptr, *child_ptr, *child_idx;
  const float *JT, *JS, *jreg_rowsum;
  const int *pj_ptr, *pj_bone, *pj_joint;
  const float *pj_c0, *pj_cm;
  // joint-bone coupling records for the lane-per-pair passes, SoA by float4 so that a warp's loads are coalesced:
  //   rec4[(s0/32 * pr_nq + q) * 32 + lane], q0 = (bone bits | joint bits | c0.w | 0), q1 = (c0.xyz, 0),
  //   q(2+3t4 .. 4+3t4) = (cm_x | cm_y | cm_z)[4 t4 .. 4 t4 + 3];  padding records carry bone -1 and zeros.
  const float4* pjrec4;  // sorted by joint (forward: segmented sum per joint)
  const float4* pbrec4;  // sorted by bone  (backward: segmented sum per bone)
  int nbp4, pr_nq;       // beta axis padded to a multiple of 4; float4s per record = 2 + 3 nbp4 / 4
  const int *pb_ptr, *pb_pair;
  const float* pbrec;
  int pbrec_floats, npairs_pad;
  const float* vconst;
  const int* ell_idx;
  const float* ell_w;
  const int* fwd_pos;
  const float* rec2;
  const int *unit_bone, *unit_ptr, *unit_tile, *tile_unit_ptr, *warp_unit_ptr, *warp_unit, *tjv_ptr, *tjv_v;
  const int *sb_ptr, *unit_slot;
  // vertex-major backward skin pass (skin_bwd_vm.cuh)
  int vm_ok, vm_chunk, vm_nphase, vm_warps, vm_ntile, vm_recf, vm_nslots, h_vm_max_tile_recs, h_vm_max_tile_segs;
  const float* vm_rec;
  const int4* vm_seg;
  const int *vm_tile_rec_ptr, *vm_warp_seg_ptr, *vm_sb_ptr, *vm_slot_bone;
  // lane-per-sample pose kernels (pose_kernels3.cuh): ONE blob of 4-byte words per rest-joint mode, built at model
  // creation and staged into shared memory with a single pass of 16-byte copies.  Word offsets in ls_o[]:
  //   LS_O_SCHED  chain schedule: [chain_ptr (nchain+1) | chain_joint (J-1) | chain_parent (nchain) |
  //               sched_ptr (nround*LS_NW+1) | sched_chain (nchain) | dep_ptr (J+1) | dep_chain (nchain) |
  //               pair_jb (LS_NW+1) | bone_jb (LS_NW+1)], sub-offsets in ls_off[9]
  //   LS_O_PAR parents (J) | LS_O_PJP pj_ptr (J+1) | LS_O_PJB pj_bone (np) | LS_O_PBP pb_ptr (J+1) | LS_O_PBQ pb_pair (np) |
  //   LS_O_PJJ pj_joint (np) | LS_O_RSUM jreg_rowsum (J) | LS_O_JST [3J][ls_jsw] (JT | JS_0 .. | 0) per joint component
  //   (static rigs: (J_static | 0)) | LS_O_PREC [np][ls_prf] (c0.xyz, c0.w | cm[nB][3] | 0) per (joint, bone) pair
  int ls_ok, ls_nchain, ls_nround, ls_jsw, ls_prf, ls_words;
  int ls_off[9];
  int ls_o[10];
  const float* ls_blob[2];   // [0] closed-form rest joints, [1] static rest joints (null when the rig has none)
  // host copies needed by launch logic
  int h_max_tile_pairs, h_max_jv, h_max_je, h_max_tile_units;
  SmilTuning tune;   // per-model kernel variant switches (host-side launch logic only)
  void* arena;  // single device allocation that backs every table
};

// Pair-block layout (the per-sample-pair parameter blocks shared by the pose and skin passes): every bone owns
// PPB floats, the first 24 are its 3x4 transform interleaved over the two samples of the pair.  28 instead
// of 24 makes 8 consecutive bones start in 8 different 16-byte bank groups, so the per-thread float4
// gathers of the forward skin pass collide only for bones that are congruent mod 8 (not mod 4).
constexpr int PPB = 28;
enum { LS_O_SCHED = 0, LS_O_PAR, LS_O_PJP, LS_O_PJB, LS_O_PBP, LS_O_PBQ, LS_O_PJJ, LS_O_RSUM, LS_O_JST, LS_O_PREC };

// ---- small device math ------------------------------------------------------------------------
struct f3 {
  float x, y, z;
};
__device__ __forceinline__ f3 mk3(float x, float y, float z) {
  f3 r;
  r.x = x; r.y = y; r.z = z;
  return r;
}
__device__ __forceinline__ f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(float* p, f3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// 3x3 row-major in registers
struct m3 {
  float a[9];
};
__device__ __forceinline__ f3 mul(const m3& M, f3 v) {  // M v
  return mk3(fmaf(M.a[0], v.x, fmaf(M.a[1], v.y, M.a[2] * v.z)), fmaf(M.a[3], v.x, fmaf(M.a[4], v.y, M.a[5] * v.z)),
             fmaf(M.a[6], v.x, fmaf(M.a[7], v.y, M.a[8] * v.z)));
}
__device__ __forceinline__ f3 mulT(const m3& M, f3 v) {  // M^T v
  return mk3(fmaf(M.a[0], v.x, fmaf(M.a[3], v.y, M.a[6] * v.z)), fmaf(M.a[1], v.x, fmaf(M.a[4], v.y, M.a[7] * v.z)),
             fmaf(M.a[2], v.x, fmaf(M.a[5], v.y, M.a[8] * v.z)));
}
__device__ __forceinline__ m3 matmul(const m3& A, const m3& B) {
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.a[i * 3 + j] = fmaf(A.a[i * 3], B.a[j], fmaf(A.a[i * 3 + 1], B.a[3 + j], A.a[i * 3 + 2] * B.a[6 + j]));
  return C;
}
__device__ __forceinline__ m3 matmul_AtB(const m3& A, const m3& B) {  // A^T B
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.a[i * 3 + j] = fmaf(A.a[i], B.a[j], fmaf(A.a[3 + i], B.a[3 + j], A.a[6 + i] * B.a[6 + j]));
  return C;
}
__device__ __forceinline__ m3 matmul_ABt(const m3& A, const m3& B) {  // A B^T
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.a[i * 3 + j] =
          fmaf(A.a[i * 3], B.a[j * 3], fmaf(A.a[i * 3 + 1], B.a[j * 3 + 1], A.a[i * 3 + 2] * B.a[j * 3 + 2]));
  return C;
}

// Rodrigues exactly as the reference composes it (smal_model/batch_lbs.py:31-50):
// angle = ||theta + 1e-8||, axis = theta / angle (NOT normalised), R = cos I + (1-cos) r r^T + sin [r]x.
__device__ __forceinline__ m3 rodrigues(f3 th) {
  const float e = 1e-8f;
  float tx = th.x + e, ty = th.y + e, tz = th.z + e;
  float a = sqrtf(fmaf(tx, tx, fmaf(ty, ty, tz * tz)));
  float rx = th.x / a, ry = th.y / a, rz = th.z / a;
  float s, c;
  sincosf(a, &s, &c);
  float omc = 1.0f - c;
  m3 R;
  R.a[0] = fmaf(omc, rx * rx, c);
  R.a[1] = fmaf(omc, rx * ry, -s * rz);
  R.a[2] = fmaf(omc, rx * rz, s * ry);
  R.a[3] = fmaf(omc, ry * rx, s * rz);
  R.a[4] = fmaf(omc, ry * ry, c);
  R.a[5] = fmaf(omc, ry * rz, -s * rx);
  R.a[6] = fmaf(omc, rz * rx, -s * ry);
  R.a[7] = fmaf(omc, rz * ry, s * rx);
  R.a[8] = fmaf(omc, rz * rz, c);
  return R;
}

// VJP of rodrigues(): G = dL/dR  ->  dL/dtheta (derivation: tests/emulate.py::rodrigues_bwd).
__device__ __forceinline__ f3 rodrigues_vjp(f3 th, const m3& G) {
  const float e = 1e-8f;
  float tx = th.x + e, ty = th.y + e, tz = th.z + e;
  float a = sqrtf(fmaf(tx, tx, fmaf(ty, ty, tz * tz)));
  float inv_a = 1.0f / a;
  float rx = th.x * inv_a, ry = th.y * inv_a, rz = th.z * inv_a;
  float s, c;
  sincosf(a, &s, &c);
  float kx = G.a[7] - G.a[5], ky = G.a[2] - G.a[6], kz = G.a[3] - G.a[1];
  // (G + G^T) r
  float sx = 2.f * G.a[0] * rx + (G.a[1] + G.a[3]) * ry + (G.a[2] + G.a[6]) * rz;
  float sy = (G.a[1] + G.a[3]) * rx + 2.f * G.a[4] * ry + (G.a[5] + G.a[7]) * rz;
  float sz = (G.a[2] + G.a[6]) * rx + (G.a[5] + G.a[7]) * ry + 2.f * G.a[8] * rz;
  float rGr = 0.5f * (rx * sx + ry * sy + rz * sz);
  float gc = (G.a[0] + G.a[4] + G.a[8]) - rGr;
  float gs = rx * kx + ry * ky + rz * kz;
  float omc = 1.0f - c;
  float grx = fmaf(omc, sx, s * kx), gry = fmaf(omc, sy, s * ky), grz = fmaf(omc, sz, s * kz);
  float ga = -s * gc + c * gs - (grx * th.x + gry * th.y + grz * th.z) * inv_a * inv_a;
  float gai = ga * inv_a;
  return mk3(fmaf(gai, tx, grx * inv_a), fmaf(gai, ty, gry * inv_a), fmaf(gai, tz, grz * inv_a));
}

}  // namespace smil
