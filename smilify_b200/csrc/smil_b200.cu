// C-ABI entry points of the B200-native SMAL/SMIL hot path (see include/smil_b200.h).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -shared -Xcompiler -fPIC
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "fused_kernels.cuh"
#include "generic_kernels.cuh"
#include "pose_kernels.cuh"
#include "pose_kernels2.cuh"
#include "pose_kernels3.cuh"
#include "fitter_kernels.cuh"
#include "proj_kernels.cuh"
#include "skin_bwd_vm.cuh"
#include "tri_kernels.cuh"
#include "smil_common.cuh"

using namespace smil;

struct SmilModel {
  DevModel d;
};

// ---- launch counter and optional per-kernel timing ----------------------------------------------
namespace smil {
namespace {
struct ProfSlot { const char* name; cudaEvent_t beg, end; };
std::atomic<long long> g_launches{0};
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
std::vector<ProfSlot> g_prof_slots;   // event pairs, created by smil_profile_enable
size_t g_prof_used = 0;
thread_local int t_prof_cur = -1;
thread_local cudaStream_t t_prof_stream = nullptr;
}  // namespace

void prof_begin(const char* name, cudaStream_t st) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  t_prof_cur = -1;
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_prof_used >= g_prof_slots.size()) return;
  ProfSlot& s = g_prof_slots[g_prof_used];
  s.name = name;
  if (cudaEventRecord(s.beg, st) != cudaSuccess) { cudaGetLastError(); return; }
  t_prof_cur = (int)g_prof_used++;
  t_prof_stream = st;
}
void prof_end() {
  if (t_prof_cur < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if ((size_t)t_prof_cur < g_prof_slots.size()) cudaEventRecord(g_prof_slots[t_prof_cur].end, t_prof_stream);
  t_prof_cur = -1;
}
}  // namespace smil

namespace {

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct ArenaPlan {
  struct Item { const void* src; size_t bytes; const void** dst; };
  std::vector<Item> items;
  size_t total = 0;
  template <typename T>
  void add(const T* src, size_t count, const T** dst) {
    *dst = nullptr;
    if (!src || count == 0) return;
    items.push_back({src, count * sizeof(T), reinterpret_cast<const void**>(dst)});
    total = align_up(total, 256) + count * sizeof(T);
  }
};

inline int nbt_of(int nb) {
  static const int vals[] = {1, 2, 3, 4, 5, 6, 8, 10, 13, 16};
  for (int v : vals)
    if (v >= nb) return v;
  return 0;
}

// beta-count template instances of the fused kernels (packer.NBT_VALUES).  -DSMIL_FAST_BUILD compiles only the two
// the shipped rigs use (nB = 3, 5): a development switch, never the shipped library.
#ifdef SMIL_FAST_BUILD
#define SMIL_NBT_CASES(fn) case 3: return fn<3>(m, a, st); case 5: return fn<5>(m, a, st);
#else
#define SMIL_NBT_CASES(fn)                                                                                      \
  case 1: return fn<1>(m, a, st); case 2: return fn<2>(m, a, st); case 3: return fn<3>(m, a, st);               \
  case 4: return fn<4>(m, a, st); case 5: return fn<5>(m, a, st); case 6: return fn<6>(m, a, st);               \
  case 8: return fn<8>(m, a, st); case 10: return fn<10>(m, a, st); case 13: return fn<13>(m, a, st);           \
  case 16: return fn<16>(m, a, st);
#endif

template <typename Kern>
int set_smem(Kern k, size_t bytes) {
  if (bytes > 227 * 1024) return fail(SMIL_EUNSUPPORTED, "kernel needs %zu bytes of shared memory (> 227 KB)", bytes);
  if (bytes > 48 * 1024) SMIL_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return SMIL_OK;
}

// ---- pose kernel launchers ---------------------------------------------------------------------
// the lane-per-sample forward covers the headline configuration (see pose_kernels3.cuh); everything else keeps the
// warp-per-sample kernel
static bool pose_ls_fwd_ok(const DevModel& m, const PoseArgs& a) {
  if (!m.ls_ok || !a.pp || a.A || a.ls || a.bt || (a.flags & SMIL_F_THETA_IS_ROTMAT)) return false;
  if (a.jrest_mode == 2 || !a.joints || (a.joints_mode != 1 && a.joints_mode != 2) || a.nb > 16) return false;
  if (!m.ls_blob[a.jrest_mode == 0 ? 0 : 1]) return false;
  PoseLs L;
  L.init_fwd(m);
  return (size_t)L.total * sizeof(float) <= 227 * 1024;
}

static bool pose_ls_bwd_ok(const DevModel& m, const PoseArgs& a) {
  if (!m.ls_ok || (a.flags & SMIL_F_THETA_IS_ROTMAT) || a.bt || a.g_bt || a.g_Rs || a.nb > 16) return false;
  if (!a.gA_dense || a.gA_slots || (a.gbt_part && a.ntile != 1)) return false;
  if (!(a.jrest_mode == 1 || (a.jrest_mode == 2 && a.jrest_in && a.jrest_closed))) return false;
  if (!m.ls_blob[a.jrest_mode == 1 ? 1 : 0]) return false;
  if (a.joints_mode != 1 && a.joints_mode != 2) return false;
  PoseLs L;
  L.init_bwd(m);
  return (size_t)L.total * sizeof(float) <= 227 * 1024;
}

int launch_pose(const DevModel& m, const PoseArgs& a, bool backward, cudaStream_t st) {
  PoseSmem L;
  L.J = m.J;
  L.nbp = (std::max(a.nb, 1) + 3) & ~3;
  const int pairs_mode = backward && a.joints_mode == 1 && a.g_joints && !a.joints_in_skin;
  const int fl = backward ? L.bwd_floats() : L.fwd_floats();
  const size_t smem = (size_t)((fl + 3) & ~3) * sizeof(float) * POSE_WPB;
  const int grid = (a.B + POSE_WPB - 1) / POSE_WPB;
  int rc = SMIL_OK;
  const bool lean = backward && m.tune.pose_lean != 2 && a.beta_final && a.pp_saved && !a.ls && !a.g_newJ &&
                    !a.g_jrest && !a.g_Rs2 && (a.jrest_mode == 1 || (a.jrest_mode == 2 && a.jrest_in));
  if (lean && m.tune.pose_v2 == 1 && a.nb <= 16 && L.nbp + 3 <= 32) {
    // several samples per warp (pose_kernels2.cuh): 4 when every tree level fits 8 lanes per sample, else 2
    const int spw = m.max_width <= 8 ? 4 : 2;
    const int ps = pose2_bwd_floats(m.J, L.nbp, pairs_mode ? m.npairs_pad : 0);
    const size_t smem2 = (size_t)POSE2_WPB * spw * ps * sizeof(float);
    const int grid2 = (a.B + POSE2_WPB * spw - 1) / (POSE2_WPB * spw);
    const bool nb8 = a.nb <= 8;
    prof_begin("pose_bwd_lean_kernel", st);
#define SMIL_POSE_BWD2(SPW_, NB_)                                                                   \
    do {                                                                                             \
      if ((rc = set_smem(pose_bwd2_kernel<SPW_, NB_>, smem2)) != SMIL_OK) return rc;                  \
      pose_bwd2_kernel<SPW_, NB_><<<grid2, 32 * POSE2_WPB, smem2, st>>>(m, a, L.nbp, pairs_mode);     \
    } while (0)
    if (spw == 4) { if (nb8) SMIL_POSE_BWD2(4, 8); else SMIL_POSE_BWD2(4, 16); }
    else { if (nb8) SMIL_POSE_BWD2(2, 8); else SMIL_POSE_BWD2(2, 16); }
#undef SMIL_POSE_BWD2
  } else if (lean && m.tune.pose_v2 == 0 && pose_ls_bwd_ok(m, a)) {
    // lane = sample (pose_kernels3.cuh): a CTA per 32 samples, its warps walk the chains of the tree
    PoseLs LS;
    LS.init_bwd(m);
    const size_t smem3 = (size_t)LS.total * sizeof(float);
    if (a.nb <= 8) {
      if ((rc = set_smem(pose_bwd_ls_kernel<8>, smem3)) != SMIL_OK) return rc;
      prof_begin("pose_bwd_ls_kernel", st);
      pose_bwd_ls_kernel<8><<<(a.B + 31) / 32, LS_THREADS, smem3, st>>>(m, a, LS, pairs_mode);
    } else {
      if ((rc = set_smem(pose_bwd_ls_kernel<16>, smem3)) != SMIL_OK) return rc;
      prof_begin("pose_bwd_ls_kernel", st);
      pose_bwd_ls_kernel<16><<<(a.B + 31) / 32, LS_THREADS, smem3, st>>>(m, a, LS, pairs_mode);
    }
  } else if (lean) {   // no per-joint scales: half the shared memory, the whole batch resident in one wave
    PoseSmemLean LL;
    LL.J = m.J;
    LL.nbp = L.nbp;
    const size_t smem_l = (size_t)LL.floats() * sizeof(float) * POSE_WPB;
    if ((rc = set_smem(pose_bwd_lean_kernel, smem_l)) != SMIL_OK) return rc;
    prof_begin("pose_bwd_lean_kernel", st);
    pose_bwd_lean_kernel<<<grid, POSE_THREADS, smem_l, st>>>(m, a, LL, pairs_mode);
  } else if (backward) {
    if ((rc = set_smem(pose_bwd_kernel, smem)) != SMIL_OK) return rc;
    prof_begin("pose_bwd_kernel", st);
    pose_bwd_kernel<<<grid, POSE_THREADS, smem, st>>>(m, a, L, pairs_mode);
  } else if (m.tune.pose_v2 == 0 && pose_ls_fwd_ok(m, a)) {
    // lane = sample (pose_kernels3.cuh): a CTA per 32 samples, its warps split the joints
    PoseLs LS;
    LS.init_fwd(m);
    const size_t smem3 = (size_t)LS.total * sizeof(float);
    if (a.nb <= 8) {
      if ((rc = set_smem(pose_fwd_ls_kernel<8>, smem3)) != SMIL_OK) return rc;
      prof_begin("pose_fwd_ls_kernel", st);
      pose_fwd_ls_kernel<8><<<(a.B + 31) / 32, LS_THREADS, smem3, st>>>(m, a, LS);
    } else {
      if ((rc = set_smem(pose_fwd_ls_kernel<16>, smem3)) != SMIL_OK) return rc;
      prof_begin("pose_fwd_ls_kernel", st);
      pose_fwd_ls_kernel<16><<<(a.B + 31) / 32, LS_THREADS, smem3, st>>>(m, a, LS);
    }
  } else if (m.tune.pose_v2 == 1) {
    // several samples per warp: 4 when every tree level fits 8 lanes per sample, else 2
    const int spw = m.max_width <= 8 ? 4 : 2;
    const int ps = pose2_fwd_floats(m.J, L.nbp, (a.joints && a.joints_mode == 1) ? m.npairs_pad : 0);
    const size_t smem2 = (size_t)POSE2_WPB * spw * ps * sizeof(float);
    const int grid2 = (a.B + POSE2_WPB * spw - 1) / (POSE2_WPB * spw);
    prof_begin("pose_fwd_kernel", st);
    if (spw == 4) {
      if ((rc = set_smem(pose_fwd2_kernel<4>, smem2)) != SMIL_OK) return rc;
      pose_fwd2_kernel<4><<<grid2, 32 * POSE2_WPB, smem2, st>>>(m, a, L.nbp);
    } else {
      if ((rc = set_smem(pose_fwd2_kernel<2>, smem2)) != SMIL_OK) return rc;
      pose_fwd2_kernel<2><<<grid2, 32 * POSE2_WPB, smem2, st>>>(m, a, L.nbp);
    }
  } else {
    if ((rc = set_smem(pose_fwd_kernel, smem)) != SMIL_OK) return rc;
    prof_begin("pose_fwd_kernel", st);
    pose_fwd_kernel<<<grid, POSE_THREADS, smem, st>>>(m, a, L);
  }
  SMIL_LAUNCH_CHECK(lean ? "pose_bwd_lean_kernel" : (backward ? "pose_bwd_kernel" : "pose_fwd_kernel"));
  return SMIL_OK;
}

// ---- fused skin launchers ----------------------------------------------------------------------
template <int NBT, int K, bool ALIGNED>
int launch_skin_fwd_inst2(const DevModel& m, const SkinFwdArgs& a, cudaStream_t st) {
  const size_t smem = skin_fwd_smem_bytes(a.GP, a.blk, SKIN_FWD_THREADS);
  int rc = set_smem(skin_fwd_kernel<NBT, K, ALIGNED>, smem);
  if (rc != SMIL_OK) return rc;
  // persistent CTAs: as many per vertex chunk as fit the machine in one wave; each walks its share of the pair groups
  static thread_local size_t occ_smem = 0;
  static thread_local int occ = 0;
  if (occ_smem != smem) {
    SMIL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, skin_fwd_kernel<NBT, K, ALIGNED>, SKIN_FWD_THREADS, smem));
    occ_smem = smem;
  }
  const int ngroups = ((a.B + 1) / 2 + a.GP - 1) / a.GP;
  const int cps = std::min(m.tune.fwd_cps > 0 ? m.tune.fwd_cps : 3, std::max(occ, 1));
  const int Y = std::max(1, std::min(ngroups, (m.sm_count * cps) / std::max(m.fwd_nchunk, 1)));
  dim3 grid(m.fwd_nchunk, Y);
  prof_begin("skin_fwd_kernel", st);
  skin_fwd_kernel<NBT, K, ALIGNED><<<grid, SKIN_FWD_THREADS, smem, st>>>(m, a);
  SMIL_LAUNCH_CHECK("skin_fwd_kernel");
  return SMIL_OK;
}
template <int NBT, int K>
int launch_skin_fwd_inst(const DevModel& m, const SkinFwdArgs& a, cudaStream_t st) {
  // rows are 16-byte aligned in the output tensors iff 3V % 4 == 0 (chunks are 256 vertices = 3072 bytes)
  if ((3 * m.V) % 4 == 0) return launch_skin_fwd_inst2<NBT, K, true>(m, a, st);
  return launch_skin_fwd_inst2<NBT, K, false>(m, a, st);
}
template <int NBT>
int launch_skin_fwd_k(const DevModel& m, const SkinFwdArgs& a, cudaStream_t st) {
  if (m.K <= 1) return launch_skin_fwd_inst<NBT, 1>(m, a, st);
  if (m.K <= 2) return launch_skin_fwd_inst<NBT, 2>(m, a, st);
  return launch_skin_fwd_inst<NBT, 4>(m, a, st);
}
int launch_skin_fwd(const DevModel& m, const SkinFwdArgs& a, cudaStream_t st) {
  switch (nbt_of(a.nb)) {
    SMIL_NBT_CASES(launch_skin_fwd_k)
  }
  return fail(SMIL_EUNSUPPORTED, "fused forward: %d betas not covered", a.nb);
}

template <int NBT, int STAGES>
int launch_skin_bwd_inst2(const DevModel& m, const SkinBwdArgs& a, cudaStream_t st) {
  const size_t smem = skin_bwd_smem_bytes(m.bwd_chunk, m.bwd_rec, a.max_tile_pairs, NBT, STAGES);
  int rc = set_smem(skin_bwd_kernel<NBT, STAGES>, smem);
  if (rc != SMIL_OK) return rc;
  static thread_local size_t occ_smem = 0;
  static thread_local int occ = 0;
  if (occ_smem != smem) {
    SMIL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, skin_bwd_kernel<NBT, STAGES>, 128, smem));
    occ_smem = smem;
  }
  // persistent CTAs: one wave; every CTA keeps its tile's constants and walks its share of the sample groups
  const int ngroups = (a.B + SKIN_BWD_G - 1) / SKIN_BWD_G;
  const int cps = m.tune.bwd_cps > 0 ? std::min(m.tune.bwd_cps, std::max(occ, 1)) : std::max(occ, 1);
  const int Y = std::max(1, std::min(ngroups, (m.sm_count * cps) / std::max(m.bwd_ntile, 1)));
  dim3 grid(m.bwd_ntile, Y);
  prof_begin("skin_bwd_kernel", st);
  skin_bwd_kernel<NBT, STAGES><<<grid, 128, smem, st>>>(m, a);
  SMIL_LAUNCH_CHECK("skin_bwd_kernel");
  return SMIL_OK;
}
template <int NBT>
int launch_skin_bwd_vm_inst(const DevModel& m, const SkinBwdVmArgs& a, cudaStream_t st) {
  const size_t smem = skin_bwd_vm_smem_bytes(m.vm_chunk, m.vm_nphase, m.vm_warps, m.vm_recf, a.max_tile_recs, a.max_tile_segs, NBT);
  int rc = set_smem(skin_bwd_vm_kernel<NBT>, smem);
  if (rc != SMIL_OK) return rc;
  static thread_local size_t occ_smem = 0;
  static thread_local int occ = 0;
  if (occ_smem != smem) {
    SMIL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, skin_bwd_vm_kernel<NBT>, 32 * m.vm_warps, smem));
    occ_smem = smem;
  }
  // persistent CTAs: one wave; every CTA keeps its tile's constants and walks its share of the sample groups
  const int ngroups = (a.B + SKIN_BWD_G - 1) / SKIN_BWD_G;
  const int cps = m.tune.bwd_cps > 0 ? std::min(m.tune.bwd_cps, std::max(occ, 1)) : std::max(occ, 1);
  const int Y = std::max(1, std::min(ngroups, (m.sm_count * cps) / std::max(m.vm_ntile, 1)));
  dim3 grid(m.vm_ntile, Y);
  prof_begin("skin_bwd_vm_kernel", st);
  skin_bwd_vm_kernel<NBT><<<grid, 32 * m.vm_warps, smem, st>>>(m, a);
  SMIL_LAUNCH_CHECK("skin_bwd_vm_kernel");
  return SMIL_OK;
}
int launch_skin_bwd_vm(const DevModel& m, const SkinBwdVmArgs& a, cudaStream_t st) {
  switch (nbt_of(m.nB)) {   // records are laid out for the MODEL's beta count; unused betas enter as zeros
    SMIL_NBT_CASES(launch_skin_bwd_vm_inst)
  }
  return fail(SMIL_EUNSUPPORTED, "vertex-major fused backward: %d betas not covered", m.nB);
}

template <int NBT>
int launch_skin_bwd_inst(const DevModel& m, const SkinBwdArgs& a, cudaStream_t st) {
  if (m.tune.bwd_stages != 2) return launch_skin_bwd_inst2<NBT, 1>(m, a, st);
  return launch_skin_bwd_inst2<NBT, 2>(m, a, st);
}
int launch_skin_bwd(const DevModel& m, const SkinBwdArgs& a, cudaStream_t st) {
  switch (nbt_of(a.nb)) {
    SMIL_NBT_CASES(launch_skin_bwd_inst)
  }
  return fail(SMIL_EUNSUPPORTED, "fused backward: %d betas not covered", a.nb);
}

// ---- mode selection shared by forward and backward ---------------------------------------------
struct Plan {
  bool fused;      // fused skinning + closed-form regressions
  bool is_static;
  bool rotmat;
  bool joints_only;
  int jrest_mode;  // 0 closed form, 1 static, 2 regressed from v_shaped (saved)
};

int make_plan(const DevModel& m, const SmilFwdArgs& f, Plan* p) {
  if (f.batch <= 0) return fail(SMIL_EINVAL, "batch must be positive (got %d)", f.batch);
  if (f.num_betas_used < 0 || f.num_betas_used > m.nB)
    return fail(SMIL_EINVAL, "num_betas_used=%d outside [0,%d]", f.num_betas_used, m.nB);
  if (f.num_betas_used == 0 && f.batch != 1)
    return fail(SMIL_EINVAL, "beta with 0 columns is only defined for batch 1 (smal_torch.py:250-253)");
  if (!f.theta) return fail(SMIL_EINVAL, "theta is NULL");
  if (f.num_betas_used > 0 && !f.beta) return fail(SMIL_EINVAL, "beta is NULL");
  p->is_static = (f.flags & SMIL_F_STATIC_JOINTS) != 0;
  if (p->is_static && !m.has_static) return fail(SMIL_EINVAL, "STATIC_JOINTS requested but the model has no J table");
  p->rotmat = (f.flags & SMIL_F_THETA_IS_ROTMAT) != 0;
  p->joints_only = (f.flags & SMIL_F_JOINTS_ONLY) != 0;
  const bool custom_shape = f.del_v != nullptr || f.v_template != nullptr;
  p->fused = m.fused_ok && !custom_shape && !(f.flags & SMIL_F_FORCE_GENERIC);
  p->jrest_mode = p->is_static ? 1 : (custom_shape ? 2 : 0);
  return SMIL_OK;
}

// saved = bone transforms (fused path: pair-interleaved, ceil(B/2) pairs; generic path: [B][J][12]) + rest joints
int pair_blk(const DevModel& m) { return pair_block_floats(m.J, std::max(nbt_of(m.nB), 1)); }
size_t saved_A_bytes(const DevModel& m, int B) {
  return align_up((size_t)((B + 1) / 2) * std::max(pair_blk(m), PPB * m.J) * 4, 256);
}
size_t saved_jrest_bytes(const DevModel& m, int B) { return align_up((size_t)B * m.J * 3 * 4, 256); }
size_t saved_bytes_for(const DevModel& m, int B) { return saved_A_bytes(m, B) + saved_jrest_bytes(m, B); }

PoseArgs base_pose_args(const DevModel& m, const SmilFwdArgs& f, const Plan& p) {
  PoseArgs a;
  memset(&a, 0, sizeof(a));
  a.B = f.batch;
  a.nb = f.num_betas_used;
  a.flags = f.flags;
  a.jrest_mode = p.jrest_mode;
  a.beta = f.beta; a.beta_stride = f.beta_stride;
  a.theta = f.theta; a.theta_stride = f.theta_stride;
  a.trans = f.trans; a.trans_stride = f.trans_stride;
  a.ls = f.betas_logscale; a.ls_stride = f.logscale_stride;
  a.bt = f.betas_trans; a.bt_stride = f.btrans_stride;
  (void)m;
  return a;
}

inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

}  // namespace

// =====================================================================================================
extern "C" {

int smil_abi_version(void) { return SMIL_ABI_VERSION; }
const char* smil_last_error(void) { return err_buf(); }

long long smil_launch_count(void) { return g_launches.load(); }

int smil_profile_enable(int max_launches) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on.store(0);
  for (auto& s : g_prof_slots) { cudaEventDestroy(s.beg); cudaEventDestroy(s.end); }
  g_prof_slots.clear();
  g_prof_used = 0;
  if (max_launches <= 0) return SMIL_OK;
  g_prof_slots.resize((size_t)max_launches);
  for (auto& s : g_prof_slots) {
    s.name = nullptr;
    if (cudaEventCreate(&s.beg) != cudaSuccess || cudaEventCreate(&s.end) != cudaSuccess) {
      g_prof_slots.clear();
      return fail(SMIL_ECUDA, "smil_profile_enable: cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
  }
  g_prof_on.store(1);
  return SMIL_OK;
}

int smil_profile_collect(SmilProfileEntry* out, int max_entries, int* num_entries) {
  if (!out || !num_entries || max_entries <= 0) return fail(SMIL_EINVAL, "smil_profile_collect: bad argument");
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int n = 0;
  for (size_t i = 0; i < g_prof_used; ++i) {
    ProfSlot& s = g_prof_slots[i];
    SMIL_CUDA(cudaEventSynchronize(s.end));
    float ms = 0.f;
    SMIL_CUDA(cudaEventElapsedTime(&ms, s.beg, s.end));
    int k = 0;
    for (; k < n; ++k)
      if (strncmp(out[k].name, s.name, sizeof(out[k].name) - 1) == 0) break;
    if (k == n) {
      if (n == max_entries) continue;
      memset(&out[n], 0, sizeof(out[n]));
      strncpy(out[n].name, s.name, sizeof(out[n].name) - 1);
      ++n;
    }
    out[k].launches += 1;
    out[k].total_ms += ms;
  }
  g_prof_used = 0;
  *num_entries = n;
  return SMIL_OK;
}

namespace {
struct DeviceGuard {   // the caller's current CUDA device is restored on every exit path
  int prev = -1;
  explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; cudaSetDevice(dev); }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
}  // namespace

int smil_model_create(const SmilModelDesc* s, SmilModel** out) {
  if (!s || !out) return fail(SMIL_EINVAL, "smil_model_create: NULL argument");
  *out = nullptr;
  if (s->num_joints <= 0) return fail(SMIL_EINVAL, "num_joints must be positive");
  if (!s->parents || !s->order || !s->level_ptr || !s->child_ptr)
    return fail(SMIL_EINVAL, "tree tables (parents/order/level_ptr/child_ptr) are required");
  int ndev = 0;
  SMIL_CUDA(cudaGetDeviceCount(&ndev));
  if (s->device < 0 || s->device >= ndev) return fail(SMIL_EINVAL, "device %d outside [0,%d)", s->device, ndev);
  DeviceGuard guard(s->device);
  SmilModel* M = new (std::nothrow) SmilModel();
  if (!M) return fail(SMIL_ENOMEM, "out of host memory");
  DevModel& d = M->d;
  memset(&d, 0, sizeof(d));
  d.V = s->num_verts; d.J = s->num_joints; d.nB = s->num_betas; d.P = s->num_pose_basis; d.K = s->skin_k;
  d.jreg_nnz = s->jreg_nnz; d.num_levels = s->num_levels; d.max_width = s->max_width; d.npairs = s->num_pairs;
  d.has_static = s->has_static && s->J_static;
  d.fwd_chunk = s->fwd_chunk; d.fwd_nchunk = s->fwd_nchunk; d.v_pad = s->v_pad;
  d.bwd_chunk = s->bwd_chunk; d.bwd_ntile = s->bwd_ntile; d.bwd_warps = s->bwd_warps; d.bwd_rec = s->bwd_rec;
  d.num_units = s->num_units; d.num_item_pairs = s->num_item_pairs;
  d.h_max_tile_pairs = s->bwd_max_tile_pairs;
  d.bwd_tail = 6 + 6 * std::max(nbt_of(d.nB), 1);
  d.h_max_jv = d.h_max_je = d.h_max_tile_units = 0;
  if (s->tjv_ptr && s->jr_ptr && s->tile_unit_ptr && s->bwd_chunk > 0)
    for (int t = 0; t < s->bwd_ntile; ++t) {
      const int v0 = t * s->bwd_chunk, v1 = std::min(s->num_verts, v0 + s->bwd_chunk);
      d.h_max_jv = std::max(d.h_max_jv, s->tjv_ptr[t + 1] - s->tjv_ptr[t]);
      d.h_max_je = std::max(d.h_max_je, s->jr_ptr[v1] - s->jr_ptr[v0]);
      d.h_max_tile_units = std::max(d.h_max_tile_units, s->tile_unit_ptr[t + 1] - s->tile_unit_ptr[t]);
    }
  d.fused_ok = s->fused_ok && d.V > 0 && s->fwd_chunk == SKIN_FWD_THREADS && s->bwd_warps == 4 && d.K <= 4 &&
               nbt_of(d.nB) > 0 && d.P == 0 && s->bwd_chunk > 0 &&
               s->bwd_rec == skin_bwd_rec_floats(nbt_of(d.nB)) && d.J <= 64 && d.h_max_tile_units <= SKIN_BWD_MAXU;
  d.device = s->device;
  d.tune = s->tuning;
  SMIL_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, s->device));

  const size_t V = d.V, J = d.J, nB = d.nB, K = d.K, np = std::max(d.npairs, 1);
  ArenaPlan ap;
  ap.add(s->v_template, V * 3, &d.v_template);
  ap.add(s->shapedirs, nB * 3 * V, &d.shapedirs);
  ap.add(s->posedirs, (size_t)d.P * 3 * V, &d.posedirs);
  ap.add(s->J_static, d.has_static ? J * 3 : 0, &d.J_static);
  ap.add(s->skin_idx, V * K, &d.skin_idx);
  ap.add(s->skin_w, V * K, &d.skin_w);
  ap.add(s->jc_ptr, V ? J + 1 : 0, &d.jc_ptr);
  ap.add(s->jc_vid, V ? std::max(d.jreg_nnz, 1) : 0, &d.jc_vid);
  ap.add(s->jc_val, V ? std::max(d.jreg_nnz, 1) : 0, &d.jc_val);
  ap.add(s->jr_ptr, V ? V + 1 : 0, &d.jr_ptr);
  ap.add(s->jr_jid, V ? std::max(d.jreg_nnz, 1) : 0, &d.jr_jid);
  ap.add(s->jr_val, V ? std::max(d.jreg_nnz, 1) : 0, &d.jr_val);
  ap.add(s->parents, J, &d.parents);
  ap.add(s->order, J, &d.order);
  ap.add(s->level_ptr, (size_t)d.num_levels + 1, &d.level_ptr);
  ap.add(s->child_ptr, J + 1, &d.child_ptr);
  ap.add(s->child_idx, std::max<size_t>(J - 1, 1), &d.child_idx);
  ap.add(s->JT, V ? J * 3 : 0, &d.JT);
  ap.add(s->JS, V ? nB * J * 3 : 0, &d.JS);
  ap.add(s->jreg_rowsum, V ? J : 0, &d.jreg_rowsum);
  ap.add(s->pj_ptr, V ? J + 1 : 0, &d.pj_ptr);
  ap.add(s->pj_bone, V ? np : 0, &d.pj_bone);
  ap.add(s->pj_joint, V ? np : 0, &d.pj_joint);
  ap.add(s->pj_c0, V ? np * 4 : 0, &d.pj_c0);
  ap.add(s->pj_cm, V ? np * std::max<size_t>(nB, 1) * 3 : 0, &d.pj_cm);
  d.nbp4 = (std::max(d.nB, 1) + 3) & ~3;
  d.pr_nq = 2 + 3 * d.nbp4 / 4;
  d.npairs_pad = (d.npairs + 31) / 32 * 32;
  std::vector<float> pjrec4, pbrec4;
  if (V && d.npairs > 0 && s->pj_bone && s->pj_joint && s->pj_c0 && s->pj_cm) {
    // order[] = position -> pair index; joint-sorted is the order the packer emits, bone-sorted is a stable sort of it
    std::vector<int> by_joint(d.npairs), by_bone(d.npairs);
    for (int i = 0; i < d.npairs; ++i) by_joint[i] = by_bone[i] = i;
    std::stable_sort(by_bone.begin(), by_bone.end(), [&](int x, int y) { return s->pj_bone[x] < s->pj_bone[y]; });
    auto fill = [&](const std::vector<int>& order, std::vector<float>& out) {
      const size_t nq = d.pr_nq;
      out.assign((size_t)d.npairs_pad * nq * 4, 0.f);
      for (int pos = 0; pos < d.npairs_pad; ++pos) {
        float* base = out.data() + ((size_t)(pos / 32) * nq * 32 + (pos % 32)) * 4;   // float4 q of this lane: base + q*128
        int bone = -1, joint = 0;
        if (pos < d.npairs) {
          const int pi = order[pos];
          bone = s->pj_bone[pi]; joint = s->pj_joint[pi];
          base[2] = s->pj_c0[pi * 4 + 3];
          for (int c = 0; c < 3; ++c) base[128 + c] = s->pj_c0[pi * 4 + c];
          for (size_t t = 0; t < nB; ++t)
            for (size_t c = 0; c < 3; ++c)
              base[(2 + 3 * (t / 4) + c) * 128 + (t % 4)] = s->pj_cm[((size_t)pi * nB + t) * 3 + c];
        }
        memcpy(base + 0, &bone, 4);
        memcpy(base + 1, &joint, 4);
      }
    };
    fill(by_joint, pjrec4);
    fill(by_bone, pbrec4);
    ap.add(reinterpret_cast<const float4*>(pjrec4.data()), pjrec4.size() / 4, &d.pjrec4);
    ap.add(reinterpret_cast<const float4*>(pbrec4.data()), pbrec4.size() / 4, &d.pbrec4);
  }
  // ---- lane-per-sample pose kernels: chain schedule + re-laid tables (pose_kernels3.cuh) -------------------------
  std::vector<int> ls_int;
  std::vector<float> ls_blob0, ls_blob1;
  d.ls_ok = 0;
  if (V && J >= 2) {
    const int Ji = (int)J;
    std::vector<std::vector<int>> kids(J);
    bool tree_ok = s->parents[0] < 0;
    for (int j = 1; j < Ji && tree_ok; ++j) {
      tree_ok = s->parents[j] >= 0 && s->parents[j] < j;   // parents precede children (the level tables assume it too)
      if (tree_ok) kids[s->parents[j]].push_back(j);
    }
    if (tree_ok) {
      std::vector<int> height(J, 1), heavy(J, -1);
      for (int j = Ji - 1; j >= 0; --j)
        for (int c : kids[j])
          if (height[c] + 1 > height[j]) { height[j] = height[c] + 1; heavy[j] = c; }
      // chains: every child of the root and every non-heavy child starts one and follows the heavy children
      std::vector<int> chain_ptr{0}, chain_joint, chain_parent, chain_of(J, -1), chain_round;
      for (int h = 1; h < Ji; ++h) {
        const int p = s->parents[h];
        if (p != 0 && heavy[p] == h) continue;
        const int c = (int)chain_parent.size();
        for (int i = h; i >= 0; i = heavy[i]) { chain_joint.push_back(i); chain_of[i] = c; }
        chain_ptr.push_back((int)chain_joint.size());
        chain_parent.push_back(p);
        chain_round.push_back(p == 0 ? 0 : chain_round[chain_of[p]] + 1);   // heads are visited in increasing joint order
      }
      const int nchain = (int)chain_parent.size();
      int nround = 0;
      for (int r : chain_round) nround = std::max(nround, r + 1);
      // per round: longest chains first onto the least loaded warp
      std::vector<int> sched_ptr{0}, sched_chain;
      for (int r = 0; r < nround; ++r) {
        std::vector<int> cs;
        for (int c = 0; c < nchain; ++c) if (chain_round[c] == r) cs.push_back(c);
        std::stable_sort(cs.begin(), cs.end(), [&](int x, int y) {
          return chain_ptr[x + 1] - chain_ptr[x] > chain_ptr[y + 1] - chain_ptr[y]; });
        std::vector<std::vector<int>> per(LS_NW);
        std::vector<int> load(LS_NW, 0);
        for (int c : cs) {
          const int wmin = (int)(std::min_element(load.begin(), load.end()) - load.begin());
          per[wmin].push_back(c);
          load[wmin] += chain_ptr[c + 1] - chain_ptr[c];
        }
        for (int wv = 0; wv < LS_NW; ++wv) {
          for (int c : per[wv]) sched_chain.push_back(c);
          sched_ptr.push_back((int)sched_chain.size());
        }
      }
      std::vector<int> dep_ptr(J + 1, 0), dep_chain(nchain);
      for (int c = 0; c < nchain; ++c) dep_ptr[chain_parent[c] + 1]++;
      for (int j = 0; j < Ji; ++j) dep_ptr[j + 1] += dep_ptr[j];
      {
        std::vector<int> fillp(dep_ptr.begin(), dep_ptr.end() - 1);
        for (int c = 0; c < nchain; ++c) dep_chain[fillp[chain_parent[c]]++] = c;
      }
      // closed-form joints: contiguous joint ranges per warp with about the same number of (joint, bone) pairs
      std::vector<int> pair_jb(LS_NW + 1, Ji);
      pair_jb[0] = 0;
      if (d.npairs > 0 && s->pj_ptr) {
        for (int wv = 1; wv < LS_NW; ++wv) {
          const long long target = (long long)d.npairs * wv / LS_NW;
          int j = pair_jb[wv - 1];
          while (j < Ji && s->pj_ptr[j] < target) ++j;
          pair_jb[wv] = j;
        }
      } else {
        for (int wv = 1; wv < LS_NW; ++wv) pair_jb[wv] = (int)((long long)Ji * wv / LS_NW);
      }
      // backward: contiguous bone ranges per warp with about the same number of pairs (pairs visited by bone)
      std::vector<int> bone_jb(LS_NW + 1, Ji);
      bone_jb[0] = 0;
      for (int wv = 1; wv < LS_NW; ++wv) {
        if (d.npairs > 0 && s->pb_ptr) {
          const long long target = (long long)d.npairs * wv / LS_NW;
          int j = bone_jb[wv - 1];
          while (j < Ji && s->pb_ptr[j] < target) ++j;
          bone_jb[wv] = j;
        } else {
          bone_jb[wv] = (int)((long long)Ji * wv / LS_NW);
        }
      }
      const std::vector<int>* parts[9] = {&chain_ptr, &chain_joint, &chain_parent, &sched_ptr, &sched_chain, &dep_ptr, &dep_chain, &pair_jb, &bone_jb};
      for (int t = 0; t < 9; ++t) {
        d.ls_off[t] = (int)ls_int.size();
        ls_int.insert(ls_int.end(), parts[t]->begin(), parts[t]->end());
      }
      d.ls_nchain = nchain; d.ls_nround = nround;
      d.ls_jsw = (1 + (int)nB + 3) & ~3;
      d.ls_prf = (4 + 3 * (int)nB + 3) & ~3;
      const bool have_pairs = d.npairs > 0 && s->pj_ptr && s->pj_bone && s->pj_joint && s->pj_c0 && s->pj_cm && s->pb_ptr && s->pb_pair;
      const int npi = have_pairs ? d.npairs : 0;
      auto build_blob = [&](bool is_static, std::vector<float>& blob) {
        std::vector<int> w;   // 4-byte words; floats are stored by bit pattern
        auto pad4 = [&]() { while (w.size() & 3) w.push_back(0); };
        auto put_i = [&](const int* p, size_t n) { for (size_t i = 0; i < n; ++i) w.push_back(p ? p[i] : 0); };
        auto put_f = [&](float v) { int b; memcpy(&b, &v, 4); w.push_back(b); };
        int o[10];
        o[LS_O_SCHED] = 0; put_i(ls_int.data(), ls_int.size());
        o[LS_O_PAR] = (int)w.size(); put_i(s->parents, J);
        o[LS_O_PJP] = (int)w.size(); if (have_pairs) put_i(s->pj_ptr, J + 1); else w.insert(w.end(), J + 1, 0);
        o[LS_O_PJB] = (int)w.size(); put_i(s->pj_bone, npi);
        o[LS_O_PBP] = (int)w.size(); if (have_pairs) put_i(s->pb_ptr, J + 1); else w.insert(w.end(), J + 1, 0);
        o[LS_O_PBQ] = (int)w.size(); put_i(s->pb_pair, npi);
        o[LS_O_PJJ] = (int)w.size(); put_i(s->pj_joint, npi);
        pad4();
        o[LS_O_RSUM] = (int)w.size();
        for (size_t j = 0; j < J; ++j) put_f(s->jreg_rowsum ? s->jreg_rowsum[j] : 0.f);
        pad4();
        o[LS_O_JST] = (int)w.size();
        for (size_t e = 0; e < 3 * J; ++e)
          for (int k = 0; k < d.ls_jsw; ++k) {
            float v = 0.f;
            if (is_static) v = (k == 0) ? s->J_static[e] : 0.f;
            else if (k == 0) v = s->JT[e];
            else if ((size_t)k <= nB) v = s->JS[(size_t)(k - 1) * 3 * J + e];
            put_f(v);
          }
        o[LS_O_PREC] = (int)w.size();
        for (int pi = 0; pi < npi; ++pi)
          for (int k = 0; k < d.ls_prf; ++k) {
            float v = 0.f;
            if (k < 4) v = s->pj_c0[pi * 4 + k];
            else if ((size_t)(k - 4) < nB * 3) v = s->pj_cm[(size_t)pi * nB * 3 + (k - 4)];
            put_f(v);
          }
        pad4();
        for (int t = 0; t < 10; ++t) d.ls_o[t] = o[t];
        d.ls_words = (int)w.size();
        blob.resize(w.size());
        memcpy(blob.data(), w.data(), w.size() * 4);
      };
      d.ls_blob[0] = d.ls_blob[1] = nullptr;
      if (s->JT && (nB == 0 || s->JS)) { build_blob(false, ls_blob0); ap.add(ls_blob0.data(), ls_blob0.size(), &d.ls_blob[0]); }
      if (d.has_static && s->J_static) { build_blob(true, ls_blob1); ap.add(ls_blob1.data(), ls_blob1.size(), &d.ls_blob[1]); }
      d.ls_ok = (!ls_blob0.empty() || !ls_blob1.empty()) ? 1 : 0;   // the device pointers are filled in by the upload below
    }
  }
  ap.add(s->pb_ptr, V ? J + 1 : 0, &d.pb_ptr);
  ap.add(s->pb_pair, V ? np : 0, &d.pb_pair);
  d.pbrec_floats = (6 + 3 * d.nB + 3) / 4 * 4;
  ap.add(s->pbrec, (V && s->pbrec) ? (size_t)std::max(d.npairs_pad, 1) * d.pbrec_floats : 0, &d.pbrec);
  ap.add(s->vconst, (3 + 3 * nB) * (size_t)d.v_pad, &d.vconst);
  ap.add(s->ell_idx, K * (size_t)d.v_pad, &d.ell_idx);
  ap.add(s->ell_w, K * (size_t)d.v_pad, &d.ell_w);
  ap.add(s->fwd_pos, (size_t)d.v_pad, &d.fwd_pos);
  ap.add(s->rec2, (size_t)d.num_item_pairs * d.bwd_rec, &d.rec2);
  ap.add(s->unit_bone, (size_t)d.num_units, &d.unit_bone);
  ap.add(s->unit_ptr, d.num_units ? (size_t)d.num_units + 1 : 0, &d.unit_ptr);
  ap.add(s->unit_tile, (size_t)d.num_units, &d.unit_tile);
  ap.add(s->tile_unit_ptr, d.num_units ? (size_t)d.bwd_ntile + 1 : 0, &d.tile_unit_ptr);
  ap.add(s->warp_unit_ptr, d.num_units ? (size_t)d.bwd_ntile * d.bwd_warps + 1 : 0, &d.warp_unit_ptr);
  ap.add(s->warp_unit, (size_t)d.num_units, &d.warp_unit);
  ap.add(s->tjv_ptr, d.num_units ? (size_t)d.bwd_ntile + 1 : 0, &d.tjv_ptr);
  ap.add(s->tjv_v, d.num_units ? (size_t)std::max(d.jreg_nnz, 1) : 0, &d.tjv_v);
  ap.add(s->sb_ptr, d.num_units ? J + 1 : 0, &d.sb_ptr);
  ap.add(s->unit_slot, (size_t)d.num_units, &d.unit_slot);
  d.vm_ok = s->vm_ok && d.fused_ok && d.K <= 2 && s->vm_rec && s->vm_seg && s->vm_tile_rec_ptr && s->vm_warp_seg_ptr &&
            s->vm_sb_ptr && s->vm_slot_bone && s->vm_ntile > 0 && s->vm_recf == vm_rec_floats(std::max(nbt_of(d.nB), 1)) && s->vm_warps == VM_WARPS && s->vm_nphase == VM_PHASES && s->vm_chunk % (4 * s->vm_nphase) == 0;
  d.vm_warps = s->vm_warps; d.vm_nphase = s->vm_nphase; d.vm_chunk = s->vm_chunk; d.vm_ntile = s->vm_ntile; d.vm_recf = s->vm_recf; d.vm_nslots = s->vm_nslots;
  d.h_vm_max_tile_recs = s->vm_max_tile_recs; d.h_vm_max_tile_segs = s->vm_max_tile_segs;
  if (d.vm_ok) {
    ap.add(s->vm_rec, (size_t)s->vm_tile_rec_ptr[s->vm_ntile] * d.vm_recf, &d.vm_rec);
    ap.add(reinterpret_cast<const int4*>(s->vm_seg), (size_t)s->vm_nseg * 2, &d.vm_seg);
    ap.add(s->vm_tile_rec_ptr, (size_t)d.vm_ntile + 1, &d.vm_tile_rec_ptr);
    ap.add(s->vm_warp_seg_ptr, (size_t)d.vm_ntile * d.vm_nphase * d.vm_warps + 1, &d.vm_warp_seg_ptr);
    ap.add(s->vm_sb_ptr, J + 1, &d.vm_sb_ptr);
    ap.add(s->vm_slot_bone, (size_t)std::max(d.vm_nslots, 1), &d.vm_slot_bone);
  }

  void* arena = nullptr;
  cudaError_t e = cudaMalloc(&arena, std::max<size_t>(ap.total, 256));
  if (e != cudaSuccess) {
    delete M;
    return fail(SMIL_ENOMEM, "cudaMalloc of %zu bytes for the packed model failed: %s", ap.total, cudaGetErrorString(e));
  }
  d.arena = arena;
  size_t off = 0;
  for (auto& it : ap.items) {
    off = align_up(off, 256);
    e = cudaMemcpy(static_cast<char*>(arena) + off, it.src, it.bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      cudaFree(arena);
      delete M;
      return fail(SMIL_ECUDA, "upload of a model table failed: %s", cudaGetErrorString(e));
    }
    *it.dst = static_cast<char*>(arena) + off;
    off += it.bytes;
  }
  *out = M;
  return SMIL_OK;
}

void smil_model_destroy(SmilModel* model) {
  if (!model) return;
  if (model->d.arena) cudaFree(model->d.arena);   // cudaFree takes the pointer's own device: no device switch needed
  delete model;
}

int smil_model_info(const SmilModel* model, SmilModelInfo* info) {
  if (!model || !info) return fail(SMIL_EINVAL, "smil_model_info: NULL argument");
  const DevModel& d = model->d;
  info->num_verts = d.V; info->num_joints = d.J; info->num_betas = d.nB; info->num_pose_basis = d.P;
  info->skin_k = d.K; info->jreg_nnz = d.jreg_nnz; info->tree_depth = d.num_levels - 1;
  info->max_level_width = d.max_width; info->posedirs_zero = d.P == 0; info->joint_bone_pairs = d.npairs;
  info->num_chunks = d.fwd_nchunk; info->num_ga_slots = (d.vm_ok && d.tune.bwd_vm != 2) ? d.vm_nslots : d.num_units;
  return SMIL_OK;
}

size_t smil_saved_bytes(const SmilModel* model, int32_t batch, uint32_t) {
  if (!model || batch <= 0) return 0;
  return saved_bytes_for(model->d, batch);
}

size_t smil_fwd_scratch_bytes(const SmilModel* model, const SmilFwdArgs* f) {
  if (!model || !f || f->batch <= 0) return 0;
  const DevModel& m = model->d;
  Plan p;
  if (make_plan(m, *f, &p) == SMIL_OK && p.fused) return 0;  // the fused path needs no forward scratch
  const size_t vb = align_up((size_t)f->batch * m.V * 3 * 4, 256);
  // generic path: v_posed (pose blendshapes) + v_shaped / verts stand-ins for get_skin=False
  return 3 * vb + 256;
}

size_t smil_bwd_scratch_bytes(const SmilModel* model, const SmilBwdArgs* g) {
  if (!model || !g || g->fwd.batch <= 0) return 0;
  const DevModel& m = model->d;
  const size_t B = g->fwd.batch;
  // the same path test as smil_backward: the fused kernels need only the per-(tile, bone) slots and per-tile partials
  // (DESIGN.md section 3), the generic ones three vertex-sized buffers -- never reserve both
  Plan p;
  const bool plan_ok = make_plan(m, g->fwd, &p) == SMIL_OK;
  const bool fused = plan_ok && p.fused && !g->g_v_shaped && !g->g_del_v;
  const size_t fused_b = align_up(B * std::max(std::max(m.num_units, m.vm_nslots), 1) * 48, 256) +
                         align_up(B * std::max(std::max(m.bwd_ntile, m.vm_ntile), 1) * (m.nB + 3) * 4, 256) +
                         align_up(B * m.J * 48, 256) + align_up(B * (m.nB + 3) * 4, 256);   // + summed slots / partials
  if (fused) return fused_b + 256;
  const size_t vb = align_up(B * m.V * 3 * 4, 256);
  const size_t generic = 3 * vb + 2 * align_up(B * m.J * 48, 256) + align_up(B * m.J * 36, 256) +
                         align_up(B * m.J * 12, 256) + align_up(B * 12, 256);
  return generic + 256;
}

int smil_forward(const SmilModel* model, const SmilFwdArgs* f, void* stream) {
  if (!model || !f) return fail(SMIL_EINVAL, "smil_forward: NULL argument");
  const DevModel& m = model->d;
  if (m.V <= 0) return fail(SMIL_EINVAL, "model has no vertices (tree-only model)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Plan p;
  int rc = make_plan(m, *f, &p);
  if (rc != SMIL_OK) return rc;
  const int B = f->batch;
  if (!f->saved || f->saved_bytes < saved_bytes_for(m, B)) return fail(SMIL_ENOMEM, "saved buffer too small");
  if (!f->joints) return fail(SMIL_EINVAL, "joints output is NULL");
  if (!p.joints_only && (!f->verts || !f->v_shaped)) return fail(SMIL_EINVAL, "verts / v_shaped outputs are NULL");
  if (((uintptr_t)f->verts | (uintptr_t)f->v_shaped | (uintptr_t)f->saved) & 15)
    return fail(SMIL_EINVAL, "verts, v_shaped and saved must be 16-byte aligned");
  float* A = static_cast<float*>(f->saved);
  float* jrest = reinterpret_cast<float*>(static_cast<char*>(f->saved) + saved_A_bytes(m, B));

  PoseArgs pa = base_pose_args(m, *f, p);
  pa.Rs = p.rotmat ? nullptr : f->Rs;
  pa.A = p.fused ? nullptr : A;
  pa.pp = p.fused ? A : nullptr;
  pa.pp_blk = pair_blk(m);
  pa.newJ = f->J_transformed;
  pa.joints = f->joints;

  if (p.fused) {
    pa.joints_mode = p.is_static ? 2 : 1;
    if (!p.is_static) pa.jrest_out = jrest;   // the backward reads the rest joints back instead of re-deriving them
    if ((rc = launch_pose(m, pa, false, st)) != SMIL_OK) return rc;
    if (!p.joints_only) {
      SkinFwdArgs sa;
      sa.B = B; sa.nb = f->num_betas_used;
      sa.GP = m.tune.fwd_gp > 0 ? std::min(m.tune.fwd_gp, SKIN_FWD_GPMAX) : 2;
      sa.blk = pair_blk(m);
      // measured on B200 at B = 4096: with the barrier 74.6 vs 76.5 us on STICK (J = 55: 12.5 KB of transforms per TMA
      // group), without it 224 vs 238 us on the mouse rig (J = 31: 7 KB per group) -> keep it for large groups only
      sa.top_sync = m.tune.fwd_topsync ? (m.tune.fwd_topsync == 1) : ((size_t)sa.GP * sa.blk * sizeof(float) > 10 * 1024 ? 1 : 0);
      sa.pp = A; sa.verts = f->verts; sa.v_shaped = f->v_shaped;
      if ((rc = launch_skin_fwd(m, sa, st)) != SMIL_OK) return rc;
    }
    return SMIL_OK;
  }

  // ---- generic (unfused) path ----
  const size_t vb = align_up((size_t)B * m.V * 3 * 4, 256);
  if (!f->scratch || f->scratch_bytes < 3 * vb) return fail(SMIL_ENOMEM, "forward scratch too small");
  char* sc = static_cast<char*>(f->scratch);
  float* v_posed_buf = reinterpret_cast<float*>(sc);
  float* v_shaped = p.joints_only ? reinterpret_cast<float*>(sc + vb) : f->v_shaped;
  float* verts = p.joints_only ? reinterpret_cast<float*>(sc + 2 * vb) : f->verts;
  const float* vt = f->v_template ? f->v_template : m.v_template;
  const long long nv3 = (long long)B * m.V * 3, nbv = (long long)B * m.V;
  prof_begin("shape_fwd_kernel", st);
  shape_fwd_kernel<<<blocks_for(nv3, 256), 256, 0, st>>>(m, B, f->num_betas_used, f->beta, f->beta_stride, vt, f->del_v,
                                                         f->del_v_stride, v_shaped);
  SMIL_LAUNCH_CHECK("shape_fwd_kernel");
  if (p.jrest_mode == 2) {
    prof_begin("regress_fwd_kernel", st);
    regress_fwd_kernel<<<blocks_for((long long)B * m.J * 32, 256), 256, 0, st>>>(m, B, v_shaped, jrest);
    SMIL_LAUNCH_CHECK("regress_fwd_kernel");
    pa.jrest_in = jrest;
  }
  pa.joints_mode = p.is_static ? 2 : 0;
  if (m.P > 0 && !p.rotmat && !f->Rs) return fail(SMIL_EINVAL, "Rs output required for a rig with pose blendshapes");
  if ((rc = launch_pose(m, pa, false, st)) != SMIL_OK) return rc;
  const float* v_posed = v_shaped;
  if (m.P > 0) {
    const float* Rs = p.rotmat ? f->theta : f->Rs;
    if (p.rotmat && f->theta_stride != (long long)m.J * 9)
      return fail(SMIL_EUNSUPPORTED, "rotation-matrix theta must be contiguous for rigs with pose blendshapes");
    dim3 grid(blocks_for(3LL * m.V, 256), B);
    const size_t sh = (size_t)m.P * 4;
    if ((rc = set_smem(posedirs_fwd_kernel, sh)) != SMIL_OK) return rc;
    prof_begin("posedirs_fwd_kernel", st);
    posedirs_fwd_kernel<<<grid, 256, sh, st>>>(m, B, Rs, v_shaped, v_posed_buf);
    SMIL_LAUNCH_CHECK("posedirs_fwd_kernel");
    v_posed = v_posed_buf;
  }
  prof_begin("lbs_fwd_kernel", st);
  lbs_fwd_kernel<<<blocks_for(nbv, 256), 256, 0, st>>>(m, B, A, v_posed, f->trans, f->trans_stride, verts);
  SMIL_LAUNCH_CHECK("lbs_fwd_kernel");
  if (!p.is_static) {
    prof_begin("regress_fwd_kernel", st);
    regress_fwd_kernel<<<blocks_for((long long)B * m.J * 32, 256), 256, 0, st>>>(m, B, verts, f->joints);
    SMIL_LAUNCH_CHECK("regress_fwd_kernel");
  }
  return SMIL_OK;
}

int smil_backward(const SmilModel* model, const SmilBwdArgs* g, void* stream) {
  if (!model || !g) return fail(SMIL_EINVAL, "smil_backward: NULL argument");
  const DevModel& m = model->d;
  const SmilFwdArgs* f = &g->fwd;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Plan p;
  int rc = make_plan(m, *f, &p);
  if (rc != SMIL_OK) return rc;
  const int B = f->batch, nb = f->num_betas_used;
  if (!f->saved || f->saved_bytes < saved_bytes_for(m, B)) return fail(SMIL_ENOMEM, "saved buffer too small");
  if (!g->scratch || g->scratch_bytes < smil_bwd_scratch_bytes(model, g)) return fail(SMIL_ENOMEM, "backward scratch too small");
  const float* A = static_cast<const float*>(f->saved);
  const float* jrest = reinterpret_cast<const float*>(static_cast<const char*>(f->saved) + saved_A_bytes(m, B));
  char* sc = static_cast<char*>(g->scratch);
  const bool fused = p.fused && !g->g_v_shaped && !g->g_del_v;

  PoseArgs pa = base_pose_args(m, *f, p);
  pa.jrest_in = jrest;
  pa.g_joints = g->g_joints;
  pa.g_Rs = g->g_Rs;
  pa.g_theta = g->g_theta;
  pa.g_ls = f->betas_logscale ? g->g_logscale : nullptr;
  pa.g_bt = f->betas_trans ? g->g_btrans : nullptr;

  if (fused) {
    float* slots = reinterpret_cast<float*>(sc);
    float* part = reinterpret_cast<float*>(sc + align_up((size_t)B * std::max(std::max(m.num_units, m.vm_nslots), 1) * 48, 256));
    pa.sb_ptr = m.sb_ptr;
    if (g->g_verts && m.vm_ok && m.tune.bwd_vm != 2) {
      if (((uintptr_t)g->g_verts) & 15) return fail(SMIL_EINVAL, "g_verts must be 16-byte aligned");
      SkinBwdVmArgs sa;
      sa.B = B; sa.nb = nb;
      sa.max_tile_recs = m.h_vm_max_tile_recs; sa.max_tile_segs = m.h_vm_max_tile_segs;
      sa.beta = f->beta; sa.beta_stride = f->beta_stride;
      sa.pp = A; sa.blk = pair_blk(m); sa.g_verts = g->g_verts;
      sa.total_floats = (long long)B * m.V * 3;
      sa.gA_slots = slots; sa.gbt_part = part;
      sa.gA_dense = nullptr; sa.gbt_sum = nullptr; sa.J = m.J;
      sa.stagger_ns = m.tune.bwd_stagger_ns < 0 ? 0 : (m.tune.bwd_stagger_ns > 0 ? m.tune.bwd_stagger_ns : 0);
      char* q = reinterpret_cast<char*>(part) + align_up((size_t)B * std::max(std::max(m.bwd_ntile, m.vm_ntile), 1) * (m.nB + 3) * 4, 256);
      const size_t dense_b = align_up((size_t)B * m.J * 48, 256);
      if (m.tune.bwd_atomic != 2) {   // dense dL/dA and (g_beta, g_trans) sums accumulated by reductions: zero them first
        sa.gA_dense = reinterpret_cast<float*>(q);
        sa.gbt_sum = reinterpret_cast<float*>(q + dense_b);
        SMIL_CUDA(cudaMemsetAsync(q, 0, dense_b + (size_t)B * (nb + 3) * 4, st));
      }
      if ((rc = launch_skin_bwd_vm(m, sa, st)) != SMIL_OK) return rc;
      if (sa.gA_dense) {
        pa.gA_dense = sa.gA_dense; pa.gbt_part = sa.gbt_sum; pa.ntile = 1;
      } else {
        pa.gA_slots = slots; pa.nslots = m.vm_nslots; pa.sb_ptr = m.vm_sb_ptr;
        pa.slot_bstride = 3; pa.slot_sstride = 3LL * B;
        pa.gbt_part = part; pa.ntile = m.vm_ntile;
      }
    } else if (g->g_verts) {
      if (((uintptr_t)g->g_verts) & 15) return fail(SMIL_EINVAL, "g_verts must be 16-byte aligned");
      SkinBwdArgs sa;
      sa.B = B; sa.nb = nb;
      sa.max_tile_pairs = m.h_max_tile_pairs;
      sa.beta = f->beta; sa.beta_stride = f->beta_stride;
      sa.pp = A; sa.blk = pair_blk(m); sa.g_verts = g->g_verts;
      sa.total_floats = (long long)B * m.V * 3;
      sa.gA_slots = slots; sa.gbt_part = part;
      if ((rc = launch_skin_bwd(m, sa, st)) != SMIL_OK) return rc;
      pa.gA_slots = slots; pa.nslots = m.num_units;
      pa.slot_bstride = 3LL * m.num_units; pa.slot_sstride = 3;
      pa.gbt_part = part; pa.ntile = m.bwd_ntile;
    }
    if (pa.gA_slots && m.tune.bwd_slot_reduce != 2) {
      // the partial slots and per-tile partials are summed by a streaming pass; the pose backward reads dense sums
      char* q = reinterpret_cast<char*>(part) + align_up((size_t)B * std::max(std::max(m.bwd_ntile, m.vm_ntile), 1) * (m.nB + 3) * 4, 256);
      float* dense = reinterpret_cast<float*>(q);
      float* gsum = reinterpret_cast<float*>(q + align_up((size_t)B * m.J * 48, 256));
      SlotReduceArgs ra;
      ra.B = B; ra.J = m.J; ra.ne = nb + 3; ra.ntile = pa.ntile;
      ra.gA_slots = pa.gA_slots; ra.slot_bstride = pa.slot_bstride; ra.slot_sstride = pa.slot_sstride;
      ra.sb_ptr = pa.sb_ptr; ra.gbt_part = pa.gbt_part; ra.gA_dense = dense; ra.gbt_sum = gsum;
      const long long nthr = std::max(3LL * B, (long long)B * ra.ne);
      dim3 grid(blocks_for(nthr, 256), m.J + 1);
      prof_begin("slot_reduce_kernel", st);
      slot_reduce_kernel<<<grid, 256, 0, st>>>(ra);
      SMIL_LAUNCH_CHECK("slot_reduce_kernel");
      pa.gA_slots = nullptr; pa.gA_dense = dense;
      pa.gbt_part = gsum; pa.ntile = 1;
    }
    pa.joints_mode = p.is_static ? 2 : 1;
    pa.g_beta = g->g_beta; pa.g_trans = g->g_trans;
    pa.beta_final = 1;
    pa.jrest_closed = p.jrest_mode == 0;
    if (p.jrest_mode == 0) pa.jrest_mode = 2;   // rest joints saved by the forward
    pa.pp_saved = A; pa.pp_blk = pair_blk(m);
    return launch_pose(m, pa, true, st);
  }

  // ---- generic (unfused) path ----
  const size_t vb = align_up((size_t)B * m.V * 3 * 4, 256);
  float* g_vp = reinterpret_cast<float*>(sc);
  float* v_posed_buf = reinterpret_cast<float*>(sc + vb);
  const float* v_shaped = f->v_shaped;
  if (!v_shaped) {  // get_skin=False forward kept v_shaped in its own scratch: recompute it here
    float* vs = reinterpret_cast<float*>(sc + 2 * vb);
    prof_begin("shape_fwd_kernel", st);
    shape_fwd_kernel<<<blocks_for((long long)B * m.V * 3, 256), 256, 0, st>>>(
        m, B, nb, f->beta, f->beta_stride, f->v_template ? f->v_template : m.v_template, f->del_v, f->del_v_stride, vs);
    SMIL_LAUNCH_CHECK("shape_fwd_kernel");
    v_shaped = vs;
  }
  float* gA = reinterpret_cast<float*>(sc + 3 * vb);
  float* gRs2 = reinterpret_cast<float*>(reinterpret_cast<char*>(gA) + align_up((size_t)B * m.J * 48, 256));
  float* g_jrest = reinterpret_cast<float*>(reinterpret_cast<char*>(gRs2) + align_up((size_t)B * m.J * 36, 256));
  float* g_trans_scr = reinterpret_cast<float*>(reinterpret_cast<char*>(g_jrest) + align_up((size_t)B * m.J * 12, 256));
  float* g_trans = g->g_trans ? g->g_trans : g_trans_scr;
  if (p.fused) {  // the forward of this call was fused and saved the transforms pair-interleaved
    float* A_plain = reinterpret_cast<float*>(reinterpret_cast<char*>(g_trans_scr) + align_up((size_t)B * 12, 256));
    prof_begin("deinterleave_A_kernel", st);
    deinterleave_A_kernel<<<blocks_for((long long)B * m.J * 12, 256), 256, 0, st>>>(B, m.J, pair_blk(m), A, A_plain);
    SMIL_LAUNCH_CHECK("deinterleave_A_kernel");
    A = A_plain;
  }
  const long long nbv = (long long)B * m.V;
  const float* v_posed = v_shaped;
  const float* Rs = p.rotmat ? f->theta : f->Rs;
  if (m.P > 0) {
    if (!Rs) return fail(SMIL_EINVAL, "generic backward needs the forward's Rs output for pose blendshapes");
    dim3 grid(blocks_for(3LL * m.V, 256), B);
    const size_t sh = (size_t)m.P * 4;
    if ((rc = set_smem(posedirs_fwd_kernel, sh)) != SMIL_OK) return rc;
    prof_begin("posedirs_fwd_kernel", st);
    posedirs_fwd_kernel<<<grid, 256, sh, st>>>(m, B, Rs, v_shaped, v_posed_buf);
    SMIL_LAUNCH_CHECK("posedirs_fwd_kernel");
    v_posed = v_posed_buf;
    SMIL_CUDA(cudaMemsetAsync(gRs2, 0, (size_t)B * m.J * 36, st));
  }
  SMIL_CUDA(cudaMemsetAsync(gA, 0, (size_t)B * m.J * 48, st));
  SMIL_CUDA(cudaMemsetAsync(g_trans, 0, (size_t)B * 12, st));
  const float* gj_regress = (!p.is_static) ? g->g_joints : nullptr;
  prof_begin("lbs_bwd_kernel", st);
  lbs_bwd_kernel<<<blocks_for(nbv, 256), 256, 0, st>>>(m, B, A, v_posed, g->g_verts, gj_regress, g_vp, gA, g_trans);
  SMIL_LAUNCH_CHECK("lbs_bwd_kernel");
  if (m.P > 0) {
    prof_begin("posedirs_bwd_kernel", st);
    posedirs_bwd_kernel<<<blocks_for((long long)B * m.P * 32, 256), 256, 0, st>>>(m, B, g_vp, gRs2);
    SMIL_LAUNCH_CHECK("posedirs_bwd_kernel");
    pa.g_Rs2 = gRs2;
  }
  pa.gA_dense = gA;
  pa.joints_mode = p.is_static ? 2 : 0;
  pa.g_jrest = p.is_static ? nullptr : g_jrest;
  pa.beta_final = 0;
  pa.A_saved = A;
  if ((rc = launch_pose(m, pa, true, st)) != SMIL_OK) return rc;
  prof_begin("vshaped_grad_kernel", st);
  vshaped_grad_kernel<<<blocks_for(nbv, 256), 256, 0, st>>>(m, B, g_vp, g->g_v_shaped, p.is_static ? nullptr : g_jrest,
                                                            g->g_del_v);   // dL/d(v_shaped) per sample: the gradient of del_v and, summed over the batch by the caller, of a v_template override
  SMIL_LAUNCH_CHECK("vshaped_grad_kernel");
  if (g->g_beta && nb > 0) {
    prof_begin("shape_bwd_kernel", st);
    shape_bwd_kernel<<<blocks_for((long long)B * nb * 32, 256), 256, 0, st>>>(m, B, nb, g_vp, g->g_beta);
    SMIL_LAUNCH_CHECK("shape_bwd_kernel");
  }
  return SMIL_OK;
}

// ---- stand-alone pieces ---------------------------------------------------------------------------
int smil_rodrigues_forward(const float* theta, float* R, int64_t n, void* stream) {
  if (n <= 0) return SMIL_OK;
  if (!theta || !R) return fail(SMIL_EINVAL, "smil_rodrigues_forward: NULL argument");
  prof_begin("rodrigues_fwd_kernel", static_cast<cudaStream_t>(stream));
  rodrigues_fwd_kernel<<<blocks_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(theta, R, n);
  SMIL_LAUNCH_CHECK("rodrigues_fwd_kernel");
  return SMIL_OK;
}
int smil_rodrigues_backward(const float* theta, const float* g_R, float* g_theta, int64_t n, void* stream) {
  if (n <= 0) return SMIL_OK;
  if (!theta || !g_R || !g_theta) return fail(SMIL_EINVAL, "smil_rodrigues_backward: NULL argument");
  prof_begin("rodrigues_bwd_kernel", static_cast<cudaStream_t>(stream));
  rodrigues_bwd_kernel<<<blocks_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(theta, g_R, g_theta, n);
  SMIL_LAUNCH_CHECK("rodrigues_bwd_kernel");
  return SMIL_OK;
}

static PoseArgs chain_pose_args(const DevModel& m, const SmilChainArgs* c) {
  PoseArgs a;
  memset(&a, 0, sizeof(a));
  a.B = c->batch; a.nb = 0;
  a.flags = (c->flags & SMIL_F_PROPAGATE_SCALING) | SMIL_F_THETA_IS_ROTMAT;
  a.jrest_mode = 2;
  a.theta = c->Rs; a.theta_stride = (long long)m.J * 9;
  a.ls = c->betas_logscale; a.ls_stride = (long long)m.J * 3;
  a.bt = c->betas_trans; a.bt_stride = (long long)m.J * 3;
  a.jrest_in = c->Js;
  return a;
}
int smil_chain_forward(const SmilModel* model, const SmilChainArgs* c, void* stream) {
  if (!model || !c || !c->Rs || !c->Js) return fail(SMIL_EINVAL, "smil_chain_forward: NULL argument");
  if (c->batch <= 0) return fail(SMIL_EINVAL, "batch must be positive");
  PoseArgs a = chain_pose_args(model->d, c);
  a.A = c->A; a.newJ = c->new_J;
  return launch_pose(model->d, a, false, static_cast<cudaStream_t>(stream));
}
int smil_chain_backward(const SmilModel* model, const SmilChainArgs* c, void* stream) {
  if (!model || !c || !c->Rs || !c->Js) return fail(SMIL_EINVAL, "smil_chain_backward: NULL argument");
  if (c->batch <= 0) return fail(SMIL_EINVAL, "batch must be positive");
  PoseArgs a = chain_pose_args(model->d, c);
  a.gA_dense = c->g_A; a.g_newJ = c->g_new_J;
  a.g_theta = c->g_Rs; a.g_jrest = c->g_Js;
  a.g_ls = c->betas_logscale ? c->g_logscale : nullptr;
  a.g_bt = c->betas_trans ? c->g_btrans : nullptr;
  return launch_pose(model->d, a, true, static_cast<cudaStream_t>(stream));
}

// ---- reprojection losses ----------------------------------------------------------------------------
static int check_proj(const SmilProjArgs* a) {
  if (!a) return fail(SMIL_EINVAL, "SmilProjArgs is NULL");
  if (a->kind < SMIL_LOSS_FITTER || a->kind > SMIL_LOSS_SINGLEVIEW) return fail(SMIL_EINVAL, "unknown loss kind %d", a->kind);
  if (a->batch <= 0 || a->num_views <= 0 || a->num_joints <= 0)
    return fail(SMIL_EINVAL, "batch/num_views/num_joints must be positive");
  if (!a->joints || !a->cam_R || !a->cam_T || !a->fov_deg || !a->target || !a->partials)
    return fail(SMIL_EINVAL, "joints, cam_R, cam_T, fov_deg, target and partials are required");
  if (a->kind != SMIL_LOSS_FITTER && !(a->norm_size > 0.f)) return fail(SMIL_EINVAL, "norm_size must be positive");
  return SMIL_OK;
}
int smil_project_loss_forward(const SmilProjArgs* a, void* stream) {
  int rc = check_proj(a);
  if (rc != SMIL_OK) return rc;
  if (!a->loss) return fail(SMIL_EINVAL, "loss output is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long warps = (long long)a->batch * a->num_views;
  prof_begin("proj_fwd_kernel", st);
  proj_fwd_kernel<<<blocks_for(warps * 32, 256), 256, 0, st>>>(*a);
  SMIL_LAUNCH_CHECK("proj_fwd_kernel");
  prof_begin("proj_finalize_kernel", st);
  proj_finalize_kernel<<<1, 256, 0, st>>>(*a);
  SMIL_LAUNCH_CHECK("proj_finalize_kernel");
  return SMIL_OK;
}
int smil_project_loss_backward(const SmilProjArgs* a, void* stream) {
  int rc = check_proj(a);
  if (rc != SMIL_OK) return rc;
  const int threads = std::min(128, std::max(32, ((std::max(a->num_joints, a->num_views) + 31) / 32) * 32));
  prof_begin("proj_bwd_kernel", static_cast<cudaStream_t>(stream));
  proj_bwd_kernel<<<a->batch, threads, 0, static_cast<cudaStream_t>(stream)>>>(*a);
  SMIL_LAUNCH_CHECK("proj_bwd_kernel");
  return SMIL_OK;
}
int smil_keypoint3d_loss_forward(const SmilKp3dArgs* a, void* stream) {
  if (!a || !a->joints || !a->target || !a->loss || !a->partials) return fail(SMIL_EINVAL, "smil_keypoint3d_loss_forward: NULL argument");
  if (a->batch <= 0 || a->num_joints <= 0) return fail(SMIL_EINVAL, "batch/num_joints must be positive");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_begin("kp3d_fwd_kernel", st);
  kp3d_fwd_kernel<<<blocks_for((long long)a->batch * 32, 256), 256, 0, st>>>(*a);
  SMIL_LAUNCH_CHECK("kp3d_fwd_kernel");
  prof_begin("kp3d_finalize_kernel", st);
  kp3d_finalize_kernel<<<1, 256, 0, st>>>(*a);
  SMIL_LAUNCH_CHECK("kp3d_finalize_kernel");
  return SMIL_OK;
}
int smil_keypoint3d_loss_backward(const SmilKp3dArgs* a, void* stream) {
  if (!a || !a->joints || !a->target || !a->g_joints || !a->partials) return fail(SMIL_EINVAL, "smil_keypoint3d_loss_backward: NULL argument");
  if (a->batch <= 0 || a->num_joints <= 0) return fail(SMIL_EINVAL, "batch/num_joints must be positive");
  prof_begin("kp3d_bwd_kernel", static_cast<cudaStream_t>(stream));
  kp3d_bwd_kernel<<<blocks_for((long long)a->batch * a->num_joints, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(*a);
  SMIL_LAUNCH_CHECK("kp3d_bwd_kernel");
  return SMIL_OK;
}

// ---- allocation-free variant of the 3-D keypoint loss forward ---------------------------------------------------
// The per-sample partials (the kernel pair's scratch) live in a file-scope device array instead of a caller buffer, so
// a caller that must not allocate (graph capture, serving loops) needs no scratch tensor.  Every GPU that loads the
// library holds these SMIL_KP3D_STATIC_MAX_BATCH * 12 + 16 bytes until the process exits.  All calls on a device share
// the array: calls that overlap on the GPU (different streams) race on it, and it keeps the last call's partials.
static __device__ __align__(256) float g_kp3d_static_partials[3 * SMIL_KP3D_STATIC_MAX_BATCH + 4];

namespace {
constexpr int KP3D_MAX_DEVICES = 64;
std::mutex g_kp3d_mu;
float* g_kp3d_dev_ptr[KP3D_MAX_DEVICES] = {};   // device address of the array per device ordinal (host code must not
                                                // pass the symbol itself: that is a host-side stand-in)
int kp3d_static_ptr(float** out) {
  int dev = -1;
  SMIL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= KP3D_MAX_DEVICES) return fail(SMIL_EUNSUPPORTED, "device ordinal %d beyond %d", dev, KP3D_MAX_DEVICES);
  std::lock_guard<std::mutex> lk(g_kp3d_mu);
  if (!g_kp3d_dev_ptr[dev]) {   // once per device, with that device current
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_kp3d_static_partials);
    if (e != cudaSuccess || !p)
      return fail(SMIL_ECUDA, "cudaGetSymbolAddress(static keypoint-loss scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    g_kp3d_dev_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_kp3d_dev_ptr[dev];
  return SMIL_OK;
}
}  // namespace

int smil_keypoint3d_loss_forward_static(const SmilKp3dArgs* a, void* stream) {
  if (!a || !a->joints || !a->target || !a->loss) return fail(SMIL_EINVAL, "smil_keypoint3d_loss_forward_static: NULL argument");
  if (a->batch <= 0 || a->num_joints <= 0) return fail(SMIL_EINVAL, "batch/num_joints must be positive");
  if (a->batch > SMIL_KP3D_STATIC_MAX_BATCH)
    return fail(SMIL_EUNSUPPORTED, "smil_keypoint3d_loss_forward_static: batch %d exceeds the static scratch (max %d samples); "
                "use smil_keypoint3d_loss_forward with a caller buffer", a->batch, SMIL_KP3D_STATIC_MAX_BATCH);
  float* scratch = nullptr;
  int rc = kp3d_static_ptr(&scratch);
  if (rc != SMIL_OK) return rc;
  SmilKp3dArgs b = *a;
  b.partials = scratch;   // every slot the finalize pass reads ([0, 3 batch)) is written by the first kernel of THIS call,
                          // and the tail by the finalize pass: nothing has to start at zero
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_begin("kp3d_fwd_kernel", st);
  kp3d_fwd_kernel<<<blocks_for((long long)b.batch * 32, 256), 256, 0, st>>>(b);
  SMIL_LAUNCH_CHECK("kp3d_fwd_kernel");
  prof_begin("kp3d_finalize_kernel", st);
  kp3d_finalize_kernel<<<1, 256, 0, st>>>(b);
  SMIL_LAUNCH_CHECK("kp3d_finalize_kernel");
  return SMIL_OK;
}

// ---- fitter parameter-space terms, fused Adam -----------------------------------------------------------------
int smil_fitter_prior_terms(const SmilFitterPriorArgs* a, void* stream) {
  if (!a || !a->global_rot || !a->joint_rot || !a->loss || !a->partials || !a->g_global_rot || !a->g_joint_rot)
    return fail(SMIL_EINVAL, "smil_fitter_prior_terms: NULL argument");
  if (a->num_frames <= 0 || a->num_joints < 2) return fail(SMIL_EINVAL, "num_frames must be positive and num_joints >= 2");
  if (a->frame_offset < 0 || a->frame_offset + a->num_frames > a->total_frames)
    return fail(SMIL_EINVAL, "frames [%lld, %lld) outside the sequence of %lld frames", (long long)a->frame_offset,
                (long long)(a->frame_offset + a->num_frames), (long long)a->total_frames);
  if (a->w_temp > 0.f) {
    if (!a->trans) return fail(SMIL_EINVAL, "the temporal term needs trans");
    if (a->frame_offset > 0 && !a->halo_prev) return fail(SMIL_EINVAL, "w_temp > 0: halo_prev is required for a shard that starts at frame %lld", (long long)a->frame_offset);
    if (a->frame_offset + a->num_frames < a->total_frames && !a->halo_next) return fail(SMIL_EINVAL, "w_temp > 0: halo_next is required for a shard that ends before the sequence does");
  }
  if (a->w_betas > 0.f && a->betas_prec && (!a->betas || a->num_betas <= 0 || a->num_betas > 64))
    return fail(SMIL_EINVAL, "the betas prior needs betas and 1 <= num_betas <= 64");
  if (a->g_trans && !a->trans) return fail(SMIL_EINVAL, "g_trans without trans");
  SmilFitterPriorArgs b = *a;
  if (!b.trans) { b.g_trans = nullptr; b.trans = b.global_rot; }   // never dereferenced for the translation then
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = (size_t)(6 * a->num_joints + FIT_THREADS) * sizeof(float);
  prof_begin("fitter_prior_kernel", st);
  fitter_prior_kernel<<<a->num_frames, FIT_THREADS, smem, st>>>(b);
  SMIL_LAUNCH_CHECK("fitter_prior_kernel");
  prof_begin("fitter_prior_finalize_kernel", st);
  fitter_prior_finalize_kernel<<<1, 256, 0, st>>>(b);
  SMIL_LAUNCH_CHECK("fitter_prior_finalize_kernel");
  return SMIL_OK;
}

int smil_adam_step(const SmilAdamArgs* a, void* stream) {
  if (!a || !a->step) return fail(SMIL_EINVAL, "smil_adam_step: NULL argument");
  if (a->num_tensors <= 0 || a->num_tensors > SMIL_ADAM_MAX_TENSORS)
    return fail(SMIL_EINVAL, "num_tensors must be in [1, %d]", SMIL_ADAM_MAX_TENSORS);
  long long total = 0;
  for (int t = 0; t < a->num_tensors; ++t) {
    if (!a->param[t] || !a->grad[t] || !a->exp_avg[t] || !a->exp_avg_sq[t] || a->numel[t] <= 0)
      return fail(SMIL_EINVAL, "tensor %d: NULL pointer or empty", t);
    total += a->numel[t];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_begin("adam_tick_kernel", st);
  adam_tick_kernel<<<1, 1, 0, st>>>(a->step);
  SMIL_LAUNCH_CHECK("adam_tick_kernel");
  prof_begin("adam_step_kernel", st);
  adam_step_kernel<<<blocks_for(total, 256), 256, 0, st>>>(*a, total);
  SMIL_LAUNCH_CHECK("adam_step_kernel");
  return SMIL_OK;
}

// ---- DLT triangulation, 6-D rotations ---------------------------------------------------------------------
static int check_tri(const SmilTriArgs* a) {
  if (!a) return fail(SMIL_EINVAL, "SmilTriArgs is NULL");
  if (a->batch <= 0 || a->num_views <= 0 || a->num_joints <= 0) return fail(SMIL_EINVAL, "batch/num_views/num_joints must be positive");
  if (!a->keypoints || !a->visibility || !a->cam_R || !a->cam_T || !a->fov_deg)
    return fail(SMIL_EINVAL, "keypoints, visibility, cam_R, cam_T and fov_deg are required");
  if (!(a->image_size > 0.f)) return fail(SMIL_EINVAL, "image_size must be positive");
  return SMIL_OK;
}
int smil_triangulate_dlt_forward(const SmilTriArgs* a, void* stream) {
  int rc = check_tri(a);
  if (rc != SMIL_OK) return rc;
  if (!a->tri) return fail(SMIL_EINVAL, "tri output is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_begin("tri_fwd_kernel", st);
  tri_fwd_kernel<<<blocks_for((long long)a->batch * a->num_joints, 128), 128, 0, st>>>(*a);
  SMIL_LAUNCH_CHECK("tri_fwd_kernel");
  return SMIL_OK;
}
int smil_triangulate_dlt_backward(const SmilTriArgs* a, void* stream) {
  int rc = check_tri(a);
  if (rc != SMIL_OK) return rc;
  if (!a->g_tri) return fail(SMIL_EINVAL, "g_tri is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int threads = std::min(128, std::max(32, ((std::max(a->num_joints, a->num_views) + 31) / 32) * 32));
  const size_t smem = (size_t)a->num_joints * 6 * sizeof(float);
  if ((rc = set_smem(tri_bwd_kernel, smem)) != SMIL_OK) return rc;
  prof_begin("tri_bwd_kernel", st);
  tri_bwd_kernel<<<a->batch, threads, smem, st>>>(*a);
  SMIL_LAUNCH_CHECK("tri_bwd_kernel");
  return SMIL_OK;
}
int smil_rot6d_to_rotmat_forward(const float* d6, float* R, int64_t n, int32_t columns, void* stream) {
  if (n <= 0) return SMIL_OK;
  if (!d6 || !R) return fail(SMIL_EINVAL, "smil_rot6d_to_rotmat_forward: NULL argument");
  prof_begin("rot6d_fwd_kernel", static_cast<cudaStream_t>(stream));
  rot6d_fwd_kernel<<<blocks_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d6, R, n, columns);
  SMIL_LAUNCH_CHECK("rot6d_fwd_kernel");
  return SMIL_OK;
}
int smil_rot6d_to_rotmat_backward(const float* d6, const float* g_R, float* g_d6, int64_t n, int32_t columns, void* stream) {
  if (n <= 0) return SMIL_OK;
  if (!d6 || !g_R || !g_d6) return fail(SMIL_EINVAL, "smil_rot6d_to_rotmat_backward: NULL argument");
  prof_begin("rot6d_bwd_kernel", static_cast<cudaStream_t>(stream));
  rot6d_bwd_kernel<<<blocks_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d6, g_R, g_d6, n, columns);
  SMIL_LAUNCH_CHECK("rot6d_bwd_kernel");
  return SMIL_OK;
}

}  // extern "C"
