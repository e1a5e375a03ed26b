/*
 * smil_b200.h -- C ABI of the B200-native SMAL/SMIL hot path.
 *
 * Drop-in boundary for ONE path of FabianPlum/SMILify: the differentiable forward and
 * backward of `smal_model.smal_torch.SMAL.__call__` and the 2D joint reprojection losses
 * built on it.  Plain pointers and sizes only; no torch types.  Every pointer in the hot
 * calls is a DEVICE pointer owned by the caller; the library owns only the packed
 * constants behind `SmilModel`.  All work is enqueued on the `stream` argument (a
 * `cudaStream_t` passed as void*); nothing synchronises and nothing allocates in the hot
 * calls.  Functions return 0 on success, a negative SMIL_E* code otherwise, and never
 * abort: the reference's callers catch exceptions and continue
 * (/root/reference/smal_fitter/neuralSMIL/multiview_smil_regressor.py:1044-1046).
 * `smil_last_error()` returns a thread-local message for the last failing call.
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   smil_model_create           <- SMAL.__init__                     smal_model/smal_torch.py:88-196
 *   smil_forward / smil_backward<- SMAL.__call__ (+ its autograd)    smal_model/smal_torch.py:198-370
 *       incl. batch_rodrigues                                        smal_model/batch_lbs.py:31-50
 *             batch_global_rigid_transformation                      smal_model/batch_lbs.py:75-197
 *   smil_rodrigues_*            <- batch_rodrigues (standalone)      smal_model/batch_lbs.py:31-50
 *   smil_project_loss_*         <- Renderer.forward projection slice smal_fitter/p3d_renderer.py:130-140
 *                                  SMALFitter joint loss             smal_fitter/fitter.py:280-296
 *                                  _batch_project_joints_to_views    smal_fitter/neuralSMIL/multiview_smil_regressor.py:1623-1705
 *                                  _compute_batched_keypoint_loss    smal_fitter/neuralSMIL/multiview_smil_regressor.py:2319-2388
 *                                  _compute_batch_keypoint_loss      smal_fitter/neuralSMIL/smil_image_regressor.py:3032-3115
 *   smil_keypoint3d_loss_*      <- _compute_batch_keypoint_3d_loss   smal_fitter/neuralSMIL/smil_image_regressor.py:3334-3407
 *   smil_triangulate_dlt_*      <- _triangulate_joints_dlt           smal_fitter/neuralSMIL/multiview_smil_regressor.py:1707-1838
 *   smil_rot6d_to_rotmat_*      <- _rotation_6d_to_matrix            smal_fitter/neuralSMIL/multiview_smil_regressor.py:250-272
 *                                  rotation_6d_to_axis_angle (1st half) smal_fitter/neuralSMIL/smil_image_regressor.py:87-96
 */
#ifndef SMIL_B200_H
#define SMIL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMIL_ABI_VERSION 7

enum {
  SMIL_OK = 0,
  SMIL_EINVAL = -1,   /* bad argument (NULL, size, flag combination) */
  SMIL_ECUDA = -2,    /* a CUDA runtime call failed; message has the CUDA error string */
  SMIL_ENOMEM = -3,   /* workspace too small / allocation of the packed model failed */
  SMIL_EUNSUPPORTED = -4
};

typedef struct SmilModel SmilModel; /* opaque; device-resident packed constants */

/* Per-model tuning switches (all 0 = the library's defaults, so a zero-initialised desc needs no attention).  They
 * select between measured kernel variants and never change results beyond fp32 rounding order. */
typedef struct SmilTuning {
  int32_t fwd_gp;       /* sample pairs per TMA group of the fused forward skin pass (0 = 2) */
  int32_t fwd_topsync;  /* CTA barrier at the top of every forward group: 0 = by group size, 1 = always, 2 = never */
  int32_t fwd_cps;      /* resident CTAs per SM the forward grid is sized for (0 = 3) */
  int32_t bwd_stages;   /* g_verts row buffers of the backward skin pass: 0/1 = single, 2 = double */
  int32_t bwd_cps;      /* resident CTAs per SM the backward grid is sized for (0 = occupancy limit) */
  int32_t bwd_vm;       /* vertex-major backward skin pass for K <= 2 rigs: 0 = on, 2 = off (item-major) */
  int32_t pose_v2;      /* pose kernel family: 0 = lane-per-sample where eligible (default), 1 = several samples per warp (measured slower), 2 = warp per sample everywhere */
  int32_t pose_lean;    /* log-scale-free backward pose kernel where it applies: 0 = on, 2 = off */
  int32_t bwd_atomic;      /* vertex-major pass adds its partials to a dense dL/dA with vector reductions instead of
                              writing per-(tile, bone) slots: 0 = on, 2 = off (deterministic slots + summing pass) */
  int32_t bwd_slot_reduce; /* sum the partial slots in a streaming pass before the pose backward: 0 = on, 2 = off */
  int32_t bwd_stagger_ns; /* start delay per resident-CTA slot of the backward skin pass (de-phases the CTAs of an SM so that
                             one's TMA wait hides under the others' arithmetic): 0 = default, -1 = none */
} SmilTuning;

/* Host-side description of a rig: the packed tables `smilify_b200/packer.py` derives from
 * exactly the arrays SMAL.__init__ holds (smal_torch.py:120-196).  All pointers are HOST
 * pointers; they are copied to the device during smil_model_create and not retained.
 * Index arrays are int32, bit-exact with kintree_table[0].astype(int32) (smal_torch.py:193). */
typedef struct SmilModelDesc {
  int32_t device;          /* CUDA device ordinal the constants live on */
  int32_t num_verts;       /* V */
  int32_t num_joints;      /* J  = J_regressor.shape[1] */
  int32_t num_betas;       /* nB = shapedirs.shape[0] */
  int32_t num_pose_basis;  /* P: 9*(J-1), or 0 when posedirs is empty / all-zero (smal_torch.py:181-190) */
  int32_t skin_k;          /* K: max non-zero skinning weights per vertex */
  int32_t jreg_nnz;
  int32_t num_levels;      /* tree levels incl. the root level */
  int32_t max_width;       /* widest level */
  int32_t num_pairs;       /* (joint,bone) couplings of the closed-form posed-joint regression */
  int32_t has_static;      /* J_static present */
  int32_t fwd_chunk, fwd_nchunk, v_pad;            /* fused forward pass tiling */
  int32_t bwd_chunk, bwd_ntile, bwd_warps, bwd_rec;/* fused backward pass tiling; bwd_rec = floats per item-pair record */
  int32_t num_units, num_item_pairs;
  int32_t bwd_max_tile_pairs;
  int32_t fused_ok;        /* K, nB, P within the fused kernels' limits */
  /* dense / generic tables */
  const float* v_template;   /* [V][3] */
  const float* shapedirs;    /* [nB][3V]   smal_torch.py:124 */
  const float* posedirs;     /* [P][3V] or NULL */
  const float* J_static;     /* [J][3] or NULL   smal_torch.py:175-176 */
  const int32_t* skin_idx;   /* [V][K] ELL bone ids (padding: id 0, weight 0) */
  const float* skin_w;       /* [V][K] */
  const int32_t* jc_ptr; const int32_t* jc_vid; const float* jc_val; /* J_regressor CSC by joint */
  const int32_t* jr_ptr; const int32_t* jr_jid; const float* jr_val; /* J_regressor CSR by vertex */
  /* kinematic tree */
  const int32_t* parents;    /* [J], root = -1, parents[i] < i */
  const int32_t* order;      /* [J] joints sorted by depth */
  const int32_t* level_ptr;  /* [num_levels+1] into order */
  const int32_t* child_ptr;  /* [J+1] */
  const int32_t* child_idx;  /* [J-1] */
  /* closed-form regressions */
  const float* JT;           /* [J][3]      Jreg^T v_template */
  const float* JS;           /* [nB][J][3]  Jreg^T S_m */
  const float* jreg_rowsum;  /* [J] */
  const int32_t* pj_ptr;     /* [J+1] pairs grouped by joint */
  const int32_t* pj_bone;    /* [np] */
  const int32_t* pj_joint;   /* [np] */
  const float* pj_c0;        /* [np][4] */
  const float* pj_cm;        /* [np][nB][3] */
  const int32_t* pb_ptr;     /* [J+1] the same pairs grouped by bone */
  const int32_t* pb_pair;    /* [np] */
  const float* pbrec;        /* [np padded to 32][(6+3nB) padded to 4] the pairs bone-sorted, one record each:
                                bone (int bits, -1 = padding) | joint (int bits) | c0[4] | cm[nB][3] */
  /* fused forward pass */
  const float* vconst;       /* [3+3nB][v_pad] */
  const int32_t* ell_idx;    /* [K][v_pad] */
  const float* ell_w;        /* [K][v_pad] */
  const int32_t* fwd_pos;    /* [v_pad] vertex (local to its chunk) thread t of the forward pass owns, -1 = none;
                                vconst / ell_* are in THREAD order (packer.py: bone-sorted, bank-conflict-free) */
  /* fused backward pass (item-major; an item is one non-zero (vertex, bone, weight) of the skinning matrix) */
  const float* rec2;         /* [num_item_pairs][bwd_rec] two consecutive items of a unit per record, every constant
                                as an (item, item+1) pair: T (3 pairs) | S (3*NBT pairs) | weight | first-item flag |
                                3*local vertex (int bits) | pad; odd units end with a zero-weight twin */
  const int32_t* unit_bone;  /* [num_units] */
  const int32_t* unit_ptr;   /* [num_units+1] into item pairs */
  const int32_t* unit_tile;  /* [num_units] */
  const int32_t* tile_unit_ptr; /* [bwd_ntile+1] */
  const int32_t* warp_unit_ptr; /* [bwd_ntile*bwd_warps+1] */
  const int32_t* warp_unit;  /* [num_units] */
  const int32_t* tjv_ptr;    /* [bwd_ntile+1] vertices of a tile that carry J_regressor entries ... */
  const int32_t* tjv_v;      /* ... as indices local to the tile */
  const int32_t* sb_ptr;     /* [J+1] bone k owns the g_A partial slots sb_ptr[k] .. sb_ptr[k+1] */
  const int32_t* unit_slot;  /* [num_units] slot a unit's partial is written to (bone-major numbering) */
  /* fused backward pass, vertex-major (rigs with <= 2 weights per vertex; packer.py::pack_backward_vm): a tile's
   * vertices in the order a warp with two register-resident bones walks them, cut into bwd_warps shares */
  int32_t vm_ok, vm_chunk, vm_ntile, vm_recf;   /* vm_recf = floats per vertex record */
  int32_t vm_nslots, vm_nseg, vm_max_tile_recs, vm_max_tile_segs;
  int32_t vm_warps;               /* warps per CTA of the vertex-major pass that share one tile (1, 2 or 4) */
  int32_t vm_nphase;              /* sub-ranges of a tile with their own g_verts row buffer (1..4); vm_chunk % (4 vm_nphase) == 0 */
  const float* vm_rec;            /* [sum of tile records][vm_recf]: wX wY 3*local-vertex(int bits) T[3] S[nB][3] pad */
  const int32_t* vm_seg;          /* [vm_nseg][8]: load X bone|-1, load Y bone|-1, first record (tile-local), records,
                                     flush X slot|-1, flush Y slot|-1, mode (0 both slots, 1 X only), the warp's next segment (tile-local)|-1 */
  const int32_t* vm_tile_rec_ptr; /* [vm_ntile+1] */
  const int32_t* vm_warp_seg_ptr; /* [vm_ntile*vm_nphase*vm_warps+1], index (tile * vm_nphase + phase) * vm_warps + warp */
  const int32_t* vm_sb_ptr;       /* [J+1] bone k owns the vertex-major partial slots vm_sb_ptr[k] .. vm_sb_ptr[k+1] */
  const int32_t* vm_slot_bone;    /* [vm_nslots] bone of every slot (atomic mode adds a flush to its bone's dense entry) */
  SmilTuning tuning;              /* per-model kernel variant switches; zeros = defaults */
} SmilModelDesc;

typedef struct SmilModelInfo {
  int32_t num_verts, num_joints, num_betas, num_pose_basis;
  int32_t skin_k;          /* max non-zero skinning weights per vertex */
  int32_t jreg_nnz;        /* non-zeros of J_regressor */
  int32_t tree_depth;      /* number of levels below the root */
  int32_t max_level_width;
  int32_t posedirs_zero;   /* 1 when posedirs is absent or all-zero */
  int32_t joint_bone_pairs;/* nnz of the closed-form joint regression table */
  int32_t num_chunks;      /* vertex chunks of the fused skinning pass */
  int32_t num_ga_slots;    /* per-sample (chunk,bone) partial slots written by the backward skinning pass */
} SmilModelInfo;

int smil_abi_version(void);
const char* smil_last_error(void);

/* Number of kernels this library has launched since load (all threads). */
long long smil_launch_count(void);

/* Optional per-kernel device timing for bench.py's roofline line: while enabled every kernel
 * launch is bracketed by a cudaEvent pair on the launching stream (up to `max_launches` launches
 * between two collects; 0 disables and frees the events).  Not for use during stream capture. */
typedef struct SmilProfileEntry {
  char name[48];
  int32_t launches;
  float total_ms;
} SmilProfileEntry;
int smil_profile_enable(int max_launches);
int smil_profile_collect(SmilProfileEntry* out, int max_entries, int* num_entries);

int smil_model_create(const SmilModelDesc* desc, SmilModel** out);
void smil_model_destroy(SmilModel* model);
int smil_model_info(const SmilModel* model, SmilModelInfo* info);

/* flags shared by forward/backward */
#define SMIL_F_THETA_IS_ROTMAT    (1u << 0) /* theta is [B][J][3][3]           smal_torch.py:288-289 */
#define SMIL_F_STATIC_JOINTS      (1u << 1) /* config.STATIC_JOINT_LOCATIONS   smal_torch.py:257,343 */
#define SMIL_F_PROPAGATE_SCALING  (1u << 2) /* propagate_scaling=True          batch_lbs.py:163-167 */
#define SMIL_F_JOINTS_ONLY        (1u << 3) /* get_skin=False: verts/v_shaped/Rs outputs may be NULL */
#define SMIL_F_FORCE_GENERIC      (1u << 4) /* testing: take the unfused kernels even when the fused path applies */

/* Per-sample strides are in ELEMENTS between consecutive samples; 0 means the same row is
 * shared by every sample (e.g. SMALFitter's `betas.expand(B, nB)`, fitter.py:244). */
typedef struct SmilFwdArgs {
  int32_t batch;            /* B */
  int32_t num_betas_used;   /* beta.shape[1] <= nB; 0 allowed only with batch==1 (smal_torch.py:250-253) */
  uint32_t flags;
  int32_t _pad;
  const float* beta;            int64_t beta_stride;      /* [B][nb] */
  const float* theta;           int64_t theta_stride;     /* [B][J][3] or [B][J][9] */
  const float* trans;           int64_t trans_stride;     /* [B][3] or NULL (=0) */
  const float* betas_logscale;  int64_t logscale_stride;  /* [B][J][3] or NULL; caller passes NULL when
                                                             config.ALLOW_LIMB_SCALING is False (batch_lbs.py:123) */
  const float* betas_trans;     int64_t btrans_stride;    /* [B][J][3] or NULL */
  const float* del_v;           int64_t del_v_stride;     /* [B][V][3], [V][3] with stride 0, or NULL */
  const float* v_template;                                /* [V][3] override or NULL (smal_torch.py:219-220) */
  /* outputs (contiguous) */
  float* verts;          /* [B][V][3] */
  float* joints;         /* [B][J][3]; static rigs: equals J_transformed (no trans, smal_torch.py:343-346) */
  float* Rs;             /* [B][J][3][3]; written only when theta is axis-angle (else the caller returns theta) */
  float* v_shaped;       /* [B][V][3] */
  float* J_transformed;  /* [B][J][3] posed joint positions from the chain (smal_torch.py:307) */
  /* scratch that must stay alive and untouched until the matching smil_backward has run */
  void* saved;  size_t saved_bytes;     /* >= smil_saved_bytes(model, B, flags) */
  void* scratch; size_t scratch_bytes;  /* >= smil_fwd_scratch_bytes(...); free after the call's stream work */
} SmilFwdArgs;

typedef struct SmilBwdArgs {
  SmilFwdArgs fwd;            /* the same inputs/flags/saved buffer as the forward call (outputs may be NULL) */
  /* upstream gradients, any may be NULL (= zero) */
  const float* g_verts;       /* [B][V][3] */
  const float* g_joints;      /* [B][J][3] */
  const float* g_Rs;          /* [B][J][9] */
  const float* g_v_shaped;    /* [B][V][3] */
  /* input gradients, any may be NULL (= not needed); always per-sample, contiguous */
  float* g_beta;              /* [B][nb] */
  float* g_theta;             /* [B][J][3] or [B][J][9] */
  float* g_trans;             /* [B][3] */
  float* g_logscale;          /* [B][J][3] */
  float* g_btrans;            /* [B][J][3] */
  float* g_del_v;             /* [B][V][3] dL/d(v_shaped) per sample: gradient of del_v; summed over B = gradient of a v_template override */
  void* scratch; size_t scratch_bytes;  /* >= smil_bwd_scratch_bytes(...) */
} SmilBwdArgs;

size_t smil_saved_bytes(const SmilModel* model, int32_t batch, uint32_t flags);
size_t smil_fwd_scratch_bytes(const SmilModel* model, const SmilFwdArgs* args);
size_t smil_bwd_scratch_bytes(const SmilModel* model, const SmilBwdArgs* args);

int smil_forward(const SmilModel* model, const SmilFwdArgs* args, void* stream);
int smil_backward(const SmilModel* model, const SmilBwdArgs* args, void* stream);

/* Stand-alone batch_rodrigues (batch_lbs.py:31-50): theta [N][3] -> R [N][9]; and its VJP. */
int smil_rodrigues_forward(const float* theta, float* R, int64_t n, void* stream);
int smil_rodrigues_backward(const float* theta, const float* g_R, float* g_theta, int64_t n, void* stream);

/* Stand-alone batch_global_rigid_transformation (batch_lbs.py:75-197) over the kinematic tree
 * of `model` (a model created from a desc that carries only the tree tables is enough):
 * Rs [B][J][9], Js [B][J][3] -> new_J [B][J][3], A [B][J][12] (rows 0..2 of the 4x4; row 3 is 0 0 0 1). */
typedef struct SmilChainArgs {
  int32_t batch;
  uint32_t flags;              /* SMIL_F_PROPAGATE_SCALING */
  const float* Rs; const float* Js;
  const float* betas_logscale; /* [B][J][3] or NULL */
  const float* betas_trans;    /* [B][J][3] or NULL */
  float* new_J; float* A;
  /* backward (all optional) */
  const float* g_new_J; const float* g_A;  /* [B][J][3], [B][J][12] */
  float* g_Rs; float* g_Js; float* g_logscale; float* g_btrans;
} SmilChainArgs;
int smil_chain_forward(const SmilModel* model, const SmilChainArgs* args, void* stream);
int smil_chain_backward(const SmilModel* model, const SmilChainArgs* args, void* stream);

/* ---- reprojection + keypoint losses -------------------------------------------------------- */
enum {
  SMIL_LOSS_FITTER = 0,     /* fitter.py:280-296: pixels, invisible -> -1 on both sides, w * mean over N*Jc*2 */
  SMIL_LOSS_MULTIVIEW = 1,  /* multiview_smil_regressor.py:1623-1705, 2319-2388 */
  SMIL_LOSS_SINGLEVIEW = 2  /* smil_image_regressor.py:2727-2732, 3032-3115 */
};

typedef struct SmilProjArgs {
  int32_t kind;           /* SMIL_LOSS_* */
  int32_t batch;          /* B frames */
  int32_t num_views;      /* V cameras per frame (1 for FITTER / SINGLEVIEW) */
  int32_t num_joints;     /* Jc joints per frame after canonical selection */
  float image_size;       /* S: square render size (renderer.image_size) */
  float norm_size;        /* divisor before clamping: S+1e-8 (multiview), 512+1e-8 (single view); unused for FITTER */
  float clamp_lo, clamp_hi; /* (0,1) multiview, (-10,10) single view; unused for FITTER */
  float weight;           /* FITTER: w_j2d */
  int32_t _pad;
  const float* joints;  int64_t joints_stride;   /* [B][Jc][3] world joints (caller has applied +trans / scaling) */
  /* cameras: element strides per frame and per view; 0 = shared */
  const float* cam_R;   int64_t R_bstride, R_vstride;       /* [..][3][3] row-vector convention X@R+T */
  const float* cam_T;   int64_t T_bstride, T_vstride;       /* [..][3] */
  const float* fov_deg; int64_t fov_bstride, fov_vstride;   /* degrees */
  const float* aspect;  int64_t asp_bstride, asp_vstride;   /* or NULL (=1) */
  const float* target;      /* [B][V][Jc][2] in (y,x) order (pixels for FITTER, normalised otherwise) */
  const float* visibility;  /* [B][V][Jc] non-zero = visible */
  const float* view_mask;   /* [B][V] or NULL */
  const float* joint_weights; /* [Jc] or NULL */
  const float* sample_mask; /* [B] or NULL (SINGLEVIEW: samples whose keypoint_data is None are skipped) */
  /* outputs */
  float* proj;         /* [B][V][Jc][2] projected (y,x) after normalise/clamp (FITTER: after the -1 fill) or NULL */
  float* loss;         /* [1] */
  float* partials;     /* [B*V*3 + 4] per (frame,view) (numerator, weight sum, valid count) + 4 scalars;
                          written by forward, read by backward */
  /* backward */
  const float* g_loss; /* [1] device scalar (upstream) */
  float* g_joints;     /* [B][Jc][3] or NULL */
  float* g_cam_R;      /* [B][V][9] or NULL (per frame & view; caller reduces shared cameras) */
  float* g_cam_T;      /* [B][V][3] or NULL */
  float* g_fov;        /* [B][V]    or NULL */
} SmilProjArgs;
int smil_project_loss_forward(const SmilProjArgs* args, void* stream);
int smil_project_loss_backward(const SmilProjArgs* args, void* stream);

typedef struct SmilKp3dArgs {
  int32_t batch, num_joints;
  const float* joints;      /* [B][J][3] */
  const float* target;      /* [B][J][3] */
  const float* visibility;  /* [B][J] */
  const float* joint_weights; /* [J] or NULL */
  const float* sample_mask; /* [B] or NULL */
  float* loss; float* partials; /* [1], [B*3 + 4] */
  const float* g_loss; float* g_joints;
} SmilKp3dArgs;
int smil_keypoint3d_loss_forward(const SmilKp3dArgs* args, void* stream);
/* Allocation-free variant of smil_keypoint3d_loss_forward.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same inputs and the same `loss` output; `args->partials` is IGNORED (may be NULL): the per-sample
 * partials live in a static device array of the library, 3 * SMIL_KP3D_STATIC_MAX_BATCH + 4 floats (192 KB) that every
 * GPU loading the library holds until exit.  batch > SMIL_KP3D_STATIC_MAX_BATCH returns SMIL_EUNSUPPORTED.  The call
 * makes no cudaMalloc / cudaFree / allocating library call; the first call per device resolves the array's device
 * address (cudaGetSymbolAddress) with that device current.  All calls on a device share the array and it keeps the last
 * call's contents: calls that overlap on the GPU (different streams) race; the backward of such a call is not provided
 * (use smil_keypoint3d_loss_forward + _backward with a caller buffer). */
#define SMIL_KP3D_STATIC_MAX_BATCH 16384
int smil_keypoint3d_loss_forward_static(const SmilKp3dArgs* args, void* stream);
int smil_keypoint3d_loss_backward(const SmilKp3dArgs* args, void* stream);

/* ---- DLT triangulation from multi-view 2-D keypoints (multiview_smil_regressor.py:1707-1838) ----------------------
 * Cameras as in SmilProjArgs (PyTorch3D FoV, row vectors, per frame & view strides, 0 = shared).  Per (frame, joint):
 * rows u c3 - c0, v c3 - c1 of every view that sees the joint, w = 1 normalisation, damped (1e-6) 3x3 normal equations;
 * valid where >= 2 views contribute.  Backward: gradients to the cameras (the keypoints are data). */
typedef struct SmilTriArgs {
  int32_t batch, num_views, num_joints;
  float image_size;
  const float* keypoints;   /* [B][V][J][2] normalised (y,x) */
  const float* visibility;  /* [B][V][J] non-zero = visible */
  const float* view_mask;   /* [B][V] or NULL */
  const float* cam_R;   int64_t R_bstride, R_vstride;
  const float* cam_T;   int64_t T_bstride, T_vstride;
  const float* fov_deg; int64_t fov_bstride, fov_vstride;
  const float* aspect;  int64_t asp_bstride, asp_vstride;   /* or NULL (=1) */
  float* tri;           /* [B][J][3] world-space joints */
  float* valid;         /* [B][J] 1.0 where >= 2 views contributed, else 0.0; or NULL */
  /* backward */
  const float* g_tri;   /* [B][J][3] */
  float* g_cam_R;       /* [B][V][9] or NULL */
  float* g_cam_T;       /* [B][V][3] or NULL */
  float* g_fov;         /* [B][V]    or NULL */
} SmilTriArgs;
int smil_triangulate_dlt_forward(const SmilTriArgs* args, void* stream);
int smil_triangulate_dlt_backward(const SmilTriArgs* args, void* stream);

/* Parameter-space terms of the fitter objective, value AND gradient in one call:
 *   joint limits, pose prior, splay, betas prior   smal_fitter/fitter.py:298-331
 *   temporal smoothness                            smal_fitter/fitter.py:337-350 (get_temporal)
 * evaluated per window of `window` consecutive frames and summed over the windows, as optimize_to_joints.py:154-172
 * drives them (every `mean` of fitter.py runs over ONE window).  The frames of a call may be a shard
 * [frame_offset, frame_offset + num_frames) of a sequence of total_frames frames; with w_temp > 0 the parameters of
 * the frames just outside the shard arrive as halos (layout [3 global | 3(J-1) joint | 3 trans]); a missing halo for an
 * existing neighbour is an error.  Masks multiply the parameters (fitter.py:242-243).  pose_precs NULL = identity,
 * pose_mean NULL = zero, pose_use NULL = all but the root's three angles (the prior fitter.py:25-52 builds for rigs
 * without hard-coded body data).  Gradients are written (accumulate = 0) or added (accumulate = 1), scaled by g_scale. */
typedef struct SmilFitterPriorArgs {
  int32_t num_frames, num_joints, num_betas, window;
  int64_t frame_offset, total_frames;
  const float* global_rot;    /* [n][3] */
  const float* joint_rot;     /* [n][J-1][3] */
  const float* trans;         /* [n][3] */
  const float* betas;         /* [nB], shared by the frames (fitter.py:240-241) or NULL */
  const float* halo_prev;     /* [3J + 3] parameters of frame frame_offset - 1, or NULL */
  const float* halo_next;     /* [3J + 3] parameters of frame frame_offset + n, or NULL */
  const float* max_limits;    /* [J-1][3] or NULL */
  const float* min_limits;    /* [J-1][3] or NULL */
  const float* rotation_mask; /* [J-1][3] or NULL */
  const float* global_mask;   /* [3] or NULL */
  const float* pose_mean;     /* [3J] or NULL */
  const float* pose_precs;    /* [3J][3J] or NULL */
  const float* pose_use;      /* [3J] or NULL */
  const float* betas_mean;    /* [nB] or NULL */
  const float* betas_prec;    /* [nB][nB] or NULL (no betas prior) */
  float w_limit, w_pose, w_splay, w_betas, w_temp, g_scale;
  int32_t accumulate;
  float* loss;                /* [1] sum of the terms of this call */
  float* partials;            /* [n] scratch (per-frame values) */
  float* g_global_rot;        /* [n][3] */
  float* g_joint_rot;         /* [n][J-1][3] */
  float* g_trans;             /* [n][3] or NULL */
  float* g_betas;             /* [nB] or NULL */
} SmilFitterPriorArgs;
int smil_fitter_prior_terms(const SmilFitterPriorArgs* args, void* stream);

/* Fused Adam step over up to SMIL_ADAM_MAX_TENSORS parameter tensors (torch.optim.Adam defaults: no weight decay, no
 * amsgrad), the optimiser of smal_fitter/optimize_to_joints.py:117-126, :175.  `step` is a device counter that the call
 * increments BEFORE the update (bias corrections use the incremented value), so that the call can be replayed from a
 * CUDA graph. */
#define SMIL_ADAM_MAX_TENSORS 8
typedef struct SmilAdamArgs {
  int32_t num_tensors;
  float beta1, beta2, eps;
  float* param[SMIL_ADAM_MAX_TENSORS];
  const float* grad[SMIL_ADAM_MAX_TENSORS];
  float* exp_avg[SMIL_ADAM_MAX_TENSORS];
  float* exp_avg_sq[SMIL_ADAM_MAX_TENSORS];
  int64_t numel[SMIL_ADAM_MAX_TENSORS];
  float lr[SMIL_ADAM_MAX_TENSORS];
  int32_t* step;
} SmilAdamArgs;
int smil_adam_step(const SmilAdamArgs* args, void* stream);

/* 6-D rotation representation -> rotation matrix by Gram-Schmidt: d6 [N][6] -> R [N][9] row-major.
 * columns = 0: b1, b2, b3 are the ROWS (pytorch3d.transforms.rotation_6d_to_matrix, smil_image_regressor.py:87-96);
 * columns = 1: the COLUMNS (the reference's own _rotation_6d_to_matrix for cameras, multiview_smil_regressor.py:250-272). */
int smil_rot6d_to_rotmat_forward(const float* d6, float* R, int64_t n, int32_t columns, void* stream);
int smil_rot6d_to_rotmat_backward(const float* d6, const float* g_R, float* g_d6, int64_t n, int32_t columns, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SMIL_B200_H */
