// 2D joint reprojection (PyTorch3D FoV perspective camera, screen space, (y,x) order) fused with
// the three keypoint losses built on it, forward and backward, plus the 3D keypoint loss.
//
// Replaces (paths relative to /root/reference):
//   Renderer.forward projection slice                smal_fitter/p3d_renderer.py:130-140
//   SMALFitter joint loss                            smal_fitter/fitter.py:280-296
//   _batch_project_joints_to_views                   smal_fitter/neuralSMIL/multiview_smil_regressor.py:1623-1705
//   _compute_batched_keypoint_loss                   smal_fitter/neuralSMIL/multiview_smil_regressor.py:2319-2388
//   normalise/clamp + _compute_batch_keypoint_loss   smal_fitter/neuralSMIL/smil_image_regressor.py:2727-2732, 3032-3115
//   _compute_batch_keypoint_3d_loss                  smal_fitter/neuralSMIL/smil_image_regressor.py:3334-3407
//
// Camera algebra (pytorch3d 0.7.x, row vectors):  Xv = X R + T;  f = tan(fov/2);
//   x_ndc = Xv.x / (f * aspect * Xv.z),  y_ndc = Xv.y / (f * Xv.z);
//   screen = S/2 * (1 - ndc)  (NDC->screen with xy flip, square image);  output order (y, x).
#pragma once
#include "smil_common.cuh"

namespace smil {

struct Cam {
  float R[9], T[3], f, fa;  // f = tan(fov/2), fa = f * aspect
};

__device__ __forceinline__ Cam load_cam(const SmilProjArgs& a, int b, int v) {
  Cam c;
  const float* R = a.cam_R + (long long)b * a.R_bstride + (long long)v * a.R_vstride;
  const float* T = a.cam_T + (long long)b * a.T_bstride + (long long)v * a.T_vstride;
#pragma unroll
  for (int q = 0; q < 9; ++q) c.R[q] = R[q];
  c.T[0] = T[0]; c.T[1] = T[1]; c.T[2] = T[2];
  const float fov = a.fov_deg[(long long)b * a.fov_bstride + (long long)v * a.fov_vstride];
  // tan in fp64: an error of a few ulp in f scales EVERY ndc of the view the same way and does not average out over
  // the joints (tanf's 4 ulp showed up as 2e-5 relative on the shared-camera gradients)
  c.f = (float)tan((double)fov * 0.00872664625997164788);  // pi/360
  const float asp = a.aspect ? a.aspect[(long long)b * a.asp_bstride + (long long)v * a.asp_vstride] : 1.0f;
  c.fa = c.f * asp;
  return c;
}

struct Proj {
  float Xv[3], xn, yn;  // view-space point and NDC
  float p0, p1;         // (y, x) screen
};
__device__ __forceinline__ Proj project(const Cam& c, const float* X, float half) {
  Proj r;
  r.Xv[0] = fmaf(X[0], c.R[0], fmaf(X[1], c.R[3], fmaf(X[2], c.R[6], c.T[0])));
  r.Xv[1] = fmaf(X[0], c.R[1], fmaf(X[1], c.R[4], fmaf(X[2], c.R[7], c.T[1])));
  r.Xv[2] = fmaf(X[0], c.R[2], fmaf(X[1], c.R[5], fmaf(X[2], c.R[8], c.T[2])));
  r.xn = r.Xv[0] / (c.fa * r.Xv[2]);
  r.yn = r.Xv[1] / (c.f * r.Xv[2]);
  r.p1 = half - half * r.xn;
  r.p0 = half - half * r.yn;
  return r;
}

// Per-element loss terms shared by forward and backward.
struct Term {
  float q0, q1;      // value entering the squared error (after normalise / clamp / nan_to_num / -1 fill)
  float e0, e1;      // residual q - target, evaluated WITHOUT the ~256 px intermediate: q = k - k ndc (k = S/2 for pixels,
                     // S/2/norm otherwise), so e = (k - t) - k ndc is one fused rounding of a small quantity instead of
                     // the difference of two fp32 pixel coordinates (3x less error on the gradients, PIN_REPORT.json)
  float d0, d1;      // d q / d p   (0 where clamped, non-finite or masked)
  float w;           // mask * joint weight
  bool valid;        // mask alone (a zero joint weight still counts as a valid joint)
};
__device__ __forceinline__ Term loss_term(const SmilProjArgs& a, int b, int v, int j, float half, float xn, float yn) {
  Term t;
  const long long e = ((long long)b * a.num_views + v) * a.num_joints + j;
  const float vis = a.visibility ? a.visibility[e] : 1.f;
  const float t0 = a.target[2 * e], t1 = a.target[2 * e + 1];
  if (a.kind == SMIL_LOSS_FITTER) {
    const bool on = vis != 0.f;
    t.q0 = on ? fmaf(-half, yn, half) : -1.f; t.q1 = on ? fmaf(-half, xn, half) : -1.f;
    t.e0 = on ? fmaf(-half, yn, half - t0) : 0.f; t.e1 = on ? fmaf(-half, xn, half - t1) : 0.f;   // -1 - (-1) = 0
    t.d0 = on ? 1.f : 0.f; t.d1 = t.d0;
    t.w = 1.f;
    t.valid = true;
    return t;
  }
  const float inv = 1.0f / a.norm_size;
  const float k = half * inv;
  float n0 = fmaf(-k, yn, k), n1 = fmaf(-k, xn, k);
  float e0 = fmaf(-k, yn, k - t0), e1 = fmaf(-k, xn, k - t1);
  t.d0 = (n0 >= a.clamp_lo && n0 <= a.clamp_hi) ? inv : 0.f;
  t.d1 = (n1 >= a.clamp_lo && n1 <= a.clamp_hi) ? inv : 0.f;
  if (n0 < a.clamp_lo) { n0 = a.clamp_lo; e0 = n0 - t0; } else if (n0 > a.clamp_hi) { n0 = a.clamp_hi; e0 = n0 - t0; }
  if (n1 < a.clamp_lo) { n1 = a.clamp_lo; e1 = n1 - t1; } else if (n1 > a.clamp_hi) { n1 = a.clamp_hi; e1 = n1 - t1; }
  // NaN propagates through torch.clamp; nan_to_num(nan=0, posinf=0, neginf=0) then zeroes it (infinities were clamped)
  if (!isfinite(n0)) { n0 = 0.f; e0 = -t0; t.d0 = 0.f; }
  if (!isfinite(n1)) { n1 = 0.f; e1 = -t1; t.d1 = 0.f; }
  t.q0 = n0; t.q1 = n1;
  t.e0 = e0; t.e1 = e1;
  bool valid = vis != 0.f && t0 >= 0.f && t0 <= 1.f && t1 >= 0.f && t1 <= 1.f;
  if (a.view_mask) valid = valid && a.view_mask[(long long)b * a.num_views + v] != 0.f;
  if (a.sample_mask) valid = valid && a.sample_mask[b] != 0.f;
  t.w = valid ? (a.joint_weights ? a.joint_weights[j] : 1.f) : 0.f;
  t.valid = valid;
  return t;
}

// forward: one warp per (frame, view); lanes over joints; writes proj and the (num, wsum) partial.
__global__ void proj_fwd_kernel(SmilProjArgs a) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)a.batch * a.num_views) return;
  const int v = (int)(w % a.num_views), b = (int)(w / a.num_views);
  const Cam c = load_cam(a, b, v);
  const float half = 0.5f * a.image_size;
  float num = 0.f, ws = 0.f, cnt = 0.f;
  for (int j = lane; j < a.num_joints; j += 32) {
    const float* X = a.joints + (long long)b * a.joints_stride + 3 * j;
    const Proj pr = project(c, X, half);
    const Term t = loss_term(a, b, v, j, half, pr.xn, pr.yn);
    if (a.proj) {
      const long long e = ((long long)b * a.num_views + v) * a.num_joints + j;
      a.proj[2 * e] = t.q0; a.proj[2 * e + 1] = t.q1;
    }
    num = fmaf(t.w, fmaf(t.e0, t.e0, t.e1 * t.e1), num);
    ws += t.w;
    cnt += t.valid ? 1.f : 0.f;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    ws += __shfl_xor_sync(0xffffffffu, ws, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) { a.partials[3 * w] = num; a.partials[3 * w + 1] = ws; a.partials[3 * w + 2] = cnt; }
}

// finalize: single block, fixed-order reduction of the partials -> loss and the scalars backward needs.
// tail layout after the n*3 partials: [0] = 1/den (FITTER, MULTIVIEW) or 1/nvalid (SINGLEVIEW), [1] = any valid
__global__ void proj_finalize_kernel(SmilProjArgs a) {
  __shared__ double s_num[256], s_ws[256], s_cnt[256];
  const long long n = (long long)a.batch * a.num_views;
  const float eps = 1e-8f;
  double num = 0.0, ws = 0.0, cnt = 0.0;
  for (long long i = threadIdx.x; i < n; i += 256) {
    const float pn = a.partials[3 * i], pw = a.partials[3 * i + 1], pc = a.partials[3 * i + 2];
    if (a.kind == SMIL_LOSS_SINGLEVIEW) {
      if (pc > 0.f) { num += (double)(pn / (pw * 2.f + eps)); cnt += 1.0; }
    } else {
      num += pn; ws += pw; cnt += pc;
    }
  }
  s_num[threadIdx.x] = num; s_ws[threadIdx.x] = ws; s_cnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      s_num[threadIdx.x] += s_num[threadIdx.x + o];
      s_ws[threadIdx.x] += s_ws[threadIdx.x + o];
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float* tail = a.partials + 3 * n;
    float loss, scale, any = 1.f;
    if (a.kind == SMIL_LOSS_FITTER) {
      const float cntall = (float)n * a.num_joints * 2.f;
      loss = a.weight * (float)(s_num[0] / cntall);
      scale = a.weight / cntall;
    } else if (a.kind == SMIL_LOSS_MULTIVIEW) {
      if (s_cnt[0] > 0.0) {
        const float den = (float)s_ws[0] * 2.f + eps;
        loss = (float)s_num[0] / den + eps;
        scale = 1.f / den;
      } else { loss = eps; scale = 0.f; any = 0.f; }
    } else {
      if (s_cnt[0] > 0.0) { loss = (float)(s_num[0] / s_cnt[0]) + eps; scale = (float)(1.0 / s_cnt[0]); }
      else { loss = eps; scale = 0.f; any = 0.f; }
    }
    a.loss[0] = loss;
    tail[0] = scale; tail[1] = any;
  }
}

// d loss / d p for one element, given the upstream scalar and the finalize scalars
__device__ __forceinline__ void term_grad(const SmilProjArgs& a, const Term& t, float gl, float scale, long long w,
                                          float& g0, float& g1) {
  float s = gl * scale * 2.f * t.w;
  if (a.kind == SMIL_LOSS_SINGLEVIEW) {
    const float pw = a.partials[3 * w + 1], pc = a.partials[3 * w + 2];
    s = pc > 0.f ? s / (pw * 2.f + 1e-8f) : 0.f;
  }
  g0 = s * t.e0 * t.d0;
  g1 = s * t.e1 * t.d1;
}

// (g_p0, g_p1) -> gradient on the view-space point and on f = tan(fov/2)
__device__ __forceinline__ void proj_vjp(const Cam& c, const Proj& pr, float half, float g0, float g1, float* gXv,
                                         float& gf) {
  const float gxn = -half * g1, gyn = -half * g0;
  const float iz = 1.0f / pr.Xv[2];
  gXv[0] = gxn / (c.fa * pr.Xv[2]);
  gXv[1] = gyn / (c.f * pr.Xv[2]);
  gXv[2] = -(gxn * pr.xn + gyn * pr.yn) * iz;
  gf = -(gxn * pr.xn + gyn * pr.yn) / c.f;
}

// backward: one block per frame.  Pass A: thread per joint loops views -> g_joints (no atomics).
// Pass B: thread per view loops joints -> camera gradients (no atomics).
__global__ void proj_bwd_kernel(SmilProjArgs a) {
  const int b = blockIdx.x;
  const long long n = (long long)a.batch * a.num_views;
  const float* tail = a.partials + 3 * n;
  const float scale = tail[0];
  const float gl = a.g_loss ? a.g_loss[0] : 1.f;
  const float half = 0.5f * a.image_size;
  if (a.g_joints) {
    for (int j = threadIdx.x; j < a.num_joints; j += blockDim.x) {
      const float* X = a.joints + (long long)b * a.joints_stride + 3 * j;
      double gx = 0.0, gy = 0.0, gz = 0.0;   // fp64 sums: the terms are differences of ~256 px coordinates and cancel
      for (int v = 0; v < a.num_views; ++v) {
        const Cam c = load_cam(a, b, v);
        const Proj pr = project(c, X, half);
        const Term t = loss_term(a, b, v, j, half, pr.xn, pr.yn);
        float g0, g1, gXv[3], gf;
        term_grad(a, t, gl, scale, (long long)b * a.num_views + v, g0, g1);
        if (g0 == 0.f && g1 == 0.f) continue;
        proj_vjp(c, pr, half, g0, g1, gXv, gf);
        gx += (double)(c.R[0] * gXv[0] + c.R[1] * gXv[1] + c.R[2] * gXv[2]);
        gy += (double)(c.R[3] * gXv[0] + c.R[4] * gXv[1] + c.R[5] * gXv[2]);
        gz += (double)(c.R[6] * gXv[0] + c.R[7] * gXv[1] + c.R[8] * gXv[2]);
      }
      float* o = a.g_joints + ((long long)b * a.num_joints + j) * 3;
      o[0] = (float)gx; o[1] = (float)gy; o[2] = (float)gz;
    }
  }
  if (a.g_cam_R || a.g_cam_T || a.g_fov) {
    for (int v = threadIdx.x; v < a.num_views; v += blockDim.x) {
      const Cam c = load_cam(a, b, v);
      double gR[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, gT[3] = {0, 0, 0}, gfs = 0.0;   // fp64 sums over the joints (see above)
      for (int j = 0; j < a.num_joints; ++j) {
        const float* X = a.joints + (long long)b * a.joints_stride + 3 * j;
        const Proj pr = project(c, X, half);
        const Term t = loss_term(a, b, v, j, half, pr.xn, pr.yn);
        float g0, g1, gXv[3], gf;
        term_grad(a, t, gl, scale, (long long)b * a.num_views + v, g0, g1);
        if (g0 == 0.f && g1 == 0.f) continue;
        proj_vjp(c, pr, half, g0, g1, gXv, gf);
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int q = 0; q < 3; ++q) gR[3 * r + q] += (double)X[r] * (double)gXv[q];
        gT[0] += (double)gXv[0]; gT[1] += (double)gXv[1]; gT[2] += (double)gXv[2];
        gfs += (double)gf;
      }
      const long long w = (long long)b * a.num_views + v;
      if (a.g_cam_R) {
#pragma unroll
        for (int q = 0; q < 9; ++q) a.g_cam_R[9 * w + q] = (float)gR[q];
      }
      if (a.g_cam_T) { a.g_cam_T[3 * w] = (float)gT[0]; a.g_cam_T[3 * w + 1] = (float)gT[1]; a.g_cam_T[3 * w + 2] = (float)gT[2]; }
      // f = tan(fov * pi/360): df/dfov = pi/360 * (1 + f^2); the aspect factor cancels in d/df of x_ndc
      if (a.g_fov) a.g_fov[w] = (float)(gfs * 0.00872664625997164788 * (1.0 + (double)c.f * (double)c.f));
    }
  }
}

// ---- 3D keypoint loss (smil_image_regressor.py:3334-3407) -------------------------------------------
__device__ __forceinline__ float kp3d_weight(const SmilKp3dArgs& a, int b, int j, const float* p, const float* t) {
  const long long e = (long long)b * a.num_joints + j;
  bool valid = (a.visibility ? a.visibility[e] != 0.f : true) && isfinite(p[0]) && isfinite(p[1]) && isfinite(p[2]) &&
               isfinite(t[0]) && isfinite(t[1]) && isfinite(t[2]) && (fabsf(t[0]) + fabsf(t[1]) + fabsf(t[2]) > 0.f);
  if (a.sample_mask) valid = valid && a.sample_mask[b] != 0.f;
  return valid ? (a.joint_weights ? a.joint_weights[j] : 1.f) : 0.f;
}
__global__ void kp3d_fwd_kernel(SmilKp3dArgs a) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= a.batch) return;
  const int b = (int)w;
  float num = 0.f, ws = 0.f, cnt = 0.f;
  for (int j = lane; j < a.num_joints; j += 32) {
    const float* p = a.joints + ((long long)b * a.num_joints + j) * 3;
    const float* t = a.target + ((long long)b * a.num_joints + j) * 3;
    const float wt = kp3d_weight(a, b, j, p, t);
    if (a.sample_mask && a.sample_mask[b] == 0.f) continue;   // keypoint_data is None: the reference skips the sample
    // the reference multiplies the squared error by the 0/1 mask (smil_image_regressor.py:3383-3391), so a masked-out
    // NON-FINITE joint or target still turns the sample's loss into NaN (inf * 0); the same arithmetic here
    const float e0 = p[0] - t[0], e1 = p[1] - t[1], e2 = p[2] - t[2];
    num = fmaf(wt, e0 * e0 + e1 * e1 + e2 * e2, num);
    if (wt != 0.f) cnt += 1.f;
    ws += wt;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    ws += __shfl_xor_sync(0xffffffffu, ws, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) { a.partials[3 * w] = num; a.partials[3 * w + 1] = ws; a.partials[3 * w + 2] = cnt; }
}
__global__ void kp3d_finalize_kernel(SmilKp3dArgs a) {
  __shared__ double s_num[256], s_cnt[256];
  double num = 0.0, cnt = 0.0;
  for (int i = threadIdx.x; i < a.batch; i += 256) {
    const float pn = a.partials[3 * i], pw = a.partials[3 * i + 1], pc = a.partials[3 * i + 2];
    if (pc > 0.f) { num += (double)(pn / (pw * 3.f + 1e-8f)); cnt += 1.0; }
  }
  s_num[threadIdx.x] = num; s_cnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { s_num[threadIdx.x] += s_num[threadIdx.x + o]; s_cnt[threadIdx.x] += s_cnt[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float* tail = a.partials + 3LL * a.batch;
    if (s_cnt[0] > 0.0) { a.loss[0] = (float)(s_num[0] / s_cnt[0]) + 1e-8f; tail[0] = (float)(1.0 / s_cnt[0]); }
    else { a.loss[0] = 1e-8f; tail[0] = 0.f; }
  }
}
__global__ void kp3d_bwd_kernel(SmilKp3dArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)a.batch * a.num_joints) return;
  const int j = (int)(i % a.num_joints), b = (int)(i / a.num_joints);
  const float* p = a.joints + i * 3;
  const float* t = a.target + i * 3;
  const float wt = kp3d_weight(a, b, j, p, t);
  const float pw = a.partials[3 * b + 1], pc = a.partials[3 * b + 2];
  const float scale = a.partials[3LL * a.batch];
  const float gl = a.g_loss ? a.g_loss[0] : 1.f;
  float s = 0.f;
  if (pc > 0.f && wt != 0.f) s = gl * scale * 2.f * wt / (pw * 3.f + 1e-8f);
  if (pc <= 0.f || (a.sample_mask && a.sample_mask[b] == 0.f)) {   // sample not in the reference's graph at all
    a.g_joints[i * 3] = 0.f; a.g_joints[i * 3 + 1] = 0.f; a.g_joints[i * 3 + 2] = 0.f;
    return;
  }
  a.g_joints[i * 3] = s * (p[0] - t[0]);
  a.g_joints[i * 3 + 1] = s * (p[1] - t[1]);
  a.g_joints[i * 3 + 2] = s * (p[2] - t[2]);
}

}  // namespace smil
