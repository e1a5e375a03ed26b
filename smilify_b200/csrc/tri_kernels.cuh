// DLT triangulation of joints from multi-view 2-D keypoints, and 6-D rotations -> rotation matrices, forward and backward.
//
// Replaces (paths relative to /root/reference):
//   _triangulate_joints_dlt                    smal_fitter/neuralSMIL/multiview_smil_regressor.py:1707-1838
//   _rotation_6d_to_matrix (cameras, columns)  smal_fitter/neuralSMIL/multiview_smil_regressor.py:250-272
//   rotation_6d_to_axis_angle's first half     smal_fitter/neuralSMIL/smil_image_regressor.py:87-96
//       (pytorch3d.transforms.rotation_6d_to_matrix, rows; its output feeds SMIL_F_THETA_IS_ROTMAT directly, which skips
//        the reference's matrix -> axis-angle -> matrix round trip, smal_torch.py:288-289)
//
// Triangulation algebra (PyTorch3D FoV camera, row vectors, P = world->NDC 4x4): with c_k = column k of P,
//   c0 = sx (R00 R10 R20 T0), c1 = sy (R01 R11 R21 T1), c3 = (R02 R12 R22 T2), sx = 1 / (tan(fov/2) aspect), sy = 1 / tan(fov/2),
// a visible (view, joint) contributes the rows  u c3 - c0,  v c3 - c1  (u, v = NDC of the keypoint);  with w = 1 the 3x3
// system (sum a a^T + 1e-6 I) p = - sum a_w a  is solved per (frame, joint).  The sums and the solve run in fp64 (the
// normal equations square the condition number; the work is ~30 V fused multiply-adds per joint).
#pragma once
#include "smil_common.cuh"

namespace smil {

struct TriCam {
  float c0[4], c1[4], c3[4], sx, sy, th;
};
__device__ __forceinline__ TriCam tri_cam(const SmilTriArgs& a, int b, int v) {
  TriCam c;
  const float* R = a.cam_R + (long long)b * a.R_bstride + (long long)v * a.R_vstride;
  const float* T = a.cam_T + (long long)b * a.T_bstride + (long long)v * a.T_vstride;
  const float fov = a.fov_deg[(long long)b * a.fov_bstride + (long long)v * a.fov_vstride];
  const float asp = a.aspect ? a.aspect[(long long)b * a.asp_bstride + (long long)v * a.asp_vstride] : 1.0f;
  c.th = (float)tan((double)fov * 0.00872664625997164788);  // pi/360, fp64 (see proj_kernels.cuh::load_cam)
  c.sx = 1.0f / (c.th * asp);
  c.sy = 1.0f / c.th;
  c.c0[0] = c.sx * R[0]; c.c0[1] = c.sx * R[3]; c.c0[2] = c.sx * R[6]; c.c0[3] = c.sx * T[0];
  c.c1[0] = c.sy * R[1]; c.c1[1] = c.sy * R[4]; c.c1[2] = c.sy * R[7]; c.c1[3] = c.sy * T[1];
  c.c3[0] = R[2]; c.c3[1] = R[5]; c.c3[2] = R[8]; c.c3[3] = T[2];
  return c;
}
__device__ __forceinline__ bool tri_visible(const SmilTriArgs& a, int b, int v, int j) {
  bool on = a.visibility[((long long)b * a.num_views + v) * a.num_joints + j] != 0.f;
  if (a.view_mask) on = on && a.view_mask[(long long)b * a.num_views + v] != 0.f;
  return on;
}
__device__ __forceinline__ void tri_uv(const SmilTriArgs& a, int b, int v, int j, float& u, float& w) {
  const float* kp = a.keypoints + (((long long)b * a.num_views + v) * a.num_joints + j) * 2;   // (y, x) normalised
  const float hs = a.image_size * 0.5f;
  u = 1.0f - (kp[1] * a.image_size) / hs;
  w = 1.0f - (kp[0] * a.image_size) / hs;
}
// symmetric 3x3 solve by the adjugate in fp64 (M is positive definite thanks to the damping)
__device__ __forceinline__ void solve_sym3(const double* M, const double* r, double* x) {
  const double a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5];   // [a b c; b d e; c e f]
  const double A = d * f - e * e, Bc = c * e - b * f, C = b * e - c * d;
  const double det = a * A + b * Bc + c * C;
  const double D = a * f - c * c, E = b * c - a * e, F = a * d - b * b;
  const double id = 1.0 / det;
  x[0] = (A * r[0] + Bc * r[1] + C * r[2]) * id;
  x[1] = (Bc * r[0] + D * r[1] + E * r[2]) * id;
  x[2] = (C * r[0] + E * r[1] + F * r[2]) * id;
}
__device__ __forceinline__ int tri_system(const SmilTriArgs& a, int b, int j, double* M, double* r) {
  for (int q = 0; q < 6; ++q) M[q] = 0.0;
  r[0] = r[1] = r[2] = 0.0;
  int count = 0;
  for (int v = 0; v < a.num_views; ++v) {
    if (!tri_visible(a, b, v, j)) continue;
    ++count;
    const TriCam c = tri_cam(a, b, v);
    float u, w;
    tri_uv(a, b, v, j, u, w);
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const float s = rr ? w : u;
      const float* ck = rr ? c.c1 : c.c0;
      const double x = (double)(s * c.c3[0] - ck[0]), y = (double)(s * c.c3[1] - ck[1]), z = (double)(s * c.c3[2] - ck[2]),
                   aw = (double)(s * c.c3[3] - ck[3]);
      M[0] += x * x; M[1] += x * y; M[2] += x * z; M[3] += y * y; M[4] += y * z; M[5] += z * z;
      r[0] -= aw * x; r[1] -= aw * y; r[2] -= aw * z;
    }
  }
  M[0] += 1e-6; M[3] += 1e-6; M[5] += 1e-6;
  return count;
}

__global__ void tri_fwd_kernel(SmilTriArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)a.batch * a.num_joints) return;
  const int j = (int)(i % a.num_joints), b = (int)(i / a.num_joints);
  double M[6], r[3], p[3];
  const int count = tri_system(a, b, j, M, r);
  solve_sym3(M, r, p);
  a.tri[3 * i] = (float)p[0]; a.tri[3 * i + 1] = (float)p[1]; a.tri[3 * i + 2] = (float)p[2];
  if (a.valid) a.valid[i] = count >= 2 ? 1.f : 0.f;
}

// backward: one block per frame.  Pass A: thread per joint -> lambda = M^-1 g and p into shared memory.
// Pass B: thread per view loops the joints -> camera gradients (no atomics).
__global__ void tri_bwd_kernel(SmilTriArgs a) {
  extern __shared__ float s_lp[];   // [J][6]: lambda (3), p (3)
  const int b = blockIdx.x;
  for (int j = threadIdx.x; j < a.num_joints; j += blockDim.x) {
    double M[6], r[3], p[3], lam[3], g[3];
    tri_system(a, b, j, M, r);
    solve_sym3(M, r, p);
    const float* gp = a.g_tri + ((long long)b * a.num_joints + j) * 3;
    g[0] = gp[0]; g[1] = gp[1]; g[2] = gp[2];
    solve_sym3(M, g, lam);
    float* o = s_lp + 6 * j;
    o[0] = (float)lam[0]; o[1] = (float)lam[1]; o[2] = (float)lam[2];
    o[3] = (float)p[0]; o[4] = (float)p[1]; o[5] = (float)p[2];
  }
  __syncthreads();
  for (int v = threadIdx.x; v < a.num_views; v += blockDim.x) {
    const TriCam c = tri_cam(a, b, v);
    double g0[4] = {0, 0, 0, 0}, g1[4] = {0, 0, 0, 0}, g3[4] = {0, 0, 0, 0};
    for (int j = 0; j < a.num_joints; ++j) {
      if (!tri_visible(a, b, v, j)) continue;
      const float* lp = s_lp + 6 * j;
      const double lx = lp[0], ly = lp[1], lz = lp[2], px = lp[3], py = lp[4], pz = lp[5];
      float u, w;
      tri_uv(a, b, v, j, u, w);
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const float s = rr ? w : u;
        const float* ck = rr ? c.c1 : c.c0;
        const double x = (double)(s * c.c3[0] - ck[0]), y = (double)(s * c.c3[1] - ck[1]), z = (double)(s * c.c3[2] - ck[2]),
                     aw = (double)(s * c.c3[3] - ck[3]);
        const double ap = x * px + y * py + z * pz, la = x * lx + y * ly + z * lz;
        // dL/da = -(a.p) lambda - (lambda.a) p - a_w lambda;  dL/da_w = -(lambda.a)
        const double gr[4] = {-(ap + aw) * lx - la * px, -(ap + aw) * ly - la * py, -(ap + aw) * lz - la * pz, -la};
        double* gk = rr ? g1 : g0;
#pragma unroll
        for (int q = 0; q < 4; ++q) { g3[q] += (double)s * gr[q]; gk[q] -= gr[q]; }
      }
    }
    const float* R = a.cam_R + (long long)b * a.R_bstride + (long long)v * a.R_vstride;
    const float* T = a.cam_T + (long long)b * a.T_bstride + (long long)v * a.T_vstride;
    const long long w_ = (long long)b * a.num_views + v;
    const double gsx = g0[0] * R[0] + g0[1] * R[3] + g0[2] * R[6] + g0[3] * T[0];
    const double gsy = g1[0] * R[1] + g1[1] * R[4] + g1[2] * R[7] + g1[3] * T[1];
    if (a.g_cam_R) {
      float* o = a.g_cam_R + 9 * w_;
      for (int i = 0; i < 3; ++i) {
        o[3 * i + 0] = (float)(c.sx * g0[i]);
        o[3 * i + 1] = (float)(c.sy * g1[i]);
        o[3 * i + 2] = (float)g3[i];
      }
    }
    if (a.g_cam_T) {
      a.g_cam_T[3 * w_] = (float)(c.sx * g0[3]); a.g_cam_T[3 * w_ + 1] = (float)(c.sy * g1[3]); a.g_cam_T[3 * w_ + 2] = (float)g3[3];
    }
    if (a.g_fov) {   // sx = 1 / (th aspect), sy = 1 / th, th = tan(fov pi / 360)
      const double gth = -(gsx * c.sx + gsy * c.sy) / (double)c.th;
      a.g_fov[w_] = (float)(gth * 0.00872664625997164788 * (1.0 + (double)c.th * (double)c.th));
    }
  }
}

// ---- 6-D rotation representation -> rotation matrix (Gram-Schmidt), rows (PyTorch3D) or columns (the reference's own) ----
__device__ __forceinline__ float norm_eps(float x, float y, float z) { return fmaxf(sqrtf(fmaf(x, x, fmaf(y, y, z * z))), 1e-12f); }
__global__ void rot6d_fwd_kernel(const float* __restrict__ d6, float* __restrict__ R, long long n, int columns) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* a = d6 + 6 * i;
  const float n1 = norm_eps(a[0], a[1], a[2]);
  const f3 b1 = mk3(a[0] / n1, a[1] / n1, a[2] / n1);
  const float c = b1.x * a[3] + b1.y * a[4] + b1.z * a[5];
  const f3 t = mk3(a[3] - c * b1.x, a[4] - c * b1.y, a[5] - c * b1.z);
  const float n2 = norm_eps(t.x, t.y, t.z);
  const f3 b2 = mk3(t.x / n2, t.y / n2, t.z / n2);
  const f3 b3 = mk3(b1.y * b2.z - b1.z * b2.y, b1.z * b2.x - b1.x * b2.z, b1.x * b2.y - b1.y * b2.x);
  float* o = R + 9 * i;
  if (columns) {
    o[0] = b1.x; o[1] = b2.x; o[2] = b3.x; o[3] = b1.y; o[4] = b2.y; o[5] = b3.y; o[6] = b1.z; o[7] = b2.z; o[8] = b3.z;
  } else {
    o[0] = b1.x; o[1] = b1.y; o[2] = b1.z; o[3] = b2.x; o[4] = b2.y; o[5] = b2.z; o[6] = b3.x; o[7] = b3.y; o[8] = b3.z;
  }
}
__global__ void rot6d_bwd_kernel(const float* __restrict__ d6, const float* __restrict__ gR, float* __restrict__ gd6, long long n,
                                 int columns) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* a = d6 + 6 * i;
  const float* g = gR + 9 * i;
  f3 g1, g2, g3;
  if (columns) { g1 = mk3(g[0], g[3], g[6]); g2 = mk3(g[1], g[4], g[7]); g3 = mk3(g[2], g[5], g[8]); }
  else { g1 = mk3(g[0], g[1], g[2]); g2 = mk3(g[3], g[4], g[5]); g3 = mk3(g[6], g[7], g[8]); }
  const float r1 = sqrtf(fmaf(a[0], a[0], fmaf(a[1], a[1], a[2] * a[2])));
  const float n1 = fmaxf(r1, 1e-12f);
  const f3 b1 = mk3(a[0] / n1, a[1] / n1, a[2] / n1);
  const f3 a2 = mk3(a[3], a[4], a[5]);
  const float c = b1.x * a2.x + b1.y * a2.y + b1.z * a2.z;
  const f3 t = mk3(a2.x - c * b1.x, a2.y - c * b1.y, a2.z - c * b1.z);
  const float r2 = sqrtf(fmaf(t.x, t.x, fmaf(t.y, t.y, t.z * t.z)));
  const float n2 = fmaxf(r2, 1e-12f);
  const f3 b2 = mk3(t.x / n2, t.y / n2, t.z / n2);
  // b3 = b1 x b2
  g1 = mk3(g1.x + (b2.y * g3.z - b2.z * g3.y), g1.y + (b2.z * g3.x - b2.x * g3.z), g1.z + (b2.x * g3.y - b2.y * g3.x));
  g2 = mk3(g2.x + (g3.y * b1.z - g3.z * b1.y), g2.y + (g3.z * b1.x - g3.x * b1.z), g2.z + (g3.x * b1.y - g3.y * b1.x));
  // b2 = t / max(|t|, eps)
  f3 gt;
  if (r2 > 1e-12f) {
    const float d = b2.x * g2.x + b2.y * g2.y + b2.z * g2.z;
    gt = mk3((g2.x - b2.x * d) / n2, (g2.y - b2.y * d) / n2, (g2.z - b2.z * d) / n2);
  } else {
    gt = mk3(g2.x / n2, g2.y / n2, g2.z / n2);
  }
  // t = a2 - c b1, c = b1 . a2
  const float gc = -(gt.x * b1.x + gt.y * b1.y + gt.z * b1.z);
  const f3 ga2 = mk3(gt.x + gc * b1.x, gt.y + gc * b1.y, gt.z + gc * b1.z);
  g1 = mk3(g1.x - c * gt.x + gc * a2.x, g1.y - c * gt.y + gc * a2.y, g1.z - c * gt.z + gc * a2.z);
  // b1 = a1 / max(|a1|, eps)
  f3 ga1;
  if (r1 > 1e-12f) {
    const float d = b1.x * g1.x + b1.y * g1.y + b1.z * g1.z;
    ga1 = mk3((g1.x - b1.x * d) / n1, (g1.y - b1.y * d) / n1, (g1.z - b1.z * d) / n1);
  } else {
    ga1 = mk3(g1.x / n1, g1.y / n1, g1.z / n1);
  }
  float* o = gd6 + 6 * i;
  o[0] = ga1.x; o[1] = ga1.y; o[2] = ga1.z; o[3] = ga2.x; o[4] = ga2.y; o[5] = ga2.z;
}

}  // namespace smil
