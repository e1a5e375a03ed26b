// Parameter-space terms of the fitter objective and the fused Adam update (first half).
//
// Replaces, for a batch of frames resident on the device:
//   joint limits, pose prior, splay, betas prior      /root/reference/smal_fitter/fitter.py:298-331
//   temporal smoothness (get_temporal)                /root/reference/smal_fitter/fitter.py:337-350
//   torch.optim.Adam over the fitter's parameters     /root/reference/smal_fitter/optimize_to_joints.py:117-126, :175
// The objective is evaluated per WINDOW of config.WINDOW_SIZE consecutive frames and summed over the windows
// (optimize_to_joints.py:154-157): every `mean` of fitter.py runs over one window, so a frame's weight depends on how
// many frames its window has (the last window of a sequence may be short).  Frames may be a shard of a longer sequence:
// frame_offset / total_frames place them, and the two neighbours across the shard boundary arrive as halos (the 1-frame
// halo of SURVEY.md section 8(e)).  One kernel computes value AND gradient (the terms are elementwise or tiny matvecs,
// HBM-bound on the parameters: ~2 x 4 (3J+3) bytes per frame in, the same out).
#pragma once
#include "smil_common.cuh"

namespace smil {

constexpr int FIT_THREADS = 128;

// frames in the window that holds global frame f
__device__ __forceinline__ int fit_window_frames(long long f, int window, long long total) {
  if (window <= 0) return (int)total;
  const long long w0 = (f / window) * window;
  return (int)min((long long)window, total - w0);
}

// block per frame.  Shared memory: x[P] (masked pose), r[P] (pose-prior residual * use^2), red[FIT_THREADS].
__global__ void __launch_bounds__(FIT_THREADS) fitter_prior_kernel(SmilFitterPriorArgs a) {
  extern __shared__ float sm_fit[];
  const int J = a.num_joints, P = 3 * J, NJ3 = P - 3;
  float* x = sm_fit;
  float* r = sm_fit + P;
  float* red = r + P;
  const int f = blockIdx.x, tid = threadIdx.x;
  const long long gf = a.frame_offset + f;
  const int nw = fit_window_frames(gf, a.window, a.total_frames);
  const float* gr = a.global_rot + (long long)f * 3;
  const float* jr = a.joint_rot + (long long)f * NJ3;
  for (int i = tid; i < P; i += FIT_THREADS) {
    const float m = (i < 3) ? (a.global_mask ? a.global_mask[i] : 1.f) : (a.rotation_mask ? a.rotation_mask[i - 3] : 1.f);
    x[i] = ((i < 3) ? gr[i] : jr[i - 3]) * m;
  }
  __syncthreads();
  float val = 0.f;
  // pose prior: res = ((x - mean) precs) * use;  value w/(nw P) sum res^2;  dL/dx = 2 w/(nw P) precs (use^2 res)
  const bool pose = a.w_pose > 0.f;
  const float cp = pose ? a.w_pose / ((float)nw * (float)P) : 0.f;
  if (pose) {
    for (int k = tid; k < P; k += FIT_THREADS) {
      float res;
      if (a.pose_precs) {
        res = 0.f;
        for (int i = 0; i < P; ++i) res = fmaf(x[i] - (a.pose_mean ? a.pose_mean[i] : 0.f), a.pose_precs[(size_t)i * P + k], res);
      } else {
        res = x[k] - (a.pose_mean ? a.pose_mean[k] : 0.f);
      }
      const float u = a.pose_use ? a.pose_use[k] : (k < 3 ? 0.f : 1.f);
      res *= u;
      val = fmaf(cp * res, res, val);
      r[k] = res * u;
    }
  }
  __syncthreads();
  const float cl = (a.w_limit > 0.f && a.max_limits && a.min_limits) ? a.w_limit / ((float)nw * (float)NJ3) : 0.f;
  const float ct_g = a.w_temp > 0.f ? a.w_temp / 3.f : 0.f, ct_j = a.w_temp > 0.f ? a.w_temp / (float)NJ3 : 0.f;
  // temporal neighbours (masked like x): global frames gf-1 / gf+1 from the shard or from the halos
  const bool has_prev = a.w_temp > 0.f && gf > 0, has_next = a.w_temp > 0.f && gf + 1 < a.total_frames;
  // (halo layout: [3 global | NJ3 joint | 3 trans], i.e. element i of the loop below)
  const bool prev_in = f > 0, next_in = f + 1 < a.num_frames;
  for (int i = tid; i < P + 3; i += FIT_THREADS) {
    // element i: 0..2 global rotation, 3..P-1 joint rotations, P..P+2 translation
    float g = 0.f;
    float m = 1.f, xi;
    if (i < P) {
      m = (i < 3) ? (a.global_mask ? a.global_mask[i] : 1.f) : (a.rotation_mask ? a.rotation_mask[i - 3] : 1.f);
      xi = x[i];
    } else {
      xi = a.trans[(long long)f * 3 + (i - P)];
    }
    if (i < P && pose) {   // pose prior gradient
      float s;
      if (a.pose_precs) {
        s = 0.f;
        for (int k = 0; k < P; ++k) s = fmaf(a.pose_precs[(size_t)i * P + k], r[k], s);
      } else {
        s = r[i];
      }
      g = fmaf(2.f * cp, s, g);
    }
    if (i >= 3 && i < P) {
      const int e = i - 3;
      if (cl > 0.f) {     // joint limits: mean(max(x - hi, 0) + max(lo - x, 0)); ties split like torch.max(a, b)
        const float hi = xi - a.max_limits[e], lo = a.min_limits[e] - xi;
        val = fmaf(cl, fmaxf(hi, 0.f) + fmaxf(lo, 0.f), val);
        g += cl * ((hi > 0.f ? 1.f : (hi == 0.f ? 0.5f : 0.f)) - (lo > 0.f ? 1.f : (lo == 0.f ? 0.5f : 0.f)));
      }
      if (a.w_splay > 0.f && (e % 3) != 1) {   // splay: sum of squares of the x / z angles
        val = fmaf(a.w_splay * xi, xi, val);
        g = fmaf(2.f * a.w_splay, xi, g);
      }
    }
    if (has_prev || has_next) {   // temporal: w mean((x_f - x_f+1)^2) per pair and parameter group
      const float c = (i < 3 || i >= P) ? ct_g : ct_j;
      auto neighbour = [&](bool inside, int df, const float* halo) {
        if (inside) {
          if (i < 3) return a.global_rot[(long long)(f + df) * 3 + i] * m;
          if (i < P) return a.joint_rot[(long long)(f + df) * NJ3 + (i - 3)] * m;
          return a.trans[(long long)(f + df) * 3 + (i - P)];
        }
        return halo[i] * m;
      };
      if (has_next && (next_in || a.halo_next)) {
        const float d = xi - neighbour(next_in, 1, a.halo_next);
        val = fmaf(c * d, d, val);              // the pair (f, f+1) is counted by the owner of f
        g = fmaf(2.f * c, d, g);
      }
      if (has_prev && (prev_in || a.halo_prev)) {
        const float d = xi - neighbour(prev_in, -1, a.halo_prev);
        g = fmaf(2.f * c, d, g);
      }
    }
    g *= m * a.g_scale;
    float* out = (i < 3) ? a.g_global_rot + (long long)f * 3 + i
                         : (i < P ? a.g_joint_rot + (long long)f * NJ3 + (i - 3) : a.g_trans + (long long)f * 3 + (i - P));
    if (i < P || a.g_trans) {
      if (a.accumulate) *out += g; else *out = g;
    }
  }
  // frame value -> partials[f] (fixed-order tree)
  red[tid] = val;
  __syncthreads();
  for (int o = FIT_THREADS / 2; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  if (tid == 0) a.partials[f] = red[0];
}

// one block: betas prior (shared betas: one value per window, fitter.py:322-331) + the sum of the frame partials.
//   res = (betas - mean) prec;  per window w_betas mean(res^2)  ->  nwin * w_betas / nB * sum res^2
__global__ void __launch_bounds__(256) fitter_prior_finalize_kernel(SmilFitterPriorArgs a) {
  __shared__ double s_red[256];
  __shared__ float s_res[64];
  const int tid = threadIdx.x, nB = a.num_betas;
  // windows that START inside this shard
  long long nwin = 0;
  if (a.window <= 0) nwin = (a.frame_offset == 0) ? 1 : 0;
  else {
    const long long first = (a.frame_offset + a.window - 1) / a.window, last = (a.frame_offset + a.num_frames + a.window - 1) / a.window;
    nwin = last - first;
  }
  const bool bet = a.w_betas > 0.f && a.betas && a.betas_prec && nB > 0 && nB <= 64;
  double acc = 0.0;
  if (bet) {
    if (tid < nB) {
      float res = 0.f;
      for (int i = 0; i < nB; ++i) res = fmaf(a.betas[i] - (a.betas_mean ? a.betas_mean[i] : 0.f), a.betas_prec[i * nB + tid], res);
      s_res[tid] = res;
      acc = (double)nwin * (double)(a.w_betas / (float)nB) * (double)res * (double)res;
    }
    __syncthreads();
    if (tid < nB && a.g_betas) {
      float s = 0.f;
      for (int k = 0; k < nB; ++k) s = fmaf(a.betas_prec[tid * nB + k], s_res[k], s);
      const float g = (float)nwin * 2.f * a.w_betas / (float)nB * s * a.g_scale;
      if (a.accumulate) a.g_betas[tid] += g; else a.g_betas[tid] = g;
    }
  } else if (tid < nB && a.g_betas && !a.accumulate) {
    a.g_betas[tid] = 0.f;
  }
  for (int i = tid; i < a.num_frames; i += 256) acc += (double)a.partials[i];
  s_red[tid] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) s_red[tid] += s_red[tid + o];
    __syncthreads();
  }
  if (tid == 0) a.loss[0] = (float)s_red[0];
}

// ---- fused Adam (torch.optim.Adam defaults: no weight decay, no amsgrad) over up to SMIL_ADAM_MAX_TENSORS tensors ----
__global__ void adam_tick_kernel(int* step) { *step += 1; }
__global__ void __launch_bounds__(256) adam_step_kernel(SmilAdamArgs a, long long total) {
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= total) return;
  int t = 0;
  long long i = i0;
  while (t + 1 < a.num_tensors && i >= a.numel[t]) { i -= a.numel[t]; ++t; }
  const float step = (float)*a.step;
  const float bc1 = 1.f - powf(a.beta1, step), bc2 = 1.f - powf(a.beta2, step);
  const float g = a.grad[t][i];
  const float m = a.beta1 * a.exp_avg[t][i] + (1.f - a.beta1) * g;
  const float v = a.beta2 * a.exp_avg_sq[t][i] + (1.f - a.beta2) * g * g;
  a.exp_avg[t][i] = m;
  a.exp_avg_sq[t][i] = v;
  const float denom = sqrtf(v) / sqrtf(bc2) + a.eps;
  a.param[t][i] -= (a.lr[t] / bc1) * (m / denom);
}

}  // namespace smil
