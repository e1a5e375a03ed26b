// Microbenchmark (sm_100a): FMA-pipe rates that decide scalar-vs-packed in the lane-per-sample passes.
//   mode 0: scalar FFMA, 16 independent accumulators, three distinct register operands per instruction
//   mode 1: packed FFMA2 (fma.rn.f32x2), 8 independent accumulator pairs (the same number of FMAs per iteration)
//   mode 2: mode 0 + one warp-uniform LDS.128 per 8 FFMA whose result feeds the next iteration's operands
//   mode 3: mode 1 + one warp-uniform LDS.128 per 4 FFMA2
//   mode 4: scalar FFMA + 3 conflict-free LDS.32 per 16 FFMA (lane-strided reads of a transposed staging tile)
// Reports FMA / clk / SM at the nominal 1.965 GHz for several CTA sizes (warps per SM).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void fma2(float2& d, float2 a, float2 b) {
  unsigned long long dd = *reinterpret_cast<unsigned long long*>(&d), aa = *reinterpret_cast<unsigned long long*>(&a),
                     bb = *reinterpret_cast<unsigned long long*>(&b);
  asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(dd) : "l"(aa), "l"(bb));
  d = *reinterpret_cast<float2*>(&dd);
}

template <int MODE>
__global__ void k(float* out, int iters, float s) {
  __shared__ float4 tab[64];
  __shared__ float tile[33 * 32];
  if (threadIdx.x < 64) tab[threadIdx.x] = make_float4(s, s * 1.001f, s * 0.999f, s * 1.002f);
  for (int i = threadIdx.x; i < 33 * 32; i += blockDim.x) tile[i] = s + i * 1e-6f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  float acc[16];
  float2 acc2[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 0.001f + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc2[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float x[4] = {s, s * 1.01f, s * 0.99f, s * 1.02f}, y[4] = {0.999f, 1.001f, 0.998f, 1.002f};
  float2 x2[4], y2[2];
#pragma unroll
  for (int i = 0; i < 4; ++i) x2[i] = make_float2(x[i], x[(i + 1) & 3]);
  y2[0] = make_float2(y[0], y[1]); y2[1] = make_float2(y[2], y[3]);
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 2 || MODE == 4) {
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fmaf(x[i & 3], y[i >> 2], acc[i]);
      if (MODE == 2) {
        const float4 t0 = tab[it & 63], t1 = tab[(it + 7) & 63];
        x[0] = t0.x; x[1] = t0.y; x[2] = t0.z; x[3] = t0.w; y[0] = t1.x; y[1] = t1.y; y[2] = t1.z; y[3] = t1.w;
      }
      if (MODE == 4) {
        const int c = (it * 3) & 31;
        x[0] = tile[c * 33 + lane]; x[1] = tile[((c + 1) & 31) * 33 + lane]; x[2] = tile[((c + 2) & 31) * 33 + lane];
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) fma2(acc2[i], x2[i & 3], y2[i >> 2]);
      if (MODE == 3) {
        const float4 t0 = tab[it & 63], t1 = tab[(it + 7) & 63];
        x2[0] = make_float2(t0.x, t0.y); x2[1] = make_float2(t0.z, t0.w);
        x2[2] = make_float2(t1.x, t1.y); x2[3] = make_float2(t1.z, t1.w);
      }
    }
  }
  float r = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += acc[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) r += acc2[i].x + acc2[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

int main() {
  float* d;
  cudaMalloc(&d, 148 * 4 * 1024 * 4);
  const int iters = 20000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const char* names[] = {"FFMA scalar 16 acc", "FFMA2 8 acc pairs", "FFMA + 2 LDS.128 uniform / 16", "FFMA2 + 2 LDS.128 uniform / 8",
                         "FFMA + 3 LDS.32 lane-strided / 16"};
  for (int mode = 0; mode < 5; ++mode)
    for (int bs : {128, 256, 512, 1024}) {
      float ms = 0;
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (mode == 0) k<0><<<148, bs>>>(d, iters, 1.0f);
        if (mode == 1) k<1><<<148, bs>>>(d, iters, 1.0f);
        if (mode == 2) k<2><<<148, bs>>>(d, iters, 1.0f);
        if (mode == 3) k<3><<<148, bs>>>(d, iters, 1.0f);
        if (mode == 4) k<4><<<148, bs>>>(d, iters, 1.0f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
      }
      const double fma = 148.0 * bs * (double)iters * 16;
      printf("mode %d (%-34s) warps/SM %2d: %.3f ms  %.1f FMA/clk/SM @1.965GHz\n", mode, names[mode], bs / 32, ms,
             fma / (ms * 1e-3) / 148 / 1.965e9);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
