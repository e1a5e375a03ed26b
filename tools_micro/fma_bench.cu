// Microbenchmark: scalar FFMA vs packed fma.rn.f32x2 issue throughput on sm_100a.
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
  float2 acc[8];
  for (int i = 0; i < 8; ++i) acc[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float2 a = make_float2(s, s * 1.0001f), b = make_float2(0.999f, 1.001f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { acc[i].x = fmaf(acc[i].x, a.x, b.x); acc[i].y = fmaf(acc[i].y, a.y, b.y); }
      if (MODE == 1) { fma2(acc[i], a, b); }
      if (MODE == 2) { acc[i].x = fmaf(a.x, b.x, acc[i].x); acc[i].y = fmaf(a.y, b.y, acc[i].y); }
    }
  }
  float r = 0;
  for (int i = 0; i < 8; ++i) r += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 1024 * 4);
  const int iters = 20000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 3; ++mode) for (int bs : {128, 256, 512, 1024}) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<148 * 2, bs>>>(d, iters, 1.0f);
      if (mode == 1) k<1><<<148 * 2, bs>>>(d, iters, 1.0f);
      if (mode == 2) k<2><<<148 * 2, bs>>>(d, iters, 1.0f);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double fma = 148.0 * 2 * bs * (double)iters * 16;
      if (rep) printf("mode %d (%s) block %4d: %.3f ms  %.1f TFMA/s  -> %.1f FMA/clk/SM @1.965GHz\n", mode,
                      mode == 0 ? "FFMA dep-chain x16" : mode == 1 ? "FFMA2" : "FFMA acc+=a*b", bs, ms, fma / ms / 1e9,
                      fma / (ms * 1e-3) / 148 / 1.965e9);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
