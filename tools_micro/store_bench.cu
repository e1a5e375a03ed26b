// Microbenchmark: how fast can 256-vertex row tiles (3072 B) of [B][V][3] fp32 outputs leave an SM?
//   mode 0: smem ring -> cp.async.bulk (TMA) stores, R rows per stage, S stages
//   mode 1: smem double buffer -> coalesced STG.128 by the CTA's threads
//   mode 2: direct STG.32 x3 from registers (12-byte stride per thread)
// Persistent CTAs: grid = nchunk x Y, each CTA walks samples b = y, y+Y, ...
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int K> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(K) : "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int C = 256, SF = 772;

template <int MODE, int S, int R>
__global__ void __launch_bounds__(256) k(float* __restrict__ o0, float* __restrict__ o1, int B, int V, float seed) {
  extern __shared__ __align__(128) float sm[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chunk = blockIdx.x;
  const int nv = min(C, V - chunk * C), n = 3 * nv;
  float x = seed + tid, y = seed * 2 + tid, z = seed * 3 - tid;
  int sidx = 0;
  // a "pass" handles R/2 samples (two tensors each -> R rows)
  const int SPP = R / 2;
  for (int b0 = blockIdx.y * SPP; b0 < B; b0 += gridDim.y * SPP) {
    float* st = sm + sidx * (R * SF);
    if (MODE == 2) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int b = b0 + (r >> 1);
        if (b < B && tid < nv) {
          float* d = ((r & 1) ? o1 : o0) + ((long long)b * V + chunk * C + tid) * 3;
          d[0] = x + r; d[1] = y + r; d[2] = z + r;
        }
      }
      x += 1.f;
      continue;
    }
    if (tid < nv) {
#pragma unroll
      for (int r = 0; r < R; ++r) { float* p = st + r * SF + 3 * tid; p[0] = x + r; p[1] = y + r; p[2] = z + r; }
    }
    x += 1.f;
    if (MODE == 0) {
      fence_async();
      if (warp < R && lane == 0) bulk_wait_read<S - 2>();
      __syncthreads();
      if (warp < R && lane == 0) {
        const int b = b0 + (warp >> 1);
        if (b < B) bulk_s2g(((warp & 1) ? o1 : o0) + ((long long)b * V + chunk * C) * 3, st + warp * SF, n * 4);
        bulk_commit();
      }
    } else {
      __syncthreads();
      const int nq = n >> 2;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int b = b0 + (r >> 1);
        if (b < B && tid < nq) {
          float4* d = reinterpret_cast<float4*>(((r & 1) ? o1 : o0) + ((long long)b * V + chunk * C) * 3);
          d[tid] = reinterpret_cast<const float4*>(st + r * SF)[tid];
        }
      }
    }
    sidx = (sidx + 1 == S) ? 0 : sidx + 1;
  }
  if (MODE == 0 && warp < R && lane == 0) bulk_wait_read<0>();
}

template <int MODE, int S, int R>
void run(const char* name, float* o0, float* o1, int B, int V, int ctas_per_sm, int extra_smem) {
  const int nchunk = (V + C - 1) / C;
  int Y = (148 * ctas_per_sm) / nchunk;
  size_t smem = (size_t)S * R * SF * 4 + extra_smem;
  cudaFuncSetAttribute(k<MODE, S, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k<MODE, S, R>, 256, smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(e0);
    k<MODE, S, R><<<dim3(nchunk, Y), 256, smem>>>(o0, o1, B, V, 1.0f + rep);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (rep) best = ms < best ? ms : best;
  }
  double bytes = 2.0 * B * V * 12;
  printf("%-10s S=%d R=%d want %d CTA/SM (occ %d) Y=%3d smem %6zu: %.1f us  %.0f GB/s  %s\n", name, S, R, ctas_per_sm, occ, Y, smem,
         best * 1e3, bytes / best / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int B = 4096, V = 3020;
  float *o0, *o1;
  cudaMalloc(&o0, (size_t)B * V * 12); cudaMalloc(&o1, (size_t)B * V * 12);
  for (int c : {1, 2, 3, 4}) {
    run<0, 2, 4>("tma", o0, o1, B, V, c, 0);
    run<0, 3, 4>("tma", o0, o1, B, V, c, 0);
    run<0, 4, 4>("tma", o0, o1, B, V, c, 0);
    run<0, 3, 8>("tma", o0, o1, B, V, c, 0);
    run<0, 6, 2>("tma", o0, o1, B, V, c, 0);
    run<1, 2, 4>("stg128", o0, o1, B, V, c, 0);
    run<1, 2, 8>("stg128", o0, o1, B, V, c, 0);
    run<2, 1, 4>("direct", o0, o1, B, V, c, 0);
  }
  run<0, 3, 4>("tma+42K", o0, o1, B, V, 2, 42 * 1024);
  run<2, 1, 4>("direct", o0, o1, B, V, 8, 0);
  run<1, 2, 4>("stg128", o0, o1, B, V, 6, 0);
  cudaMemset(o0, 0, (size_t)B * V * 12);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) {
    cudaEventRecord(e0); cudaMemsetAsync(o0, 0, (size_t)B * V * 12); cudaMemsetAsync(o1, 0, (size_t)B * V * 12); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("memset x2: %.1f us %.0f GB/s\n", ms * 1e3, 2.0 * B * V * 12 / ms / 1e6);
  }
  return 0;
}
