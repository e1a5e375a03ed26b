// Microbenchmark: shared-memory wavefronts of warp gathers with few distinct addresses (the bone-transform
// gathers of skin_fwd).  Each launch = one (vector width, byte stride between bones, lane->bone pattern);
// read the wavefront counts with
//   ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum ./lds_gather
#include <cstdio>
#include <cuda_runtime.h>
template <int W>  // W = floats per load: 1, 2, 4
__global__ void k(const int* __restrict__ pat, int stride_floats, float* out, int iters) {
  extern __shared__ __align__(16) float sm[];
  for (int i = threadIdx.x; i < 12288; i += blockDim.x) sm[i] = (float)i;
  __syncthreads();
  const int bone = pat[threadIdx.x & 31];
  const float* p = sm + bone * stride_floats;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    const unsigned ad = (unsigned)__cvta_generic_to_shared(p) + 4u * W * (it & 3);
    if (W == 1) { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(ad)); acc += x; }
    if (W == 2) { float x, y; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(ad)); acc += x + y; }
    if (W == 4) { float x, y, z, w; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(ad)); acc += x + y + z + w; }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
  int h[8][32];
  const char* names[8] = {"uniform", "2 bones halves", "2 bones interleaved", "3 bones runs", "4 bones random", "8 bones random",
                          "16 bones random", "32 distinct"};
  unsigned s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return s >> 8; };
  for (int l = 0; l < 32; ++l) {
    h[0][l] = 5; h[1][l] = l < 16 ? 3 : 9; h[2][l] = (l & 1) ? 3 : 9; h[3][l] = l < 10 ? 2 : (l < 21 ? 11 : 30);
    static const int b4[4] = {1, 6, 12, 23}; h[4][l] = b4[rnd() & 3];
    static const int b8[8] = {0, 3, 5, 14, 17, 22, 36, 41}; h[5][l] = b8[rnd() & 7];
    h[6][l] = (rnd() % 16) * 3 + 1; h[7][l] = (l * 7) % 32 + 4;
  }
  int* d; cudaMalloc(&d, sizeof(h)); cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
  float* o; cudaMalloc(&o, 32 * 4);
  const int strides[] = {1, 2, 4, 24, 28, 32};
  for (int w : {1, 2, 4}) for (int st : strides) {
    if (st < w) continue;
    for (int p = 0; p < 8; ++p) {
      // one warp, 1000 loads: wavefronts/1000 = wavefronts per load
      if (w == 1) k<1><<<1, 32, 49152>>>(d + 32 * p, st, o, 1000);
      if (w == 2) k<2><<<1, 32, 49152>>>(d + 32 * p, st, o, 1000);
      if (w == 4) k<4><<<1, 32, 49152>>>(d + 32 * p, st, o, 1000);
      cudaDeviceSynchronize();
      printf("launch: W=%d stride=%d floats pattern=%s\n", w, st, names[p]);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
