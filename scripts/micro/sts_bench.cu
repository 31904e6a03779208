// Micro-benchmark: cost of scattered shared-memory stores of 2 and 4 bytes (4 warps, like the converter warps of
// sparse_mma_kernel).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sts_bench sts_bench.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int BYTES, int MODE>   // MODE 0: bank-distinct positions per warp instruction, 1: random positions
__global__ void k(const uint32_t* pos, unsigned long long* out, int iters) {
  extern __shared__ uint8_t sm[];
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm);
  const int t = threadIdx.x;
  uint32_t p[8];
  for (int e = 0; e < 8; ++e) p[e] = pos[(MODE * 8 + e) * blockDim.x + t];
  __syncthreads();
  const long long c0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      if (BYTES == 2) asm volatile("st.shared.u16 [%0], %1;" ::"r"(base + 2 * p[e]), "h"((unsigned short)it) : "memory");
      else asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * (p[e] >> 1)), "r"((uint32_t)it) : "memory");
    }
  }
  __syncthreads();
  const long long c1 = clock64();
  if (t == 0) out[0] = (unsigned long long)(c1 - c0);
}

int main() {
  const int T = 128, iters = 1000;
  uint32_t h[16 * T];
  // mode 0: store e of lane l of warp w hits element (bank l) in row group ...: 32 distinct banks per warp instruction
  for (int e = 0; e < 8; ++e)
    for (int t = 0; t < T; ++t) {
      const int lane = t & 31, w = t >> 5;
      h[e * T + t] = (uint32_t)(((e * 4 + w) * 997 % 500) * 64 + lane * 2 + ((e + w) & 1));        // word = row*32 + lane -> bank = lane
      h[(8 + e) * T + t] = (uint32_t)((1103515245u * (uint32_t)(e * T + t + 1) + 12345u) >> 8) % 32000u;   // random element in 64 KB
    }
  uint32_t* d; unsigned long long* o;
  cudaMalloc(&d, sizeof(h)); cudaMalloc(&o, 8);
  cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(k<2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<4, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(k<4, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  unsigned long long r;
  const char* names[4] = {"u16 bank-distinct", "u16 random", "u32 bank-distinct", "u32 random"};
  for (int v = 0; v < 4; ++v) {
    for (int rep = 0; rep < 2; ++rep) {
      if (v == 0) k<2, 0><<<1, T, 65536>>>(d, o, iters);
      if (v == 1) k<2, 1><<<1, T, 65536>>>(d, o, iters);
      if (v == 2) k<4, 0><<<1, T, 65536>>>(d, o, iters);
      if (v == 3) k<4, 1><<<1, T, 65536>>>(d, o, iters);
      cudaDeviceSynchronize();
    }
    cudaMemcpy(&r, o, 8, cudaMemcpyDeviceToHost);
    printf("%-20s %.2f clk per warp store instruction (4 warps, 8 stores per thread per iteration)\n", names[v], (double)r / iters / 32.0);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
