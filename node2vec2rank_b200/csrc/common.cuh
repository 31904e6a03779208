// Shared host/device helpers for the n2v2r_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/n2v2r_b200.h"

namespace n2v2r {

// ---- error plumbing -------------------------------------------------------------------
void set_error(const char* fmt, ...);          // api.cu
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define N2V2R_CUDA(call)                                                            \
  do {                                                                              \
    cudaError_t _e = (call);                                                        \
    if (_e != cudaSuccess) return n2v2r::cuda_fail(_e, #call, __FILE__, __LINE__);   \
  } while (0)

#define N2V2R_CHECK_LAUNCH() N2V2R_CUDA(cudaGetLastError())

#define N2V2R_REQUIRE(cond, ...)                 \
  do {                                           \
    if (!(cond)) {                               \
      n2v2r::set_error(__VA_ARGS__);             \
      return N2V2R_EINVAL;                       \
    }                                            \
  } while (0)

int sm_count();                                // cached, current device

// ---- counter-based RNG (splitmix64 finaliser); NumPy twin in synthetic.py --------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// hash of (seed, stream id a, b, c): three chained mixes
__host__ __device__ __forceinline__ uint64_t hash4(uint64_t seed, uint64_t a, uint64_t b, uint64_t c) {
  uint64_t h = mix64(seed ^ (a * 0xD6E8FEB86659FD93ull));
  h = mix64(h ^ b);
  h = mix64(h ^ c);
  return h;
}
// uniform in [0,1) with 24 bits (exact in fp32)
__host__ __device__ __forceinline__ float u24(uint64_t h) { return (float)(uint32_t)(h >> 40) * (1.0f / 16777216.0f); }

// ---- bf16 splitting ---------------------------------------------------------------------
// x = hi + mid + lo with each term a bf16 (round-to-nearest); |x - (hi+mid+lo)| <= 2^-24 |x|.
__device__ __forceinline__ void split3(float x, __nv_bfloat16& hi, __nv_bfloat16& mid, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  float r = x - __bfloat162float(hi);
  mid = __float2bfloat16_rn(r);
  r -= __bfloat162float(mid);
  lo = __float2bfloat16_rn(r);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// bump allocator over the caller's workspace
struct Arena {
  char* base; size_t size; size_t off;
  Arena(void* p, size_t n) : base((char*)p), size(n), off(0) {}
  template <class T> T* take(size_t count) {
    size_t bytes = align_up(count * sizeof(T));
    T* r = (T*)(base ? base + off : nullptr);
    off += bytes;
    return r;
  }
  bool ok() const { return off <= size; }
};

}  // namespace n2v2r
