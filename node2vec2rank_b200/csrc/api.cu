// Error plumbing, device queries and the K0 synthetic SBM generator.
#include "common.cuh"
#include <string.h>
#include <vector>

namespace n2v2r {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
  return N2V2R_ECUDA;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    cached[dev] = v;
  }
  return cached[dev];
}

// ---- K0: SBM generator ------------------------------------------------------------------
// One thread per 4 consecutive columns of a row; entries depend only on (seed, layer, min, max).
__global__ void __launch_bounds__(256) sbm_kernel(float* __restrict__ out, int64_t n, int64_t ld, int K,
                                                  const float* __restrict__ B, uint64_t seed, int layer,
                                                  int weighted) {
  extern __shared__ float sB[];
  for (int t = threadIdx.x; t < K * K; t += blockDim.x) sB[t] = B[t];
  __syncthreads();
  const int64_t quads = ld / 4;
  const int64_t total = n * quads;
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < total;
       q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = q / quads;
    const int64_t j0 = (q - i * quads) * 4;
    float v[4];
    const int zi = (int)(i % K);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int64_t j = j0 + e;
      float x = 0.f;
      if (j < n && j != i) {
        const uint64_t a = (uint64_t)(i < j ? i : j), b = (uint64_t)(i < j ? j : i);
        const uint64_t h = hash4(seed, (uint64_t)layer, a, b);
        const float p = sB[zi * K + (int)(j % K)];
        if (u24(h) < p) {
          if (weighted == 2) {
            const uint64_t h2 = mix64(h ^ 0xA5A5A5A5A5A5A5A5ull);
            x = u24(h2) + (1.0f / 16777216.0f);                            // 24-bit weight fp32(u') + 2^-24 in (0,1]
          } else if (weighted) {
            const uint64_t h2 = mix64(h ^ 0xA5A5A5A5A5A5A5A5ull);
            x = (float)((uint32_t)(h2 >> 48) + 1u) * (1.0f / 65536.0f);   // 16-bit weight in (0,1]
          } else {
            x = 1.f;
          }
        }
      }
      v[e] = x;
    }
    *reinterpret_cast<float4*>(out + i * ld + j0) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" int n2v2r_abi_version(void) { return N2V2R_ABI_VERSION; }
extern "C" const char* n2v2r_last_error(void) { return g_err; }

extern "C" int n2v2r_device_info(int device, int* sms, int* cc, size_t* total_mem) {
  cudaDeviceProp p;
  N2V2R_CUDA(cudaGetDeviceProperties(&p, device));
  if (sms) *sms = p.multiProcessorCount;
  if (cc) *cc = p.major * 10 + p.minor;
  if (total_mem) *total_mem = p.totalGlobalMem;
  return N2V2R_OK;
}

extern "C" int n2v2r_sbm_generate(void* stream, float* d_out, int64_t n, int64_t ld, int K,
                                  const float* h_B, uint64_t seed, int layer, int weighted) {
  N2V2R_REQUIRE(d_out && h_B && n > 0 && K > 0 && K <= 64, "sbm_generate: bad arguments");
  N2V2R_REQUIRE(ld >= n && ld % 4 == 0, "sbm_generate: ld must be >= n and a multiple of 4");
  cudaStream_t s = (cudaStream_t)stream;
  float* dB = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&dB, sizeof(float) * K * K, s));
  N2V2R_CUDA(cudaMemcpyAsync(dB, h_B, sizeof(float) * K * K, cudaMemcpyHostToDevice, s));
  const int grid = sm_count() * 8;
  sbm_kernel<<<grid, 256, sizeof(float) * K * K, s>>>(d_out, n, ld, K, dB, seed, layer, weighted);
  N2V2R_CHECK_LAUNCH();
  N2V2R_CUDA(cudaFreeAsync(dB, s));
  return N2V2R_OK;
}
