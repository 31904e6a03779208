// K1a / K1b: percentile order statistics, the fused load pass, plane utilities.
// Replaces preprocessing_utils.py:116-141 (network_transform) -- see include/n2v2r_b200.h.
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <algorithm>

namespace n2v2r {

// ---- element transform shared by select and apply (must be identical in both) ----------
template <class T>
__device__ __forceinline__ T pre_cut(T x, int absolute, int has_thr, T thr) {
  T y = x;
  if (absolute) y = (y < (T)0) ? -y : y;      // np.abs (also maps -0 -> +0 like fabs for ordering purposes)
  if (has_thr && y < thr) y = (T)0;           // network_copy[network_copy < threshold] = 0
  return y;
}

// order-preserving key of a float / double (ascending)
__device__ __forceinline__ uint64_t okey(float f) {
  uint32_t b = __float_as_uint(f);
  b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return (uint64_t)b;
}
__device__ __forceinline__ uint64_t okey(double f) {
  uint64_t b = (uint64_t)__double_as_longlong(f);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
template <class T> __device__ __forceinline__ T from_okey(uint64_t k);
template <> __device__ __forceinline__ float from_okey<float>(uint64_t k) {
  uint32_t b = (uint32_t)k;
  b = (b & 0x80000000u) ? (b & 0x7fffffffu) : ~b;
  return __uint_as_float(b);
}
template <> __device__ __forceinline__ double from_okey<double>(uint64_t k) {
  k = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)k);
}

constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kReplicas = 8;

// Histogram of the `width`-bit digit at bit offset `shift` over elements whose key matches
// `prefix` on the bits above shift+width.  Zeros (after the transform) are skipped.
template <class T>
__global__ void __launch_bounds__(512) select_hist_kernel(const T* __restrict__ in, int64_t count, int absolute,
                                                          int has_thr, T thr, uint64_t prefix, int shift, int width,
                                                          int first_pass, unsigned long long* __restrict__ ghist) {
  extern __shared__ unsigned int sh[];   // kReplicas * kBins
  for (int t = threadIdx.x; t < kReplicas * kBins; t += blockDim.x) sh[t] = 0;
  __syncthreads();
  unsigned int* my = sh + (threadIdx.x & (kReplicas - 1)) * kBins;
  const uint64_t dmask = (1ull << width) - 1ull;
  const int hi_shift = shift + width;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const T y = pre_cut<T>(in[i], absolute, has_thr, thr);
    if (y == (T)0) continue;               // x.nonzero(): NaN != 0 is kept, like NumPy
    const uint64_t k = okey(y);
    if (!first_pass && (hi_shift >= 64 ? 0ull : (k >> hi_shift)) != prefix) continue;
    atomicAdd(&my[(k >> shift) & dmask], 1u);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < kBins; b += blockDim.x) {
    unsigned long long s = 0;
#pragma unroll
    for (int r = 0; r < kReplicas; ++r) s += sh[r * kBins + b];
    if (s) atomicAdd(&ghist[b], s);
  }
}

// count(y <= v) and min(y > v) over the nonzeros (keys compared)
template <class T>
__global__ void __launch_bounds__(512) select_next_kernel(const T* __restrict__ in, int64_t count, int absolute,
                                                          int has_thr, T thr, uint64_t vkey,
                                                          unsigned long long* __restrict__ out /* [0]=cnt_le, [1]=min_gt key */) {
  unsigned long long cnt = 0, mn = ~0ull;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const T y = pre_cut<T>(in[i], absolute, has_thr, thr);
    if (y == (T)0) continue;
    const uint64_t k = okey(y);
    if (k <= vkey) ++cnt; else if (k < mn) mn = k;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    unsigned long long m2 = __shfl_xor_sync(0xffffffffu, mn, o);
    mn = m2 < mn ? m2 : mn;
  }
  if ((threadIdx.x & 31) == 0) {
    if (cnt) atomicAdd(&out[0], cnt);
    atomicMin(&out[1], mn);
  }
}

template <class T>
static int select_impl(cudaStream_t s, const T* in, int64_t count, int absolute, int has_thr, double thr_d,
                       double frac, int index_f32, void* ws, int64_t* h_nnz, double* h_lo, double* h_hi, double* h_pos) {
  // virtual index exactly as NumPy's 'linear' method computes it: (n-1)*quantile, in float32 when the
  // data (hence the quantile) is float32 (numpy/lib/_function_base_impl.py: percentile -> _quantile)
  auto virtual_index = [&](int64_t cnt) -> double {
    if (index_f32) { const float v = (float)(cnt - 1) * (float)frac; return (double)v; }
    return (double)(cnt - 1) * frac;
  };
  unsigned long long* ghist = (unsigned long long*)ws;          // kBins + 2
  static thread_local unsigned long long hh[kBins + 2];
  const int key_bits = sizeof(T) * 8;
  const T thr = (T)thr_d;
  const int grid = sm_count() * 4;
  const size_t smem = sizeof(unsigned int) * kReplicas * kBins;
  static bool attr_done = false;
  if (!attr_done) {
    N2V2R_CUDA(cudaFuncSetAttribute(select_hist_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    N2V2R_CUDA(cudaFuncSetAttribute(select_hist_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  uint64_t prefix = 0;
  int shift = key_bits;
  int64_t rank = 0, nnz = 0;
  bool first = true;
  while (shift > 0) {
    const int width = shift >= kDigitBits ? kDigitBits : shift;
    shift -= width;
    N2V2R_CUDA(cudaMemsetAsync(ghist, 0, sizeof(unsigned long long) * kBins, s));
    select_hist_kernel<T><<<grid, 512, smem, s>>>(in, count, absolute, has_thr, thr, prefix, shift, width,
                                                  first ? 1 : 0, ghist);
    N2V2R_CHECK_LAUNCH();
    N2V2R_CUDA(cudaMemcpyAsync(hh, ghist, sizeof(unsigned long long) * kBins, cudaMemcpyDeviceToHost, s));
    N2V2R_CUDA(cudaStreamSynchronize(s));
    if (first) {
      for (int b = 0; b < kBins; ++b) nnz += (int64_t)hh[b];
      *h_nnz = nnz;
      if (nnz == 0) { *h_lo = *h_hi = NAN; if (h_pos) *h_pos = 0.0; return N2V2R_OK; }
      double pos = virtual_index(nnz);
      if (h_pos) *h_pos = pos;
      rank = (int64_t)floor(pos);
      if (rank < 0) rank = 0;
      if (rank > nnz - 1) rank = nnz - 1;
      first = false;
    }
    int64_t acc = 0;
    int b = 0;
    for (; b < (1 << width); ++b) {
      if (acc + (int64_t)hh[b] > rank) break;
      acc += (int64_t)hh[b];
    }
    rank -= acc;
    prefix = (prefix << width) | (uint64_t)b;
  }
  const uint64_t vkey = prefix;
  // adjacent order statistic
  unsigned long long init[2] = {0ull, ~0ull};
  N2V2R_CUDA(cudaMemcpyAsync(ghist, init, sizeof(init), cudaMemcpyHostToDevice, s));
  select_next_kernel<T><<<grid, 512, 0, s>>>(in, count, absolute, has_thr, thr, vkey, ghist);
  N2V2R_CHECK_LAUNCH();
  N2V2R_CUDA(cudaMemcpyAsync(hh, ghist, sizeof(unsigned long long) * 2, cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  // host-side decode (same bit tricks as the device)
  auto decode = [&](uint64_t k) -> double {
    if (sizeof(T) == 4) {
      uint32_t b = (uint32_t)k;
      b = (b & 0x80000000u) ? (b & 0x7fffffffu) : ~b;
      float f; memcpy(&f, &b, 4); return (double)f;
    } else {
      k = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
      double d; memcpy(&d, &k, 8); return d;
    }
  };
  const double frac_pos = virtual_index(nnz);
  int64_t r0 = (int64_t)floor(frac_pos);
  if (r0 < 0) r0 = 0;
  if (r0 > nnz - 1) r0 = nnz - 1;
  *h_lo = decode(vkey);
  if ((int64_t)hh[0] >= r0 + 2 || hh[1] == ~0ull) *h_hi = *h_lo;
  else *h_hi = decode(hh[1]);
  return N2V2R_OK;
}

// ---- K1b: fused load pass -------------------------------------------------------------------
// Thread t of a CTA owns columns (2t, 2t+1) of a 512-column strip and walks down a range of
// rows: loads are coalesced along the row, column sums stay in registers.
template <class T>
__global__ void __launch_bounds__(256) transform_kernel(const T* __restrict__ in, int64_t rows, int64_t cols, int64_t ld_in,
                                                        int absolute, int has_thr, T thr, int has_cut, T cut, int binarize,
                                                        float* __restrict__ out, int64_t ld_out,
                                                        __nv_bfloat16* __restrict__ planes, int n_planes, int64_t ld_p,
                                                        double* __restrict__ colsum, int rows_per_cta) {
  const int64_t j0 = (int64_t)blockIdx.x * 512 + 2 * threadIdx.x;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r_end = (rows < r_begin + rows_per_cta) ? rows : (r_begin + rows_per_cta);
  const int64_t width = planes ? ld_p : cols;
  if (j0 >= width) return;
  const bool in0 = j0 < cols, in1 = j0 + 1 < cols;
  double s0 = 0.0, s1 = 0.0;
  const size_t pstride = (size_t)rows * ld_p;
#pragma unroll 4
  for (int64_t i = r_begin; i < r_end; ++i) {
    float v[2] = {0.f, 0.f};
    const bool inb[2] = {in0, in1};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      if (inb[e]) {
        T y = pre_cut<T>(in[i * ld_in + j0 + e], absolute, has_thr, thr);
        if (has_cut && y < cut) y = (T)0;
        if (binarize && y != (T)0) y = (T)1;
        v[e] = (float)y;
      }
    }
    s0 += (double)v[0];
    s1 += (double)v[1];
    if (out) {
      if (in0) out[i * ld_out + j0] = v[0];
      if (in1) out[i * ld_out + j0 + 1] = v[1];
    }
    if (planes) {   // j0 even and ld_p even -> 4-byte aligned pair store; j0+1 < ld_p always
      float r0 = v[0], r1 = v[1];
      for (int p = 0; p < n_planes; ++p) {
        __nv_bfloat162 h;
        h.x = __float2bfloat16_rn(r0);
        h.y = __float2bfloat16_rn(r1);
        r0 -= __bfloat162float(h.x);
        r1 -= __bfloat162float(h.y);
        *reinterpret_cast<__nv_bfloat162*>(planes + p * pstride + i * ld_p + j0) = h;
      }
    }
  }
  if (colsum) {
    if (in0) atomicAdd(&colsum[j0], s0);
    if (in1) atomicAdd(&colsum[j0 + 1], s1);
  }
}

__global__ void __launch_bounds__(256) planes_needed_kernel(const float* __restrict__ in, int64_t rows, int64_t cols,
                                                            int64_t ld, int* __restrict__ need) {
  int local = 1;
  for (int64_t i = blockIdx.y; i < rows; i += gridDim.y) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < cols; j += (int64_t)gridDim.x * blockDim.x) {
      const float x = in[i * ld + j];
      float r = x - __bfloat162float(__float2bfloat16_rn(x));
      if (r != 0.f) {
        r -= __bfloat162float(__float2bfloat16_rn(r));
        local = max(local, r != 0.f ? 3 : 2);
      }
    }
  }
  local = __reduce_max_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local > 1) atomicMax(need, local);
}

// tile pair (bi <= bj): compare tile(bi,bj) with tile(bj,bi)^T
__global__ void __launch_bounds__(256) symmetric_kernel(const uint16_t* __restrict__ planes, int n_planes, int64_t n,
                                                        int64_t ld, int* __restrict__ asym) {
  __shared__ uint16_t a[32][33], b[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bi > bj) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  int bad = 0;
  for (int p = 0; p < n_planes; ++p) {
    const uint16_t* P = planes + (size_t)p * n * ld;
    for (int r = ty; r < 32; r += 8) {
      const int64_t i = (int64_t)bi * 32 + r, j = (int64_t)bj * 32 + tx;
      a[r][tx] = (i < n && j < n) ? P[i * ld + j] : 0;
      const int64_t i2 = (int64_t)bj * 32 + r, j2 = (int64_t)bi * 32 + tx;
      b[r][tx] = (i2 < n && j2 < n) ? P[i2 * ld + j2] : 0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) bad |= (a[r][tx] != b[tx][r]);
    __syncthreads();
  }
  if (__any_sync(0xffffffffu, bad) && tx == 0) atomicOr(asym, 1);
}

__global__ void __launch_bounds__(256) transpose_kernel(const uint16_t* __restrict__ in, uint16_t* __restrict__ out,
                                                        int64_t n, int64_t ld) {
  __shared__ uint16_t t[32][33];
  const size_t poff = (size_t)blockIdx.z * n * ld;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = (int64_t)blockIdx.y * 32 + r, j = (int64_t)blockIdx.x * 32 + tx;
    t[r][tx] = (i < n && j < n) ? in[poff + i * ld + j] : 0;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = (int64_t)blockIdx.x * 32 + r, j = (int64_t)blockIdx.y * 32 + tx;   // out[i][j] = in[j][i]
    if (i < n && j < ld) out[poff + i * ld + j] = (j < n) ? t[tx][r] : 0;
  }
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" size_t n2v2r_select_workspace_bytes(void) { return sizeof(unsigned long long) * (kBins + 2); }

extern "C" int n2v2r_select_nonzero(void* stream, const void* d_in, int dtype, int64_t rows, int64_t cols,
                                    int64_t ld, int absolute, int has_thr, double thr, double frac, int index_f32,
                                    void* d_ws, size_t ws_bytes, int64_t* h_nnz, double* h_lo, double* h_hi, double* h_pos) {
  N2V2R_REQUIRE(d_in && d_ws && h_nnz && h_lo && h_hi, "select_nonzero: null pointer");
  N2V2R_REQUIRE(ld == cols, "select_nonzero: input must be contiguous (ld == cols)");
  N2V2R_REQUIRE(frac >= 0.0 && frac <= 1.0, "select_nonzero: frac outside [0,1]");
  if (ws_bytes < n2v2r_select_workspace_bytes()) { set_error("select_nonzero: workspace too small"); return N2V2R_ENOMEM; }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t count = rows * cols;
  if (dtype == N2V2R_F32) return select_impl<float>(s, (const float*)d_in, count, absolute, has_thr, thr, frac, index_f32, d_ws, h_nnz, h_lo, h_hi, h_pos);
  if (dtype == N2V2R_F64) return select_impl<double>(s, (const double*)d_in, count, absolute, has_thr, thr, frac, index_f32, d_ws, h_nnz, h_lo, h_hi, h_pos);
  set_error("select_nonzero: unknown dtype %d", dtype);
  return N2V2R_EINVAL;
}

extern "C" int n2v2r_transform(void* stream, const void* d_in, int dtype, int64_t rows, int64_t cols, int64_t ld_in,
                               int absolute, int has_thr, double thr, int has_cut, double cut, int binarize,
                               float* d_out_f32, int64_t ld_out, void* d_planes, int n_planes, int64_t ld_p,
                               double* d_colsum) {
  N2V2R_REQUIRE(d_in && rows > 0 && cols > 0 && ld_in >= cols, "transform: bad input");
  N2V2R_REQUIRE(!d_out_f32 || ld_out >= cols, "transform: ld_out < cols");
  N2V2R_REQUIRE(!d_planes || (n_planes >= 1 && n_planes <= 3 && ld_p >= cols && ld_p % 64 == 0),
                "transform: planes need 1..3 planes and ld_p a multiple of 64 >= cols");
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t width = d_planes ? ld_p : cols;
  const int strips = (int)((width + 511) / 512);
  int row_ctas = (sm_count() * 8 + strips - 1) / strips;
  if (row_ctas > rows) row_ctas = (int)rows;
  if (row_ctas > 65535) row_ctas = 65535;
  const int rows_per_cta = (int)((rows + row_ctas - 1) / row_ctas);
  dim3 grid(strips, (unsigned)((rows + rows_per_cta - 1) / rows_per_cta));
  if (dtype == N2V2R_F32)
    transform_kernel<float><<<grid, 256, 0, s>>>((const float*)d_in, rows, cols, ld_in, absolute, has_thr, (float)thr, has_cut,
                                                 (float)cut, binarize, d_out_f32, ld_out, (__nv_bfloat16*)d_planes, n_planes,
                                                 ld_p, d_colsum, rows_per_cta);
  else if (dtype == N2V2R_F64)
    transform_kernel<double><<<grid, 256, 0, s>>>((const double*)d_in, rows, cols, ld_in, absolute, has_thr, thr, has_cut, cut,
                                                  binarize, d_out_f32, ld_out, (__nv_bfloat16*)d_planes, n_planes, ld_p,
                                                  d_colsum, rows_per_cta);
  else { set_error("transform: unknown dtype %d", dtype); return N2V2R_EINVAL; }
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_planes_needed(void* stream, const float* d_in, int64_t rows, int64_t cols, int64_t ld, int* h_planes) {
  N2V2R_REQUIRE(d_in && h_planes && rows > 0 && cols > 0 && ld >= cols, "planes_needed: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  int* d_need = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_need, sizeof(int), s));
  int one = 1;
  N2V2R_CUDA(cudaMemcpyAsync(d_need, &one, sizeof(int), cudaMemcpyHostToDevice, s));
  dim3 grid((unsigned)std::min<int64_t>((cols + 255) / 256, 64), (unsigned)std::min<int64_t>(rows, (int64_t)sm_count() * 2));
  planes_needed_kernel<<<grid, 256, 0, s>>>(d_in, rows, cols, ld, d_need);
  N2V2R_CHECK_LAUNCH();
  N2V2R_CUDA(cudaMemcpyAsync(h_planes, d_need, sizeof(int), cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  N2V2R_CUDA(cudaFreeAsync(d_need, s));
  return N2V2R_OK;
}

extern "C" int n2v2r_is_symmetric(void* stream, const void* d_planes, int n_planes, int64_t n, int64_t ld_p, int* h_sym) {
  N2V2R_REQUIRE(d_planes && h_sym && n > 0 && n_planes >= 1 && n_planes <= 3 && ld_p >= n, "is_symmetric: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  int* d_flag = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_flag, sizeof(int), s));
  N2V2R_CUDA(cudaMemsetAsync(d_flag, 0, sizeof(int), s));
  const unsigned tiles = (unsigned)((n + 31) / 32);
  symmetric_kernel<<<dim3(tiles, tiles), 256, 0, s>>>((const uint16_t*)d_planes, n_planes, n, ld_p, d_flag);
  N2V2R_CHECK_LAUNCH();
  int flag = 0;
  N2V2R_CUDA(cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  N2V2R_CUDA(cudaFreeAsync(d_flag, s));
  *h_sym = flag ? 0 : 1;
  return N2V2R_OK;
}

extern "C" int n2v2r_transpose_planes(void* stream, const void* d_planes, int n_planes, int64_t n, int64_t ld_p, void* d_out) {
  N2V2R_REQUIRE(d_planes && d_out && n > 0 && n_planes >= 1 && n_planes <= 3 && ld_p >= n, "transpose_planes: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned tr = (unsigned)((n + 31) / 32), tc = (unsigned)((ld_p + 31) / 32);
  // blockIdx.x walks the input columns (output rows < n), blockIdx.y the input rows (output cols < ld)
  transpose_kernel<<<dim3(tr, tc, n_planes), 256, 0, s>>>((const uint16_t*)d_planes, (uint16_t*)d_out, n, ld_p);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
