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

// ---- flat streaming read: 16-byte loads, four of them in flight per thread ----------------------
// Calls f(value) for every element of in[0..count).  The HBM-bound passes of this file (histogram, adjacent
// order statistic, plane count) are pure reads: what bounds them is bytes in flight per SM, so every thread
// issues 4 x LDG.128 before it touches the first value.  `in` must be 16-byte aligned (torch allocations are).
template <class T> struct Vec16;
template <> struct Vec16<float> { using type = float4; static constexpr int n = 4; };
template <> struct Vec16<double> { using type = double2; static constexpr int n = 2; };
template <class F> __device__ __forceinline__ void each(const float4& v, F& f) { f(v.x); f(v.y); f(v.z); f(v.w); }
template <class F> __device__ __forceinline__ void each(const double2& v, F& f) { f(v.x); f(v.y); }

template <class T, class F>
__device__ __forceinline__ void stream_read(const T* __restrict__ in, int64_t count, F f) {
  using V = typename Vec16<T>::type;
  constexpr int per = Vec16<T>::n;
  const V* __restrict__ v = reinterpret_cast<const V*>(in);
  const int64_t nv = count / per;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < nv; i += 4 * stride) {
    const V a = __ldg(v + i), b = __ldg(v + i + stride), c = __ldg(v + i + 2 * stride), d = __ldg(v + i + 3 * stride);
    each(a, f); each(b, f); each(c, f); each(d, f);
  }
  for (; i < nv; i += stride) { const V a = __ldg(v + i); each(a, f); }
  if (blockIdx.x == 0 && threadIdx.x < count - nv * per) f(in[nv * per + threadIdx.x]);   // tail (< 4 elements)
}

constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kReplicas = 4;      // 32 KB of shared histograms per CTA: four 512-thread CTAs per SM

// Native-width order-preserving keys: the passes are instruction-issue bound (a 6.5 TB/s fp32 stream leaves ~20 thread
// instructions per element), so fp32 data is keyed in 32 bits and the element transform is resolved at compile time.
template <class T> struct Key;
template <> struct Key<float> {
  using type = uint32_t;
  static __device__ __forceinline__ uint32_t of(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
};
template <> struct Key<double> {
  using type = uint64_t;
  static __device__ __forceinline__ uint64_t of(double f) { return okey(f); }
};

// Histogram of the `width`-bit digit at bit offset `shift` over elements whose key matches
// `prefix` on the bits above shift+width.  Zeros (after the transform) are skipped.
// kPre: abs / threshold present; kFirst: first pass (no prefix filter).
template <class T, bool kPre, bool kFirst>
__global__ void __launch_bounds__(512) select_hist_kernel(const T* __restrict__ in, int64_t count, int absolute,
                                                          int has_thr, T thr, uint64_t prefix64, int shift, int width,
                                                          unsigned long long* __restrict__ ghist) {
  using K = typename Key<T>::type;
  extern __shared__ unsigned int sh[];   // kReplicas * kBins
  for (int t = threadIdx.x; t < kReplicas * kBins; t += blockDim.x) sh[t] = 0;
  __syncthreads();
  unsigned int* my = sh + (threadIdx.x & (kReplicas - 1)) * kBins;
  const K dmask = (K)((1ull << width) - 1ull);
  const int hi_shift = shift + width;                  // < key bits on every pass but the first
  const K prefix = (K)prefix64;
  stream_read<T>(in, count, [&](T x) {
    const T y = kPre ? pre_cut<T>(x, absolute, has_thr, thr) : x;
    if (y == (T)0) return;                 // x.nonzero(): NaN != 0 is kept, like NumPy
    const K k = Key<T>::of(y);
    if (!kFirst && (K)(k >> hi_shift) != prefix) return;
    atomicAdd(&my[(k >> shift) & dmask], 1u);
  });
  __syncthreads();
  for (int b = threadIdx.x; b < kBins; b += blockDim.x) {
    unsigned long long s = 0;
#pragma unroll
    for (int r = 0; r < kReplicas; ++r) s += sh[r * kBins + b];
    if (s) atomicAdd(&ghist[b], s);
  }
}

// count(y <= v) and min(y > v) over the nonzeros (keys compared)
template <class T, bool kPre>
__global__ void __launch_bounds__(512) select_next_kernel(const T* __restrict__ in, int64_t count, int absolute,
                                                          int has_thr, T thr, uint64_t vkey64,
                                                          unsigned long long* __restrict__ out /* [0]=cnt_le, [1]=min_gt key */) {
  using K = typename Key<T>::type;
  const K vkey = (K)vkey64;
  unsigned int cnt32 = 0;                  // per thread: at most count / (grid * block) + 4 elements
  K mnk = (K)~(K)0;
  stream_read<T>(in, count, [&](T x) {
    const T y = kPre ? pre_cut<T>(x, absolute, has_thr, thr) : x;
    if (y == (T)0) return;
    const K k = Key<T>::of(y);
    if (k <= vkey) ++cnt32; else if (k < mnk) mnk = k;
  });
  unsigned long long cnt = cnt32;
  unsigned long long mn = (mnk == (K)~(K)0) ? ~0ull : (unsigned long long)mnk;
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
    attr_done = true;      // 32 KB of dynamic shared memory: below the 48 KB default limit, no attribute needed
  }
  uint64_t prefix = 0;
  int shift = key_bits;
  int64_t rank = 0, nnz = 0;
  bool first = true;
  while (shift > 0) {
    const int width = shift >= kDigitBits ? kDigitBits : shift;
    shift -= width;
    N2V2R_CUDA(cudaMemsetAsync(ghist, 0, sizeof(unsigned long long) * kBins, s));
    const bool pre = absolute || has_thr;
    if (first && pre) select_hist_kernel<T, true, true><<<grid, 512, smem, s>>>(in, count, absolute, has_thr, thr, prefix, shift, width, ghist);
    else if (first) select_hist_kernel<T, false, true><<<grid, 512, smem, s>>>(in, count, absolute, has_thr, thr, prefix, shift, width, ghist);
    else if (pre) select_hist_kernel<T, true, false><<<grid, 512, smem, s>>>(in, count, absolute, has_thr, thr, prefix, shift, width, ghist);
    else select_hist_kernel<T, false, false><<<grid, 512, smem, s>>>(in, count, absolute, has_thr, thr, prefix, shift, width, ghist);
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
  if (absolute || has_thr) select_next_kernel<T, true><<<grid, 512, 0, s>>>(in, count, absolute, has_thr, thr, vkey, ghist);
  else select_next_kernel<T, false><<<grid, 512, 0, s>>>(in, count, absolute, has_thr, thr, vkey, ghist);
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

// Fast path of K1b: every thread owns FOUR columns of a 1024-column strip (one 16-byte load per row for fp32, two for
// fp64) and walks its row range four rows at a time, all loads of the batch issued before the first store, so a CTA keeps
// 16 KB (fp32) of reads in flight.  Needs cols % 4 == 0, 16-byte aligned bases and leading dimensions that are multiples of
// 4 elements; the two-column kernel above is the general path.
template <class T> __device__ __forceinline__ void load4(const T* __restrict__ p, T (&v)[4]);
template <> __device__ __forceinline__ void load4<float>(const float* __restrict__ p, float (&v)[4]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <> __device__ __forceinline__ void load4<double>(const double* __restrict__ p, double (&v)[4]) {
  const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// Rows are staged through a per-thread ring in shared memory with cp.async (LDGSTS): kRing rows (16 B each for fp32, 32 B
// for fp64) are in flight per thread without holding registers, i.e. 32 / 64 KB per CTA and ~190 KB per SM -- the
// bytes-in-flight a 6.5 TB/s stream needs (Little's law at ~2 us loaded latency).  Every thread reads back only the slots
// it filled itself, so cp.async.wait_group is the only synchronisation.
constexpr int kRing = 8;

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// The element options are template flags (kPre: abs / threshold, kCut, kBin, kSum: column sums): like the select passes
// this kernel is bound by instruction issue once the loads are deep enough, so nothing is decided per element at run time.
template <class T, bool kPre, bool kCut, bool kBin, bool kSum>
__global__ void __launch_bounds__(256) transform4_kernel(const T* __restrict__ in, int64_t rows, int64_t cols, int64_t ld_in,
                                                         int absolute, int has_thr, T thr, T cut,
                                                         float* __restrict__ out, int64_t ld_out,
                                                         __nv_bfloat16* __restrict__ planes, int n_planes, int64_t ld_p,
                                                         double* __restrict__ colsum, int rows_per_cta) {
  extern __shared__ __align__(16) unsigned char ring_raw[];
  constexpr int kSlot = 4 * sizeof(T);                                    // bytes per thread per row
  constexpr int kStage = 256 * kSlot;                                     // bytes per ring stage
  unsigned char* my = ring_raw + (size_t)threadIdx.x * kSlot;             // slot s of this thread: my + s * kStage
  const int64_t j0 = (int64_t)blockIdx.x * 1024 + 4 * threadIdx.x;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r_end = (rows < r_begin + rows_per_cta) ? rows : (r_begin + rows_per_cta);
  const int64_t width = planes ? ld_p : cols;
  if (j0 >= width || r_begin >= r_end) return;
  const bool inside = j0 < cols;                       // cols % 4 == 0: a thread's four columns are all in or all padding
  const int n_rows = (int)(r_end - r_begin);
  double cs[4] = {0.0, 0.0, 0.0, 0.0};
  const size_t pstride = (size_t)rows * ld_p;
  const unsigned char* src = reinterpret_cast<const unsigned char*>(in + r_begin * ld_in + j0);
  const size_t src_step = (size_t)ld_in * sizeof(T);
  int issued = 0;
  auto issue = [&]() {
    if (inside && issued < n_rows) {
      unsigned char* dst = my + (issued & (kRing - 1)) * kStage;
      cp_async16(dst, src);
      if (sizeof(T) == 8) cp_async16(dst + 16, src + 16);
      src += src_step;
    }
    ++issued;
    cp_async_commit();                                 // one group per row, empty past the end: the wait count stays fixed
  };
#pragma unroll
  for (int sidx = 0; sidx < kRing; ++sidx) issue();
  float* po = out ? out + r_begin * ld_out + j0 : nullptr;
  __nv_bfloat16* pp = planes ? planes + r_begin * ld_p + j0 : nullptr;
  for (int r = 0; r < n_rows; ++r) {
    cp_async_wait<kRing - 1>();                        // row r has landed
    T raw[4];
    if (inside) {
      const T* slot = reinterpret_cast<const T*>(my + (r & (kRing - 1)) * kStage);
      raw[0] = slot[0]; raw[1] = slot[1]; raw[2] = slot[2]; raw[3] = slot[3];
    } else {
      raw[0] = raw[1] = raw[2] = raw[3] = (T)0;
    }
    issue();                                           // refill the slot just read
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      T y = kPre ? pre_cut<T>(raw[e], absolute, has_thr, thr) : raw[e];
      if (kCut && y < cut) y = (T)0;
      if (kBin && y != (T)0) y = (T)1;
      v[e] = (float)y;
      if (kSum) cs[e] += (double)v[e];
    }
    if (po) {
      if (inside) *reinterpret_cast<float4*>(po) = make_float4(v[0], v[1], v[2], v[3]);
      po += ld_out;
    }
    if (pp) {
      __nv_bfloat16* q = pp;
      float res[4] = {v[0], v[1], v[2], v[3]};
      for (int p = 0; p < n_planes; ++p) {
        __nv_bfloat16 h[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) { h[e] = __float2bfloat16_rn(res[e]); res[e] -= __bfloat162float(h[e]); }
        uint2 pk;
        pk.x = (uint32_t)__bfloat16_as_ushort(h[0]) | ((uint32_t)__bfloat16_as_ushort(h[1]) << 16);
        pk.y = (uint32_t)__bfloat16_as_ushort(h[2]) | ((uint32_t)__bfloat16_as_ushort(h[3]) << 16);
        *reinterpret_cast<uint2*>(q) = pk;
        q += pstride;
      }
      pp += ld_p;
    }
  }
  cp_async_wait<0>();
  if (kSum && colsum && inside) {
#pragma unroll
    for (int e = 0; e < 4; ++e) atomicAdd(&colsum[j0 + e], cs[e]);
  }
}

template <class T>
static cudaError_t launch_transform4(cudaStream_t s, dim3 grid, const T* in, int64_t rows, int64_t cols, int64_t ld_in, int absolute,
                                     int has_thr, T thr, int has_cut, T cut, int binarize, float* out, int64_t ld_out,
                                     __nv_bfloat16* planes, int n_planes, int64_t ld_p, double* colsum, int rpc) {
  const size_t smem = (size_t)kRing * 256 * 4 * sizeof(T);
  const int mode = ((absolute || has_thr) ? 1 : 0) | (has_cut ? 2 : 0) | (binarize ? 4 : 0) | (colsum ? 8 : 0);
#define N2V2R_T4(M)                                                                                                              \
  case M: {                                                                                                                      \
    auto kern = transform4_kernel<T, (M & 1) != 0, (M & 2) != 0, (M & 4) != 0, (M & 8) != 0>;                                    \
    static bool attr_set = false;            /* once per instantiation: the call is slow for > 48 KB */                          \
    if (!attr_set) {                                                                                                             \
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                        \
      if (e != cudaSuccess) return e;                                                                                            \
      attr_set = true;                                                                                                           \
    }                                                                                                                            \
    kern<<<grid, 256, smem, s>>>(in, rows, cols, ld_in, absolute, has_thr, thr, cut, out, ld_out, planes, n_planes, ld_p, colsum, rpc); \
    break;                                                                                                                       \
  }
  switch (mode) {
    N2V2R_T4(0) N2V2R_T4(1) N2V2R_T4(2) N2V2R_T4(3) N2V2R_T4(4) N2V2R_T4(5) N2V2R_T4(6) N2V2R_T4(7)
    N2V2R_T4(8) N2V2R_T4(9) N2V2R_T4(10) N2V2R_T4(11) N2V2R_T4(12) N2V2R_T4(13) N2V2R_T4(14) N2V2R_T4(15)
  }
#undef N2V2R_T4
  return cudaGetLastError();
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

// contiguous input (ld == cols): the same test over the flat array with 16-byte loads
__global__ void __launch_bounds__(256) planes_needed_flat_kernel(const float* __restrict__ in, int64_t count, int* __restrict__ need) {
  int local = 1;
  stream_read<float>(in, count, [&](float x) {
    float r = x - __bfloat162float(__float2bfloat16_rn(x));
    if (r != 0.f) {
      r -= __bfloat162float(__float2bfloat16_rn(r));
      local = max(local, r != 0.f ? 3 : 2);
    }
  });
  local = __reduce_max_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local > 1) atomicMax(need, local);
}

// ---- symmetry check / transpose of bf16 planes: 64 x 64 tiles, 16-byte global accesses -------------------------
// A tile row is 64 x 2 B = 128 B = 8 threads x uint4; 256 threads move 32 tile rows per pass.  Shared tiles have a pitch of
// 66 elements (33 words): reading a column walks consecutive banks.  ld is a multiple of 64 (zero padded), so whole tile
// rows are always in bounds column-wise; rows >= n are treated as zero.
constexpr int kTile = 64, kPitch = 66;

struct TileRegs { uint4 v[2]; };

template <bool kMaskCols>
__device__ __forceinline__ TileRegs fetch_tile(const uint16_t* __restrict__ P, int64_t n, int64_t ld, int64_t row0, int64_t col0) {
  const int c8 = (threadIdx.x & 7) * 8, r0 = threadIdx.x >> 3;
  TileRegs t;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int64_t i = row0 + r0 + 32 * h;
    t.v[h] = (i < n) ? __ldg(reinterpret_cast<const uint4*>(P + i * ld + col0 + c8)) : make_uint4(0u, 0u, 0u, 0u);
  }
  if (kMaskCols && col0 + c8 + 8 > n) {                 // last tile column only: zero the elements at columns >= n
    const int keep = (int)(n - col0 - c8);              // elements of this thread's eight that are inside (may be <= 0)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t w[4] = {t.v[h].x, t.v[h].y, t.v[h].z, t.v[h].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (2 * q >= keep) w[q] = 0u;
        else if (2 * q + 1 >= keep) w[q] &= 0xffffu;
      }
      t.v[h] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  return t;
}

__device__ __forceinline__ void store_tile(const TileRegs& t, uint16_t (*sm)[kPitch]) {
  const int c8 = (threadIdx.x & 7) * 8, r0 = threadIdx.x >> 3;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sm[r0 + 32 * h][c8]);     // pitch 66 and c8 % 8 == 0: 4-byte aligned
    dst[0] = t.v[h].x; dst[1] = t.v[h].y; dst[2] = t.v[h].z; dst[3] = t.v[h].w;
  }
}

template <bool kMaskCols>
__device__ __forceinline__ void load_tile(const uint16_t* __restrict__ P, int64_t n, int64_t ld, int64_t row0, int64_t col0,
                                          uint16_t (*sm)[kPitch]) {
  store_tile(fetch_tile<kMaskCols>(P, n, ld, row0, col0), sm);
}

// Work item = (plane, tile pair bi <= bj): compare tile(bi,bj) with tile(bj,bi)^T.  Persistent CTAs walk the items with a
// stride of gridDim.x; the global loads of item k+1 are issued (into registers) before item k is compared out of shared
// memory, so every CTA always has 16 KB of reads in flight.
__global__ void __launch_bounds__(256) symmetric_kernel(const uint16_t* __restrict__ planes, int n_planes, int64_t n,
                                                        int64_t ld, int64_t pairs, int* __restrict__ asym) {
  __shared__ __align__(16) uint16_t a[kTile][kPitch], b[kTile][kPitch];
  const int64_t items = pairs * n_planes;
  auto fetch = [&](int64_t item, TileRegs& ra, TileRegs& rb) {
    const int64_t L = item % pairs;                      // L = bj (bj + 1) / 2 + bi
    const uint16_t* P = planes + (size_t)(item / pairs) * n * ld;
    int64_t bj = (int64_t)((sqrt(8.0 * (double)L + 1.0) - 1.0) * 0.5);
    while ((bj + 1) * (bj + 2) / 2 <= L) ++bj;
    while (bj * (bj + 1) / 2 > L) --bj;
    const int64_t bi = L - bj * (bj + 1) / 2;
    ra = fetch_tile<true>(P, n, ld, bi * kTile, bj * kTile);      // rows and columns >= n read as zero,
    rb = fetch_tile<true>(P, n, ld, bj * kTile, bi * kTile);      // so the compare needs no bounds
  };
  int bad = 0;
  TileRegs ra, rb;
  int64_t item = blockIdx.x;
  if (item < items) fetch(item, ra, rb);
  while (item < items) {
    store_tile(ra, a);
    store_tile(rb, b);
    __syncthreads();
    const int64_t next = item + gridDim.x;
    if (next < items) fetch(next, ra, rb);
    for (int e = threadIdx.x; e < kTile * kTile / 2; e += 256) {                // two rows of a against one word of b
      const int c = e & 63, r2 = (e >> 6) * 2;
      const uint32_t av = (uint32_t)a[r2][c] | ((uint32_t)a[r2 + 1][c] << 16);
      const uint32_t bv = *reinterpret_cast<const uint32_t*>(&b[c][r2]);
      bad |= (av != bv);
    }
    __syncthreads();
    item = next;
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(asym, 1);
}

// out tile (bx, by) = in tile (by, bx)^T; out rows < n, out columns < ld with zeros at columns >= n
__global__ void __launch_bounds__(256) transpose_kernel(const uint16_t* __restrict__ in, uint16_t* __restrict__ out,
                                                        int64_t n, int64_t ld) {
  __shared__ __align__(16) uint16_t t[kTile][kPitch];
  const size_t poff = (size_t)blockIdx.z * n * ld;
  const int64_t in_row0 = (int64_t)blockIdx.y * kTile, in_col0 = (int64_t)blockIdx.x * kTile;
  load_tile<false>(in + poff, n, ld, in_row0, in_col0, t);       // t[r][c] = in[in_row0 + r][in_col0 + c]
  __syncthreads();
  const int c8 = (threadIdx.x & 7) * 8, r0 = threadIdx.x >> 3;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int r = r0 + 32 * h;                            // out row = in column
    const int64_t i = in_col0 + r;
    if (i >= n) continue;
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int64_t j = in_row0 + c8 + 2 * q;             // out column = in row (rows >= n were loaded as zero)
      const uint32_t lo = t[c8 + 2 * q][r], hi = t[c8 + 2 * q + 1][r];
      (void)j;
      w[q] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(out + poff + i * ld + in_row0 + c8) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// ---- K0b: weighted edge list (COO, duplicates already resolved) -> dense symmetric matrix --------------------------
// out must be zero-filled; entries with an endpoint < 0 (node dropped by the node intersection) are skipped.
template <class T>
__global__ void __launch_bounds__(256) scatter_edges_kernel(const int32_t* __restrict__ iu, const int32_t* __restrict__ iv,
                                                            const double* __restrict__ w, int64_t n_edges, T* __restrict__ out,
                                                            int64_t n, int64_t ld) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (int64_t)gridDim.x * blockDim.x) {
    const int32_t a = iu[e], b = iv[e];
    if (a < 0 || b < 0 || a >= n || b >= n) continue;
    const T v = (T)w[e];
    out[(int64_t)a * ld + b] = v;
    out[(int64_t)b * ld + a] = v;
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
  N2V2R_REQUIRE(((uintptr_t)d_in & 15) == 0, "select_nonzero: input must be 16-byte aligned");
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
  const size_t esz = dtype == N2V2R_F32 ? 4 : 8;
  const bool fast = (dtype == N2V2R_F32 || dtype == N2V2R_F64) && cols % 4 == 0 && ld_in % 4 == 0 &&
                    ((uintptr_t)d_in & 15) == 0 && (ld_in * esz) % 16 == 0 &&
                    (!d_out_f32 || (ld_out % 4 == 0 && ((uintptr_t)d_out_f32 & 15) == 0)) &&
                    (!d_planes || ((uintptr_t)d_planes & 7) == 0);
  if (fast) {
    const int strips4 = (int)((width + 1023) / 1024);
    // one resident wave: 6 CTAs per SM fit for fp32 (32 KB ring, 40 registers), 3 for fp64 (64 KB ring)
    const int per_sm = dtype == N2V2R_F32 ? 6 : 3;
    int row_ctas4 = sm_count() * per_sm / strips4;
    if (row_ctas4 < 1) row_ctas4 = 1;
    if (row_ctas4 > (rows + 3) / 4) row_ctas4 = (int)((rows + 3) / 4);
    if (row_ctas4 > 65535) row_ctas4 = 65535;
    int rpc = (int)((rows + row_ctas4 - 1) / row_ctas4);
    dim3 grid4(strips4, (unsigned)((rows + rpc - 1) / rpc));
    if (dtype == N2V2R_F32)
      N2V2R_CUDA(launch_transform4<float>(s, grid4, (const float*)d_in, rows, cols, ld_in, absolute, has_thr, (float)thr, has_cut,
                                          (float)cut, binarize, d_out_f32, ld_out, (__nv_bfloat16*)d_planes, n_planes, ld_p, d_colsum, rpc));
    else
      N2V2R_CUDA(launch_transform4<double>(s, grid4, (const double*)d_in, rows, cols, ld_in, absolute, has_thr, thr, has_cut, cut,
                                           binarize, d_out_f32, ld_out, (__nv_bfloat16*)d_planes, n_planes, ld_p, d_colsum, rpc));
    N2V2R_CHECK_LAUNCH();
    return N2V2R_OK;
  }
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
  if (ld == cols && ((uintptr_t)d_in & 15) == 0) {
    planes_needed_flat_kernel<<<sm_count() * 8, 256, 0, s>>>(d_in, rows * cols, d_need);
  } else {
    dim3 grid((unsigned)std::min<int64_t>((cols + 255) / 256, 64), (unsigned)std::min<int64_t>(rows, (int64_t)sm_count() * 2));
    planes_needed_kernel<<<grid, 256, 0, s>>>(d_in, rows, cols, ld, d_need);
  }
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
  N2V2R_REQUIRE(ld_p % 64 == 0 && ((uintptr_t)d_planes & 15) == 0, "is_symmetric: ld_p must be a multiple of 64 and the planes 16-byte aligned");
  const unsigned tiles = (unsigned)((n + kTile - 1) / kTile);
  const int64_t pairs = (int64_t)tiles * (tiles + 1) / 2;
  const int64_t want = (int64_t)sm_count() * 8;
  const unsigned ctas = (unsigned)std::min<int64_t>(pairs * n_planes, want);
  symmetric_kernel<<<ctas, 256, 0, s>>>((const uint16_t*)d_planes, n_planes, n, ld_p, pairs, d_flag);
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
  N2V2R_REQUIRE(ld_p % 64 == 0 && ((uintptr_t)d_planes & 15) == 0 && ((uintptr_t)d_out & 15) == 0,
                "transpose_planes: ld_p must be a multiple of 64 and the planes 16-byte aligned");
  const unsigned tr = (unsigned)((n + kTile - 1) / kTile), tc = (unsigned)(ld_p / kTile);
  // blockIdx.x walks the input columns (output rows < n), blockIdx.y the input rows (output cols < ld)
  transpose_kernel<<<dim3(tr, tc, n_planes), 256, 0, s>>>((const uint16_t*)d_planes, (uint16_t*)d_out, n, ld_p);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_scatter_edges(void* stream, const int32_t* d_u, const int32_t* d_v, const double* d_w, int64_t n_edges,
                                   void* d_out, int dtype, int64_t n, int64_t ld) {
  N2V2R_REQUIRE(d_out && n > 0 && ld >= n && n_edges >= 0, "scatter_edges: bad arguments");
  N2V2R_REQUIRE(n_edges == 0 || (d_u && d_v && d_w), "scatter_edges: null edge arrays");
  N2V2R_REQUIRE(dtype == N2V2R_F32 || dtype == N2V2R_F64, "scatter_edges: unknown dtype %d", dtype);
  if (n_edges == 0) return N2V2R_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = (int)std::min<int64_t>((n_edges + 255) / 256, (int64_t)sm_count() * 8);
  if (dtype == N2V2R_F32) scatter_edges_kernel<float><<<grid, 256, 0, s>>>(d_u, d_v, d_w, n_edges, (float*)d_out, n, ld);
  else scatter_edges_kernel<double><<<grid, 256, 0, s>>>(d_u, d_v, d_w, n_edges, (double*)d_out, n, ld);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
