// K3/K4 for layers stored as ONE e4m3 plane (1 byte per entry): binarised layers ({0,1}, preprocessing_utils.py:138-139) and
// any layer whose entries are exactly representable in e4m3.  Half the HBM bytes of a bf16 plane (north_star (1): "binarized
// layers are held as exact bf16/fp8 to halve HBM bytes").
//
//   D[128 x 6l] (TMEM, fp32) += A_tile[128 x 128] (smem, TMA, e4m3, SWIZZLE_128B)
//                               * [X_0 | ... | X_5]^T tile [6l x 128] (smem, TMA, e4m3)         tcgen05.mma kind::f8f6f4
//
// The skinny fp32 operand X is split into SIX e4m3 terms with a power-of-two scale per column s_c and a factor 16 per term:
//   v = x s_c (|v| <= 256),  t_0 = e4m3(v),  r_1 = v - t_0,  t_i = e4m3(16^i r_i),  r_{i+1} = r_i - t_i / 16^i
// (every subtraction is exact in fp32), so x = (sum_i t_i 16^-i) / s_c to 24 bits wherever |v| >= 2^-6, and to 2^-38 of the
// column maximum below.  Each term has its own fp32 accumulator columns in TMEM; the epilogue recombines them in Horner form.
// kind::f8f6f4 runs at twice the bf16 rate with K = 32 per instruction, so the six terms cost what the three bf16 terms of the
// bf16-plane kernel cost, and the pass is bound by the halved HBM stream.
// Same warp roles, K-split partials and fixed-order reduction as k3_tcgen05.cu.  Replaces the same reference code (ARPACK
// operator matvec / rmatvec, scipy _svds.py:442-457, under spectral_embedding.UASE at model.py:54-55).
#include "k5_orth.cuh"
#include "tc_ptx.cuh"
#include <cuda_fp8.h>
#include <algorithm>

namespace n2v2r {

// fill = 1: the A tile stays in the collector buffer for the next MMA; fill = 0: it is taken from there (last use)
template <int FILL>
__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if (FILL)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4.collector::a::fill [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4.collector::a::lastuse [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// kind::f8f6f4 instruction descriptor: D = f32, A = B = e4m3 (format code 0), both K-major, M = 128
__device__ __forceinline__ uint32_t idesc_e4m3(int N) { return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24); }

__device__ __forceinline__ float e4m3_to_float(uint8_t b) {
  return __half2float(__half(__nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, __NV_E4M3)));
}
__device__ __forceinline__ uint8_t float_to_e4m3(float x) { return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3); }

// ---- operand: column maxima, then the six-term split -------------------------------------------------
// stage 1: per-CTA column maxima of |X| (no atomics: part[cta][col]); stage 2 (colmax_final_kernel): maximum over the CTAs
__global__ void __launch_bounds__(256) colmax_kernel(const float* __restrict__ X, int64_t n, int l, float* __restrict__ part) {
  __shared__ float sm[256];
  const int rpb = 256 / l;
  const int col = threadIdx.x % l, r = threadIdx.x / l;
  float m = 0.f;
  if (r < rpb)
    for (int64_t row = (int64_t)blockIdx.x * rpb + r; row < n; row += (int64_t)gridDim.x * rpb) m = fmaxf(m, fabsf(X[row * l + col]));
  sm[threadIdx.x] = m;
  __syncthreads();
  if ((int)threadIdx.x < l) {
    for (int rr = 1; rr < rpb; ++rr) m = fmaxf(m, sm[rr * l + threadIdx.x]);
    part[(size_t)blockIdx.x * l + threadIdx.x] = m;
  }
}
__global__ void __launch_bounds__(128) colmax_final_kernel(const float* __restrict__ part, int n_parts, int l, uint32_t* __restrict__ colmax_bits) {
  const int col = threadIdx.x;
  if (col >= l) return;
  float m = 0.f;
  for (int p = 0; p < n_parts; ++p) m = fmaxf(m, part[(size_t)p * l + col]);
  colmax_bits[col] = __float_as_uint(m);
}

// X fp32 [n][l] -> Xs e4m3 [6][l][n_pad] (K-major operand), inv_scale[c] = 1 / s_c
__global__ void __launch_bounds__(256) split8_transpose_kernel(const float* __restrict__ X, int64_t n, int l, int64_t n_pad,
                                                               const uint32_t* __restrict__ colmax_bits, uint8_t* __restrict__ Xs,
                                                               float* __restrict__ inv_scale) {
  __shared__ float t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t rb = (int64_t)blockIdx.x * 32;
  const int cb = blockIdx.y * 32;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = rb + r; const int j = cb + tx;
    t[r][tx] = (i < n && j < l) ? X[i * l + j] : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int j = cb + r; const int64_t i = rb + tx;
    if (j < l && i < n_pad) {
      const float cm = __uint_as_float(colmax_bits[j]);
      int e = 0;
      if (cm > 0.f) frexpf(cm, &e);                       // cm = m 2^e, m in [0.5, 1)  ->  cm 2^(8-e) in [128, 256)
      e = e < -100 ? -100 : e;
      const float s = cm > 0.f ? ldexpf(1.f, 8 - e) : 1.f;
      if (blockIdx.x == 0 && tx == 0) inv_scale[j] = 1.f / s;
      float rres = t[tx][r] * s;                          // exact (power of two), |v| <= 256
      float up = 1.f, down = 1.f;
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        const uint8_t b = float_to_e4m3(rres * up);
        Xs[((size_t)q * l + j) * n_pad + i] = b;
        rres -= e4m3_to_float(b) * down;                  // exact
        up *= 16.f; down *= 0.0625f;
      }
    }
  }
}

// ---- layer: one bf16 plane -> e4m3 bytes (rows x ld8, padding zero), flag = 1 when some value is not representable ----
__global__ void __launch_bounds__(256) plane_to_e4m3_kernel(const __nv_bfloat16* __restrict__ plane, int64_t rows, int64_t cols, int64_t ld_p,
                                                            uint8_t* __restrict__ out, int64_t ld8, int* __restrict__ inexact) {
  const int64_t total = rows * (ld8 / 4);
  int bad = 0;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / (ld8 / 4), c0 = (e - r * (ld8 / 4)) * 4;
    uint32_t pack = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int64_t c = c0 + q;
      const float v = c < cols ? __bfloat162float(plane[r * ld_p + c]) : 0.f;
      const uint8_t b = float_to_e4m3(v);
      if (e4m3_to_float(b) != v) bad = 1;
      pack |= (uint32_t)b << (8 * q);
    }
    *reinterpret_cast<uint32_t*>(out + r * ld8 + c0) = pack;
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicMax(inexact, 1);
}

// ---- the product kernel -------------------------------------------------------------------------------
struct Stream8Args {
  int l, stages, n_tiles, n_splits, k_chunk;       // k_chunk: multiple of 128
  int64_t k_pad;                                   // contraction length, multiple of 128
  int64_t rows_valid, out_ld, split_stride;
  uint32_t stage_bytes, x_bytes, tmem_cols;
  const float* inv_scale;
  float* out;
};

__global__ void __launch_bounds__(192, 1) stream_mma8_kernel(const __grid_constant__ CUtensorMap mapA,
                                                             const __grid_constant__ CUtensorMap mapX, const Stream8Args a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * 8 + 2];
  __shared__ uint32_t tmem_base_smem;
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t bar_tfull = smem_u32(&bars[16]), bar_tempty = smem_u32(&bars[17]);
  const int n_units = a.n_tiles * a.n_splits;
  const int n3 = 3 * a.l;

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(a.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int split = u / a.n_tiles, tile = u - split * a.n_tiles;
        const int64_t k_begin = (int64_t)split * a.k_chunk;
        const int64_t k_end = (k_begin + a.k_chunk < a.k_pad) ? k_begin + a.k_chunk : a.k_pad;
        for (int64_t k0 = k_begin; k0 < k_end; k0 += 128) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
          const uint32_t base = smem0 + stage * a.stage_bytes;
          mbar_expect_tx(bar_full + 8 * stage, a.stage_bytes);
          tma_load_2d(&mapA, bar_full + 8 * stage, base, (int)k0, tile * 128);
          tma_load_3d(&mapX, bar_full + 8 * stage, base + 16384, (int)k0, 0, 0);
          if (++stage == a.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    int stage = 0; uint32_t phase = 0, uphase = 0;
    const uint32_t idesc = idesc_e4m3(n3);
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles;
      const int64_t k_begin = (int64_t)split * a.k_chunk;
      const int64_t k_end = (k_begin + a.k_chunk < a.k_pad) ? k_begin + a.k_chunk : a.k_pad;
      mbar_wait(bar_tempty, uphase ^ 1u);
      tc_fence_after();
      bool first = true;
      for (int64_t k0 = k_begin; k0 < k_end; k0 += 128) {
        mbar_wait(bar_full + 8 * stage, phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t abase = smem0 + stage * a.stage_bytes;
          const uint32_t xbase = abase + 16384;
#pragma unroll
          for (int k = 0; k < 4; ++k) {                    // K = 32 e4m3 elements = 32 bytes per instruction
            const uint32_t acc = (first && k == 0) ? 0u : 1u;
            // both term groups meet the same A k-slice: read from shared memory once, kept in the collector for the second MMA
            umma_f8<1>(tmem_base, sw128_desc(abase + k * 32), sw128_desc(xbase + k * 32), idesc, acc);                    // terms 0..2
            umma_f8<0>(tmem_base + n3, sw128_desc(abase + k * 32), sw128_desc(xbase + n3 * 128 + k * 32), idesc, acc);    // terms 3..5
          }
          umma_commit(bar_empty + 8 * stage);
        }
        __syncwarp();
        first = false;
        if (++stage == a.stages) { stage = 0; phase ^= 1u; }
      }
      if (lane == 0) umma_commit(bar_tfull);
      __syncwarp();
      uphase ^= 1u;
    }
  } else {
    const int q = warp & 3;
    uint32_t uphase = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles, tile = u - split * a.n_tiles;
      mbar_wait(bar_tfull, uphase);
      tc_fence_after();
      const int64_t row = (int64_t)tile * 128 + q * 32 + lane;
      float* dst = a.out + (size_t)split * a.split_stride + (size_t)row * a.out_ld;
      const bool row_ok = row < a.rows_valid;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      for (int c0 = 0; c0 < a.l; c0 += 8) {
        uint32_t t[6][8];
#pragma unroll
        for (int g = 0; g < 6; ++g) tmem_ld8(taddr + g * a.l + c0, t[g]);
        tmem_ld_wait();
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float s = __uint_as_float(t[5][e]);
#pragma unroll
          for (int g = 4; g >= 0; --g) s = fmaf(s, 0.0625f, __uint_as_float(t[g][e]));
          v[e] = s * a.inv_scale[c0 + e];
        }
        if (row_ok) {
          *reinterpret_cast<float4*>(dst + c0) = make_float4(v[0], v[1], v[2], v[3]);
          *reinterpret_cast<float4*>(dst + c0 + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty);
      uphase ^= 1u;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols));
  }
}

// ---- host ---------------------------------------------------------------------------------------------
int stream_apply_e4m3(cudaStream_t s, const StreamPlan& plan, const void* plane8, int64_t ld8, const float* x, void* xs_ws,
                      float* parts, float* out, int accumulate) {
  N2V2R_REQUIRE(xs_ws && parts && plane8, "e4m3 product needs its workspace");
  N2V2R_REQUIRE(ld8 % 128 == 0 && ld8 >= plan.n, "e4m3 layer: ld must be a multiple of 128 >= n");
  const int l = plan.l;
  const int64_t n_pad = round_up(plan.n, 128);
  // workspace: [6][l][n_pad] e4m3 | column maxima (l uint32) | inverse scales (l floats)
  uint8_t* xs8 = (uint8_t*)xs_ws;
  uint32_t* colmax = (uint32_t*)(xs8 + align_up((size_t)6 * l * n_pad));
  float* inv_scale = (float*)(colmax + 128);
  float* cm_part = inv_scale + 128;                       // [kColmaxCtas][l]
  constexpr int kColmaxCtas = 96;
  N2V2R_REQUIRE(align_up((size_t)6 * l * n_pad) + 1024 + sizeof(float) * kColmaxCtas * 80 <= plan.xs_bytes,
                "e4m3 product: operand workspace too small");
  colmax_kernel<<<kColmaxCtas, 256, 0, s>>>(x, plan.n, l, cm_part);
  N2V2R_CHECK_LAUNCH();
  colmax_final_kernel<<<1, 128, 0, s>>>(cm_part, kColmaxCtas, l, colmax);
  N2V2R_CHECK_LAUNCH();
  dim3 g2((unsigned)((n_pad + 31) / 32), (unsigned)((l + 31) / 32));
  split8_transpose_kernel<<<g2, 256, 0, s>>>(x, plan.n, l, n_pad, colmax, xs8, inv_scale);
  N2V2R_CHECK_LAUNCH();
  Stream8Args a;
  a.l = l;
  a.k_pad = n_pad;
  a.k_chunk = (int)round_up(plan.k_chunk, 128);
  a.n_splits = (int)((n_pad + a.k_chunk - 1) / a.k_chunk);
  N2V2R_REQUIRE(a.n_splits <= plan.n_splits + 1, "e4m3 product: K split exceeds the workspace plan");
  a.n_tiles = (int)((plan.n + 127) / 128);
  a.x_bytes = 6u * l * 128u;
  a.stage_bytes = 16384u + a.x_bytes;
  const uint32_t budget = 227u * 1024u - 2048u;
  a.stages = (int)(budget / a.stage_bytes);
  if (a.stages > 8) a.stages = 8;
  N2V2R_REQUIRE(a.stages >= 2, "e4m3 product: stage does not fit twice in shared memory");
  { int tc = 32; while (tc < 6 * l) tc <<= 1; N2V2R_REQUIRE(tc <= 512, "e4m3 product: accumulators do not fit in TMEM"); a.tmem_cols = (uint32_t)tc; }
  const int64_t rows_pad = round_up(plan.n, 256);
  a.rows_valid = rows_pad;
  a.out_ld = l;
  a.split_stride = rows_pad * l;
  a.inv_scale = inv_scale;
  a.out = parts;
  CUtensorMap mA, mX;
  {
    const uint64_t dims[2] = {(uint64_t)ld8, (uint64_t)plan.n};
    const uint64_t str[1] = {(uint64_t)ld8};
    const uint32_t box[2] = {128, 128};
    int rc = make_map_u8(&mA, plane8, 2, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)n_pad, (uint64_t)l, 6};
    const uint64_t str[2] = {(uint64_t)n_pad, (uint64_t)l * n_pad};
    const uint32_t box[3] = {128, (uint32_t)l, 6};
    int rc = make_map_u8(&mX, xs8, 3, dims, str, box);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(stream_mma8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024)));
    attr_set = true;
  }
  const size_t smem = (size_t)a.stages * a.stage_bytes + 1024;
  const int n_units = a.n_tiles * a.n_splits;
  const int grid = n_units < sm_count() ? n_units : sm_count();
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int rc = mma_events_begin(s, &e0, &e1);
  if (rc) return rc;
  stream_mma8_kernel<<<grid, 192, smem, s>>>(mA, mX, a);
  N2V2R_CHECK_LAUNCH();
  rc = mma_events_end(s, e0, e1);
  if (rc) return rc;
  return launch_reduce_parts(s, parts, a.n_splits, rows_pad, plan.n, l, out, accumulate);
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" int n2v2r_plane_to_e4m3(void* stream, const void* d_plane, int64_t rows, int64_t cols, int64_t ld_p, void* d_out,
                                   int64_t ld8, int* h_exact) {
  N2V2R_REQUIRE(d_plane && d_out && rows > 0 && cols > 0 && ld_p >= cols && ld8 >= cols && ld8 % 128 == 0, "plane_to_e4m3: bad arguments");
  N2V2R_REQUIRE(((uintptr_t)d_out & 15) == 0, "plane_to_e4m3: output must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  int* d_flag = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_flag, sizeof(int), s));
  cudaError_t le = cudaMemsetAsync(d_flag, 0, sizeof(int), s);
  const int64_t total = rows * (ld8 / 4);
  const unsigned grid = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)sm_count() * 16);
  if (le == cudaSuccess) {
    plane_to_e4m3_kernel<<<grid, 256, 0, s>>>((const __nv_bfloat16*)d_plane, rows, cols, ld_p, (uint8_t*)d_out, ld8, d_flag);
    le = cudaGetLastError();
  }
  int flag = 0;
  if (le == cudaSuccess) le = cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (le == cudaSuccess) le = cudaStreamSynchronize(s);
  cudaFreeAsync(d_flag, s);
  if (le != cudaSuccess) return cuda_fail(le, "plane_to_e4m3", __FILE__, __LINE__);
  if (h_exact) *h_exact = flag ? 0 : 1;
  return N2V2R_OK;
}
