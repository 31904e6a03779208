// K3/K4 on the 5th-gen tensor cores: out = A X for a bf16-plane layer A (n x n, K-major) and
// a skinny fp32 X (n x l) that is split into three bf16 terms concatenated along N.
//
//   D[128 x 3l] (TMEM, fp32) += A_tile[128 x 64] (smem, TMA, SWIZZLE_128B)
//                               * [X_hi | X_mid | X_lo]^T tile [3l x 64] (smem, TMA)
//
// Warp roles (192 threads): warp 0 TMA producer, warp 1 TMEM owner + MMA issuer (one elected
// lane), warps 2-5 epilogue (tcgen05.ld -> hi+mid+lo -> global partials).  Persistent CTAs walk
// (row tile, K split) units; K splits bound the length of one fp32 accumulation chain and are
// summed in fixed order by reduce_parts_kernel (deterministic, no atomics).
// Plane p of a multi-plane layer multiplies only the splits s <= 2-p (terms below 2^-24 dropped).
#include "k5_orth.cuh"
#include "tc_ptx.cuh"
#include <vector>

namespace n2v2r {

struct StreamArgs {
  int n_planes, l, mh;              // planes, block width, 128-row halves per tile (1 or 2)
  int stages, n_tiles, n_splits, k_chunk;
  int n_jb;                         // column blocks of width l of the right operand (1 for the skinny products)
  int nb;                           // operand blocks multiplied against ONE read of the A tile (seed batching; 1 otherwise)
  int upper_only;                   // SYRK: skip units entirely below the diagonal (projection)
  int col_batched;                  // projection: the nb operand blocks of a unit are the column blocks jb nb .. jb nb + nb - 1 of ONE
                                    // operand (one TMA load each); 0: seed batching, the blocks are stacked operands (one load)
  int64_t k_pad;                    // contraction length, multiple of 64
  int64_t rows_valid, cols_valid;   // extent of the output that may be written
  int64_t out_ld, split_stride;     // output leading dimension / elements between K-split partials
  uint32_t stage_bytes, a_plane_bytes, x_bytes;
  float* out;
};

__global__ void __launch_bounds__(192, 1) stream_mma_kernel(const __grid_constant__ CUtensorMap mapA0,
                                                            const __grid_constant__ CUtensorMap mapA1,
                                                            const __grid_constant__ CUtensorMap mapA2,
                                                            const __grid_constant__ CUtensorMap mapX, const StreamArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * 8 + 2];
  __shared__ uint32_t tmem_base_smem;
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t bar_tfull = smem_u32(&bars[16]), bar_tempty = smem_u32(&bars[17]);
  const int n_units = a.n_tiles * a.n_splits * a.n_jb;
  const int nacc = 3 * a.l;
  uint32_t tmem_cols = 32u;
  while (tmem_cols < (uint32_t)(a.nb * a.mh * nacc)) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      const CUtensorMap* maps[3] = {&mapA0, &mapA1, &mapA2};
      int stage = 0; uint32_t phase = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int sj = u / a.n_tiles, tile = u - sj * a.n_tiles;
        const int split = sj / a.n_jb, jb = sj - split * a.n_jb;
        const int64_t k_begin = (int64_t)split * a.k_chunk;
        const int64_t k_end = (k_begin + a.k_chunk < a.k_pad) ? k_begin + a.k_chunk : a.k_pad;
        const int row0 = tile * 128 * a.mh;
        if (a.upper_only && (jb + 1) * (a.col_batched ? a.nb : 1) * a.l <= row0) continue;
        for (int64_t k0 = k_begin; k0 < k_end; k0 += 64) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
          const uint32_t base = smem0 + stage * a.stage_bytes;
          mbar_expect_tx(bar_full + 8 * stage, a.stage_bytes);
          for (int p = 0; p < a.n_planes; ++p)
            tma_load_2d(maps[p], bar_full + 8 * stage, base + p * a.a_plane_bytes, (int)k0, row0);
          if (a.col_batched) {
            for (int bk = 0; bk < a.nb; ++bk)
              tma_load_3d(&mapX, bar_full + 8 * stage, base + a.n_planes * a.a_plane_bytes + bk * a.x_bytes, (int)k0, (jb * a.nb + bk) * a.l, 0);
          } else {
            tma_load_3d(&mapX, bar_full + 8 * stage, base + a.n_planes * a.a_plane_bytes, (int)k0, jb * a.l, 0);
          }
          if (++stage == a.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int stage = 0; uint32_t phase = 0, uphase = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int sj_ = u / a.n_tiles;
      const int split = sj_ / a.n_jb;
      if (a.upper_only && ((sj_ - split * a.n_jb) + 1) * (a.col_batched ? a.nb : 1) * a.l <= (u - sj_ * a.n_tiles) * 128 * a.mh) continue;
      const int64_t k_begin = (int64_t)split * a.k_chunk;
      const int64_t k_end = (k_begin + a.k_chunk < a.k_pad) ? k_begin + a.k_chunk : a.k_pad;
      mbar_wait(bar_tempty, uphase ^ 1u);            // epilogue has drained the accumulators
      tc_fence_after();
      bool first = true;
      for (int64_t k0 = k_begin; k0 < k_end; k0 += 64) {
        mbar_wait(bar_full + 8 * stage, phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t base = smem0 + stage * a.stage_bytes;
          const uint32_t xbase = base + a.n_planes * a.a_plane_bytes;
          if (a.nb == 1) {
            for (int h = 0; h < a.mh; ++h) {
              bool acc = !first;
              for (int p = 0; p < a.n_planes; ++p) {
                const uint32_t idesc = idesc_bf16((3 - p) * a.l);
                const uint32_t abase = base + p * a.a_plane_bytes + h * 16384;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  umma_bf16(tmem_base + h * nacc, sw128_desc(abase + k * 32), sw128_desc(xbase + k * 32), idesc, acc ? 1u : 0u);
                  acc = true;
                }
              }
            }
          } else {
            // seed batching (mh == 1): the nb operand blocks meet the same A k-slice back to back; it is read from shared memory
            // by the first MMA (collector fill) and taken from the collector buffer by the others
            for (int p = 0; p < a.n_planes; ++p) {
              const uint32_t idesc = idesc_bf16((3 - p) * a.l);
              const uint32_t abase = base + p * a.a_plane_bytes;
              const uint32_t acc0 = (first && p == 0) ? 0u : 1u;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t acc = k == 0 ? acc0 : 1u;
                const uint64_t ad = sw128_desc(abase + k * 32);
                umma_bf16_collector(tmem_base, ad, sw128_desc(xbase + k * 32), idesc, acc, kCollectorFill);
                if (a.nb == 3)
                  umma_bf16_collector(tmem_base + nacc, ad, sw128_desc(xbase + a.x_bytes + k * 32), idesc, acc, kCollectorUse);
                umma_bf16_collector(tmem_base + (a.nb - 1) * nacc, ad, sw128_desc(xbase + (a.nb - 1) * a.x_bytes + k * 32), idesc, acc,
                                    kCollectorLastUse);
              }
            }
          }
          umma_commit(bar_empty + 8 * stage);         // frees the smem stage when these MMAs retire
        }
        __syncwarp();
        first = false;
        if (++stage == a.stages) { stage = 0; phase ^= 1u; }
      }
      if (lane == 0) umma_commit(bar_tfull);          // accumulators complete
      __syncwarp();
      uphase ^= 1u;
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global partials =====
    const int q = warp & 3;                           // TMEM lane quarter this warp may access
    uint32_t uphase = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int sj = u / a.n_tiles, tile = u - sj * a.n_tiles;
      const int split = sj / a.n_jb, jb = sj - split * a.n_jb;
      if (a.upper_only && (jb + 1) * (a.col_batched ? a.nb : 1) * a.l <= tile * 128 * a.mh) continue;
      mbar_wait(bar_tfull, uphase);
      tc_fence_after();
      for (int bk = 0; bk < a.nb; ++bk)
      for (int h = 0; h < a.mh; ++h) {
        const int64_t row = (int64_t)tile * 128 * a.mh + h * 128 + q * 32 + lane;
        const int jcol = a.col_batched ? jb * a.nb + bk : jb + bk;
        float* dst = a.out + (size_t)split * a.split_stride + (size_t)row * a.out_ld + (size_t)jcol * a.l;
        const int64_t col_base = (int64_t)jcol * a.l;
        const bool row_ok = row < a.rows_valid;
        const bool vec_ok = row_ok && (a.out_ld % 4 == 0) && (col_base + a.l <= a.cols_valid);
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (bk * a.mh + h) * nacc;
        for (int c0 = 0; c0 < a.l; c0 += 16) {
          uint32_t hi[16], mid[16], lo[16];
          tmem_ld16(taddr + c0, hi);
          tmem_ld16(taddr + a.l + c0, mid);
          tmem_ld16(taddr + 2 * a.l + c0, lo);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            float4 v;
            v.x = (__uint_as_float(lo[e]) + __uint_as_float(mid[e])) + __uint_as_float(hi[e]);
            v.y = (__uint_as_float(lo[e + 1]) + __uint_as_float(mid[e + 1])) + __uint_as_float(hi[e + 1]);
            v.z = (__uint_as_float(lo[e + 2]) + __uint_as_float(mid[e + 2])) + __uint_as_float(hi[e + 2]);
            v.w = (__uint_as_float(lo[e + 3]) + __uint_as_float(mid[e + 3])) + __uint_as_float(hi[e + 3]);
            if (vec_ok) {
              *reinterpret_cast<float4*>(dst + c0 + e) = v;
            } else if (row_ok) {
              const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int t = 0; t < 4; ++t)
                if (col_base + c0 + e + t < a.cols_valid) dst[c0 + e + t] = vv[t];
            }
          }
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
  }
}

// out[row][c] (= or +=) sum_s parts[s][row][c]   (fixed order)
__global__ void __launch_bounds__(256) reduce_parts_kernel(const float* __restrict__ parts, int n_splits, int64_t rows_pad,
                                                           int64_t n, int l, float* __restrict__ out, int accumulate) {
  const int64_t e4 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t total4 = n * l / 4;
  if (e4 >= total4) return;
  float4 s = accumulate ? reinterpret_cast<const float4*>(out)[e4] : make_float4(0.f, 0.f, 0.f, 0.f);
  for (int sp = 0; sp < n_splits; ++sp) {
    const float4 v = reinterpret_cast<const float4*>(parts + (size_t)sp * rows_pad * l)[e4];
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  reinterpret_cast<float4*>(out)[e4] = s;
}


// the same with the partials of one operand block inside a wider row: parts[s][row][col_off + c], row stride ld_parts
__global__ void __launch_bounds__(256) reduce_parts_strided_kernel(const float* __restrict__ parts, int n_splits, int64_t split_stride,
                                                                   int64_t ld_parts, int col_off, int64_t n, int l,
                                                                   float* __restrict__ out, int accumulate) {
  const int64_t e4 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t total4 = n * l / 4;
  if (e4 >= total4) return;
  const int l4 = l / 4;
  const int64_t row = e4 / l4;
  const int c4 = (int)(e4 - row * l4);
  float4 s = accumulate ? reinterpret_cast<const float4*>(out)[e4] : make_float4(0.f, 0.f, 0.f, 0.f);
  for (int sp = 0; sp < n_splits; ++sp) {
    const float4 v = *reinterpret_cast<const float4*>(parts + (size_t)sp * split_stride + row * ld_parts + col_off + 4 * c4);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  reinterpret_cast<float4*>(out)[e4] = s;
}

// ---- projection epilogue: out[i][j] = out[j][i] = sum_s parts[s][i][j] for i <= j (fixed order; exactly symmetric) ----
__global__ void __launch_bounds__(256) mirror_reduce_kernel(const float* __restrict__ parts, int n_splits, int64_t split_stride,
                                                            int64_t m, float* __restrict__ out, int64_t ld_out) {
  __shared__ float t[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bi > bj) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = (int64_t)bi * 32 + r, j = (int64_t)bj * 32 + tx;
    float v = 0.f;
    if (i < m && j < m) {
      const int64_t ii = i <= j ? i : j, jj = i <= j ? j : i;      // diagonal tiles: read the upper element
      for (int sp = 0; sp < n_splits; ++sp) v += parts[(size_t)sp * split_stride + ii * m + jj];
      out[i * ld_out + j] = v;
    }
    t[r][tx] = v;
  }
  __syncthreads();
  if (bi == bj) return;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = (int64_t)bj * 32 + r, j = (int64_t)bi * 32 + tx;       // out[j_orig][i_orig]
    if (i < m && j < m) out[i * ld_out + j] = t[tx][r];
  }
}

// ---- projection operand: W (r x c, f32/f64) -> Ws bf16 [3][m][k_pad], optionally transposed ------
template <class T>
__global__ void __launch_bounds__(256) split_planes_kernel(const T* __restrict__ W, int64_t rows, int64_t cols, int64_t ld,
                                                           int transpose, int64_t m, int64_t k_pad,
                                                           __nv_bfloat16* __restrict__ Ws) {
  __shared__ double t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  // output element (i, kk): transpose ? W[kk][i] : W[i][kk]
  const int64_t ib = (int64_t)blockIdx.y * 32, kb = (int64_t)blockIdx.x * 32;
  for (int r = ty; r < 32; r += 8) {
    double v = 0.0;
    if (transpose) {
      const int64_t src_r = kb + r, src_c = ib + tx;            // coalesced along i
      if (src_r < rows && src_c < cols) v = (double)W[src_r * ld + src_c];
      t[tx][r] = v;                                             // t[i_local][k_local]
    } else {
      const int64_t src_r = ib + r, src_c = kb + tx;            // coalesced along k
      if (src_r < rows && src_c < cols) v = (double)W[src_r * ld + src_c];
      t[r][tx] = v;
    }
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = ib + r, kk = kb + tx;
    if (i < m && kk < k_pad) {
      const double x = t[r][tx];
      const __nv_bfloat16 hi = __float2bfloat16_rn((float)x);
      const double r1 = x - (double)__bfloat162float(hi);
      const __nv_bfloat16 mid = __float2bfloat16_rn((float)r1);
      const double r2 = r1 - (double)__bfloat162float(mid);
      const __nv_bfloat16 lo = __float2bfloat16_rn((float)r2);
      Ws[((size_t)0 * m + i) * k_pad + kk] = hi;
      Ws[((size_t)1 * m + i) * k_pad + kk] = mid;
      Ws[((size_t)2 * m + i) * k_pad + kk] = lo;
    }
  }
}

// ---- host ---------------------------------------------------------------------------------------
// Optional per-launch timing of the MMA kernel alone (roofline evidence): event pairs are appended to
// the sink when one is installed by the caller (uase solver / layer_apply).
static thread_local std::vector<cudaEvent_t>* g_mma_events = nullptr;
void set_mma_event_sink(std::vector<cudaEvent_t>* sink) { g_mma_events = sink; }
int mma_events_begin(cudaStream_t s, cudaEvent_t* e0, cudaEvent_t* e1) {
  *e0 = *e1 = nullptr;
  if (!g_mma_events) return N2V2R_OK;
  N2V2R_CUDA(cudaEventCreate(e0));
  N2V2R_CUDA(cudaEventCreate(e1));
  N2V2R_CUDA(cudaEventRecord(*e0, s));
  return N2V2R_OK;
}
int mma_events_end(cudaStream_t s, cudaEvent_t e0, cudaEvent_t e1) {
  if (!g_mma_events || !e0) return N2V2R_OK;
  N2V2R_CUDA(cudaEventRecord(e1, s));
  g_mma_events->push_back(e0);
  g_mma_events->push_back(e1);
  return N2V2R_OK;
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
  }
  return fn;
}

// bf16 tensor map, SWIZZLE_128B, rank 2 or 3 (inner box = 64 elements = 128 B)
static int make_map_typed(CUtensorMap* map, CUtensorMapDataType dt, const void* base, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes, const uint32_t* box) {
  PFN_cuTensorMapEncodeTiled_v12000 enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return N2V2R_ECUDA; }
  cuuint64_t d[3]; cuuint64_t st[2]; cuuint32_t bx[3]; cuuint32_t es[3] = {1, 1, 1};
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) st[i] = strides_bytes[i];
  CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void*>(base), d, st, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with %d", (int)r); return N2V2R_ECUDA; }
  return N2V2R_OK;
}
int make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
             const uint32_t* box) {
  return make_map_typed(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides_bytes, box);
}
int make_map_u8(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box) {
  return make_map_typed(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, base, rank, dims, strides_bytes, box);
}

// D[m_rows x (n_jb*l)] = A[m_rows x k] * B^T, A = n_planes planes (row stride lda), B = 3 split planes [3][b_rows][ldb]
static int launch_mma(cudaStream_t s, const void* a_planes, int n_planes, int64_t m_rows, int64_t lda, size_t a_plane_stride,
                      const void* b_planes, int64_t b_rows, int64_t ldb, int l, int n_jb, int64_t k_pad, int k_chunk,
                      float* out, int64_t out_ld, int64_t split_stride, int64_t rows_valid, int64_t cols_valid, int upper_only = 0,
                      int nb = 1, int col_batched = 0) {
  StreamArgs a;
  a.upper_only = upper_only;
  a.col_batched = col_batched;
  a.n_planes = n_planes;
  a.l = l;
  a.nb = nb;
  a.mh = (n_planes == 1 && nb == 1) ? 2 : 1;
  N2V2R_REQUIRE(nb >= 1 && nb * a.mh * 3 * l <= 512, "tcgen05 path: %d operand blocks of width %d do not fit in TMEM", nb, l);
  a.k_chunk = k_chunk;
  a.k_pad = k_pad;
  a.n_splits = (int)((k_pad + k_chunk - 1) / k_chunk);
  a.n_jb = col_batched ? (n_jb + nb - 1) / nb : n_jb;
  a.n_tiles = (int)((m_rows + 128 * a.mh - 1) / (128 * a.mh));
  a.a_plane_bytes = 16384u * a.mh;
  a.x_bytes = 3u * l * 128u;                       // multiple of 1024 because l % 16 == 0
  a.stage_bytes = n_planes * a.a_plane_bytes + nb * a.x_bytes;
  a.out = out;
  a.out_ld = out_ld;
  a.split_stride = split_stride;
  a.rows_valid = rows_valid;
  a.cols_valid = cols_valid;
  const uint32_t budget = 227u * 1024u - 2048u;
  a.stages = (int)(budget / a.stage_bytes);
  if (a.stages > 8) a.stages = 8;
  N2V2R_REQUIRE(a.stages >= 2, "tcgen05 path: stage does not fit twice in shared memory");
  CUtensorMap mA[3], mX;
  for (int p = 0; p < 3; ++p) {
    const int pp = p < n_planes ? p : 0;
    const uint64_t dims[2] = {(uint64_t)lda, (uint64_t)m_rows};
    const uint64_t str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 128u * a.mh};
    int rc = make_map(&mA[p], (const char*)a_planes + pp * a_plane_stride, 2, dims, str, box);
    if (rc) return rc;
  }
  {
    const int stacked = col_batched ? 1 : nb;                                           // seed batching: [nb][3][l][ldb], b_rows == l
    const uint64_t dims[3] = {(uint64_t)ldb, (uint64_t)b_rows, (uint64_t)(3 * stacked)};
    const uint64_t str[2] = {(uint64_t)ldb * 2, (uint64_t)b_rows * ldb * 2};
    const uint32_t box[3] = {64, (uint32_t)l, (uint32_t)(3 * stacked)};
    int rc = make_map(&mX, b_planes, 3, dims, str, box);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(stream_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024)));
    attr_set = true;
  }
  const size_t smem = (size_t)a.stages * a.stage_bytes + 1024;
  const int n_units = a.n_tiles * a.n_splits * a.n_jb;
  const int grid = n_units < sm_count() ? n_units : sm_count();
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int rc = mma_events_begin(s, &e0, &e1);
  if (rc) return rc;
  stream_mma_kernel<<<grid, 192, smem, s>>>(mA[0], mA[1], mA[2], mX, a);
  N2V2R_CHECK_LAUNCH();
  return mma_events_end(s, e0, e1);
}

int stream_apply_tcgen05(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld,
                         const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate) {
  N2V2R_REQUIRE(xs && parts, "tcgen05 path needs its workspace");
  int rc = launch_split_transpose(s, x, plan.n, plan.l, plan.n_pad, xs);
  if (rc) return rc;
  const int64_t rows_pad = round_up(plan.n, 256);
  rc = launch_mma(s, planes, n_planes, plan.n, ld, (size_t)plan.n * ld * sizeof(__nv_bfloat16), xs, plan.l, plan.n_pad, plan.l, 1,
                  plan.n_pad, plan.k_chunk, parts, plan.l, rows_pad * plan.l, rows_pad, plan.l);
  if (rc) return rc;
  return launch_reduce_parts(s, parts, plan.n_splits, rows_pad, plan.n, plan.l, out, accumulate);
}

// Rectangular product for row slabs: out (rows_out x l) (= or +=) S x, S = n_planes planes of rows_out x k_len (leading dimension
// ld), x (k_len x l).  Same kernel, same K split; the workspace of the square plan (n >= rows_out, k_len) is large enough.
int stream_apply_rect(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld, int64_t rows_out,
                      int64_t k_len, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate) {
  N2V2R_REQUIRE(xs && parts && planes, "slab product needs its workspace");
  N2V2R_REQUIRE(rows_out > 0 && k_len > 0 && rows_out <= plan.n && k_len <= plan.n && ld >= k_len && ld % 64 == 0, "slab product: bad shape");
  const int64_t k_pad = round_up(k_len, 64);
  int rc = launch_split_transpose(s, x, k_len, plan.l, k_pad, xs);
  if (rc) return rc;
  const int64_t rows_pad = round_up(rows_out, 256);
  const int n_splits = (int)((k_pad + plan.k_chunk - 1) / plan.k_chunk);
  rc = launch_mma(s, planes, n_planes, rows_out, ld, (size_t)rows_out * ld * sizeof(__nv_bfloat16), xs, plan.l, k_pad, plan.l, 1, k_pad,
                  plan.k_chunk, parts, plan.l, rows_pad * plan.l, rows_pad, plan.l);
  if (rc) return rc;
  return launch_reduce_parts(s, parts, n_splits, rows_pad, rows_out, plan.l, out, accumulate);
}

// Seed batching: out_i (= or +=) A x_i for nb operands with ONE read of the layer.  xs: [nb][3][l][n_pad] bf16, parts:
// [n_splits][rows_pad][nb * l].  Every out_i has exactly the bits of the unbatched product (same K split, same MMA order per
// row; the 128- instead of 256-row tiles do not change a row's accumulation).
int stream_apply_batched(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld, int nb,
                         const float* const* x, __nv_bfloat16* xs, float* parts, float* const* out, int accumulate) {
  N2V2R_REQUIRE(xs && parts && nb >= 1 && nb <= 3, "batched product: bad arguments");
  const size_t per = (size_t)3 * plan.l * plan.n_pad;
  for (int i = 0; i < nb; ++i) {
    int rc = launch_split_transpose(s, x[i], plan.n, plan.l, plan.n_pad, xs + i * per);
    if (rc) return rc;
  }
  const int64_t rows_pad = round_up(plan.n, 256);
  const int64_t ldp = (int64_t)nb * plan.l;
  int rc = launch_mma(s, planes, n_planes, plan.n, ld, (size_t)plan.n * ld * sizeof(__nv_bfloat16), xs, plan.l, plan.n_pad, plan.l, 1,
                      plan.n_pad, plan.k_chunk, parts, ldp, rows_pad * ldp, rows_pad, ldp, 0, nb);
  if (rc) return rc;
  for (int i = 0; i < nb; ++i) {
    const int64_t total4 = plan.n * plan.l / 4;
    reduce_parts_strided_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, s>>>(parts, plan.n_splits, rows_pad * ldp, ldp, i * plan.l,
                                                                                plan.n, plan.l, out[i], accumulate);
    N2V2R_CHECK_LAUNCH();
  }
  return N2V2R_OK;
}

int launch_reduce_parts(cudaStream_t s, const float* parts, int n_splits, int64_t rows_pad, int64_t n, int l, float* out, int accumulate) {
  const int64_t total4 = n * l / 4;
  reduce_parts_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, s>>>(parts, n_splits, rows_pad, n, l, out, accumulate);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

}  // namespace n2v2r

using namespace n2v2r;

// K2: out = W^T W (on_rows=0) or W W^T (on_rows=1) -- preprocessing_utils.py:16-32
static const int kProjChunk = 512;     // K-split of the projection: short fp32 accumulation chains in TMEM, partials summed in order
static int g_proj_chunk = kProjChunk;  // debug knobs (n2v2r__debug_project): K chunk actually used (<= kProjChunk splits fit the workspace) ...
static int g_proj_nb = 2;              // ... and column blocks per A tile read

extern "C" void n2v2r__debug_project(int k_chunk, int nb) {
  if (k_chunk >= kProjChunk && k_chunk % 64 == 0) g_proj_chunk = k_chunk;
  if (nb >= 1 && nb <= 2) g_proj_nb = nb;
}

extern "C" size_t n2v2r_project_workspace_bytes(int64_t rows, int64_t cols, int on_rows) {
  const int64_t m = on_rows ? rows : cols, k = on_rows ? cols : rows;
  const int64_t k_pad = round_up(k, 64);
  const int64_t n_splits = (k_pad + kProjChunk - 1) / kProjChunk;
  return align_up(sizeof(__nv_bfloat16) * 3 * (size_t)m * k_pad) + align_up(sizeof(float) * (size_t)n_splits * round_up(m, 256) * m) + 256;
}

extern "C" int n2v2r_project(void* stream, const void* d_W, int dtype, int64_t rows, int64_t cols, int64_t ld, int on_rows,
                             float* d_out, int64_t ld_out, void* d_ws, size_t ws_bytes) {
  N2V2R_REQUIRE(d_W && d_out && d_ws && rows > 0 && cols > 0 && ld >= cols, "project: bad arguments");
  const int64_t m = on_rows ? rows : cols, k = on_rows ? cols : rows;
  N2V2R_REQUIRE(ld_out >= m, "project: ld_out < output size");
  if (ws_bytes < n2v2r_project_workspace_bytes(rows, cols, on_rows)) { set_error("project: workspace too small"); return N2V2R_ENOMEM; }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t k_pad = round_up(k, 64);
  const int n_splits = (int)((k_pad + g_proj_chunk - 1) / g_proj_chunk);
  const int64_t m_pad = round_up(m, 256);
  __nv_bfloat16* Ws = (__nv_bfloat16*)d_ws;
  float* parts = (float*)((char*)d_ws + align_up(sizeof(__nv_bfloat16) * 3 * (size_t)m * k_pad));
  dim3 grid((unsigned)((k_pad + 31) / 32), (unsigned)((m + 31) / 32));
  if (dtype == N2V2R_F32)
    split_planes_kernel<float><<<grid, 256, 0, s>>>((const float*)d_W, rows, cols, ld, on_rows ? 0 : 1, m, k_pad, Ws);
  else if (dtype == N2V2R_F64)
    split_planes_kernel<double><<<grid, 256, 0, s>>>((const double*)d_W, rows, cols, ld, on_rows ? 0 : 1, m, k_pad, Ws);
  else { set_error("project: unknown dtype %d", dtype); return N2V2R_EINVAL; }
  N2V2R_CHECK_LAUNCH();
  const int l = 64;
  const int n_jb = (int)((m + l - 1) / l);
  // SYRK: only units touching the upper triangle are computed; the mirror kernel fills the rest
  // two column blocks per read of the A tile (A k-slices kept in the collector): the unit's A traffic per output column halves
  int rc = launch_mma(s, Ws, 3, m, k_pad, (size_t)m * k_pad * sizeof(__nv_bfloat16), Ws, m, k_pad, l, n_jb, k_pad, g_proj_chunk,
                      parts, m, m_pad * m, m, m, 1, g_proj_nb, g_proj_nb > 1 ? 1 : 0);
  if (rc) return rc;
  const unsigned tiles = (unsigned)((m + 31) / 32);
  mirror_reduce_kernel<<<dim3(tiles, tiles), 256, 0, s>>>(parts, n_splits, m_pad * m, m, d_out, ld_out);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
