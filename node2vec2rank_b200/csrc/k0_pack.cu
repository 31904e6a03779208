// K0c: device side of the sparse host->device transfer.  A layer packed on the host by
// n2v2r_host_pack_nonzeros (host_pack.cpp: bitmap + non-zero values in row order) is expanded into the dense
// (rows x ld) fp32 matrix the load pass (K1) reads.  HBM-bound on the dense write: 4 bytes per entry out,
// 1/8 byte + 4 bytes per non-zero in.
#include "common.cuh"

namespace n2v2r {

constexpr int kUnpackThreads = 256;
constexpr int kUnpackWarps = kUnpackThreads / 32;

// One CTA per row.  Pass 1: non-zeros per 32-word chunk (one warp per chunk); scan of the chunk counts;
// pass 2: every warp expands its chunks, lane j writing entry j of each word (128-byte coalesced stores).
__global__ void __launch_bounds__(kUnpackThreads) unpack_rows_kernel(const uint32_t* __restrict__ bitmap,
                                                                     const uint32_t* __restrict__ values,
                                                                     const int64_t* __restrict__ row_start, int64_t rows,
                                                                     int64_t cols, int64_t words, uint32_t* __restrict__ out,
                                                                     int64_t ld) {
  extern __shared__ uint32_t chunk_base[];          // chunks + 1
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunks = (int)((words + 31) / 32);
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
    const uint32_t* bits = bitmap + r * words;
    for (int c = warp; c < chunks; c += kUnpackWarps) {
      const int64_t w = (int64_t)c * 32 + lane;
      uint32_t cnt = w < words ? __popc(bits[w]) : 0;
      cnt = __reduce_add_sync(0xffffffffu, cnt);
      if (lane == 0) chunk_base[c + 1] = cnt;
    }
    if (threadIdx.x == 0) chunk_base[0] = 0;
    __syncthreads();
    if (warp == 0) {                                 // inclusive scan of chunk_base[1..chunks], 32 at a time
      uint32_t carry = 0;
      for (int c0 = 0; c0 < chunks; c0 += 32) {
        const int c = c0 + lane;
        uint32_t v = c < chunks ? chunk_base[c + 1] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t up = __shfl_up_sync(0xffffffffu, v, d);
          if (lane >= d) v += up;
        }
        if (c < chunks) chunk_base[c + 1] = v + carry;
        carry += __shfl_sync(0xffffffffu, v, 31);
      }
    }
    __syncthreads();
    const uint32_t* vals = values + row_start[r];
    uint32_t* orow = out + r * ld;
    for (int c = warp; c < chunks; c += kUnpackWarps) {
      const int64_t w = (int64_t)c * 32 + lane;
      const uint32_t word = w < words ? bits[w] : 0;
      uint32_t off = __popc(word);                   // exclusive scan over the lanes
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, off, d);
        if (lane >= d) off += up;
      }
      off = off - __popc(word) + chunk_base[c];
      const int in_chunk = (int)min((int64_t)32, words - (int64_t)c * 32);
      for (int j = 0; j < in_chunk; ++j) {
        const uint32_t wj = __shfl_sync(0xffffffffu, word, j);
        const uint32_t oj = __shfl_sync(0xffffffffu, off, j);
        const int64_t col = ((int64_t)c * 32 + j) * 32 + lane;
        uint32_t v = 0;
        if ((wj >> lane) & 1u) v = vals[oj + __popc(wj & ((1u << lane) - 1u))];
        if (col < cols) orow[col] = v;
      }
    }
    __syncthreads();                                 // chunk_base is reused by the next row
  }
}

// ---- K0d: reindex a parsed adjacency block into the common node order (device side of match_networks) ----
// out[i][j] = in[ri[i]][ci[j]]  (transpose: out[i][j] = in[ci[j]][ri[i]], i.e. graph_pd.T then .loc[rows, cols]).
// 32 x 32 tiles through shared memory so that the transposed read stays coalesced along rows of `in`.
__global__ void __launch_bounds__(256) gather_submatrix_kernel(const double* __restrict__ in, int64_t ld_in,
                                                               const int32_t* __restrict__ ri, const int32_t* __restrict__ ci,
                                                               int64_t out_rows, int64_t out_cols, int transpose,
                                                               double* __restrict__ out, int64_t ld_out) {
  __shared__ double t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t i0 = (int64_t)blockIdx.y * 32, j0 = (int64_t)blockIdx.x * 32;
  if (!transpose) {
    const int64_t j = j0 + tx;
    const int64_t sc = j < out_cols ? ci[j] : -1;
    for (int r = ty; r < 32; r += 8) {
      const int64_t i = i0 + r;
      if (i < out_rows && sc >= 0) out[i * ld_out + j] = in[(int64_t)ri[i] * ld_in + sc];
    }
    return;
  }
  // transposed: read rows ci[j] of `in` along i (columns ri[i]), write rows i of `out` along j
  const int64_t ii = i0 + tx;
  const int64_t scol = ii < out_rows ? ri[ii] : -1;
  for (int r = ty; r < 32; r += 8) {
    const int64_t j = j0 + r;
    t[r][tx] = (j < out_cols && scol >= 0) ? in[(int64_t)ci[j] * ld_in + scol] : 0.0;     // t[j_local][i_local]
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = i0 + r, j = j0 + tx;
    if (i < out_rows && j < out_cols) out[i * ld_out + j] = t[tx][r];
  }
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" int n2v2r_gather_submatrix(void* stream, const double* d_in, int64_t in_rows, int64_t in_cols, int64_t ld_in,
                                      const int32_t* d_row_idx, const int32_t* d_col_idx, int64_t out_rows, int64_t out_cols,
                                      int transpose, double* d_out, int64_t ld_out) {
  N2V2R_REQUIRE(d_in && d_out && d_row_idx && d_col_idx, "gather_submatrix: null pointer");
  N2V2R_REQUIRE(in_rows > 0 && in_cols > 0 && ld_in >= in_cols && out_rows >= 0 && out_cols >= 0 && ld_out >= out_cols,
                "gather_submatrix: bad shape");
  if (out_rows == 0 || out_cols == 0) return N2V2R_OK;
  dim3 grid((unsigned)((out_cols + 31) / 32), (unsigned)((out_rows + 31) / 32));
  gather_submatrix_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, ld_in, d_row_idx, d_col_idx, out_rows, out_cols, transpose,
                                                                 d_out, ld_out);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_unpack_nonzeros(void* stream, const uint32_t* d_bitmap, const void* d_values, const int64_t* d_row_start,
                                     int64_t rows, int64_t cols, void* d_out, int64_t ld) {
  N2V2R_REQUIRE(rows >= 0 && cols >= 0 && ld >= cols, "n2v2r_unpack_nonzeros: bad shape (rows=%lld cols=%lld ld=%lld)",
                (long long)rows, (long long)cols, (long long)ld);
  if (rows == 0 || cols == 0) return N2V2R_OK;
  N2V2R_REQUIRE(d_bitmap && d_values && d_row_start && d_out, "n2v2r_unpack_nonzeros: null pointer");
  const int64_t words = (cols + 31) / 32;
  const int64_t chunks = (words + 31) / 32;
  const size_t smem = (size_t)(chunks + 1) * sizeof(uint32_t);
  N2V2R_REQUIRE(smem <= 48 * 1024, "n2v2r_unpack_nonzeros: more than 12 million columns are not supported");
  const int64_t grid = rows < (int64_t)sm_count() * 64 ? rows : (int64_t)sm_count() * 64;
  unpack_rows_kernel<<<(unsigned)grid, kUnpackThreads, smem, (cudaStream_t)stream>>>(
      d_bitmap, (const uint32_t*)d_values, d_row_start, rows, cols, words, (uint32_t*)d_out, ld);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
