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

}  // namespace n2v2r

using namespace n2v2r;

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
