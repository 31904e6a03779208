// Host side of the sparse host->device transfer (n2v2r_host_pack_nonzeros, include/n2v2r_b200.h).
//
// The reference hands N2V2R dense float32 (n, n) graphs that are mostly zeros after network_transform
// (top_percent_keep = 10 leaves 10 %), and converts them to CSC on the host before the SVD
// (model.py:51-55).  The B200 path needs them dense on the device, but they need not cross PCIe dense:
// this packs a layer into a bitmap (1 bit per entry) plus the non-zero values in row order, bit-exact
// (an entry is "zero" only when all 32 bits are zero, so -0.0 and NaN payloads survive), with all the
// host threads the caller grants.  n2v2r_unpack_nonzeros (k0_pack.cu) rebuilds the dense layer in HBM.
//
// Plain C++ (g++): no CUDA in this file.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#include <immintrin.h>
#include "../../include/n2v2r_b200.h"

namespace n2v2r {
void set_error(const char* fmt, ...);
}

namespace {

constexpr int64_t kRowsPerClaim = 4;
constexpr int64_t kPrefetchAhead = 2048;      // software prefetch distance in entries (8 KB); N2V2R_PACK_PREFETCH overrides

// 32 consecutive entries -> bit j set when entry j is non-zero (any bit set)
inline uint32_t mask32_scalar(const uint32_t* p, int count) {
  uint32_t m = 0;
  for (int j = 0; j < count; ++j) m |= (uint32_t)(p[j] != 0u) << j;
  return m;
}

// compaction table: for a mask byte, the lane permutation that moves the set lanes to the front
struct CompactLut {
  alignas(32) uint32_t perm[256][8];
  CompactLut() {
    for (int m = 0; m < 256; ++m) {
      int k = 0;
      for (int j = 0; j < 8; ++j)
        if (m >> j & 1) perm[m][k++] = (uint32_t)j;
      for (; k < 8; ++k) perm[m][k] = 0;
    }
  }
};
const CompactLut& compact_lut() {
  static const CompactLut lut;
  return lut;
}

// one row: bitmap words + compacted values (branch-free: 8 entries per step, permute the non-zeros to the front, store
// all 8 lanes and advance by the popcount -- `vals` needs 8 entries of slack); returns the number of non-zeros
__attribute__((target("avx2,popcnt"))) int64_t pack_row_avx2(const uint32_t* row, int64_t cols, uint32_t* bits, uint32_t* vals,
                                                              const CompactLut& lut, int64_t ahead) {
  const int64_t full = cols / 32;
  const __m256i z = _mm256_setzero_si256();
  int64_t k = 0;
  for (int64_t w = 0; w < full; ++w) {
    const uint32_t* p = row + 32 * w;
    _mm_prefetch((const char*)(p + ahead), _MM_HINT_T0);          // two lines per word, a few KB ahead:
    _mm_prefetch((const char*)(p + ahead + 16), _MM_HINT_T0);     // more misses in flight per core
    uint32_t word = 0;
#pragma GCC unroll 4
    for (int q = 0; q < 4; ++q) {
      const __m256i v = _mm256_loadu_si256((const __m256i*)(p + 8 * q));
      const uint32_t m8 = (~(uint32_t)_mm256_movemask_ps(_mm256_castsi256_ps(_mm256_cmpeq_epi32(v, z)))) & 0xffu;
      const __m256i idx = _mm256_load_si256((const __m256i*)lut.perm[m8]);
      _mm256_storeu_si256((__m256i*)(vals + k), _mm256_permutevar8x32_epi32(v, idx));
      k += __builtin_popcount(m8);
      word |= m8 << (8 * q);
    }
    bits[w] = word;
  }
  if (cols % 32) {
    const uint32_t* p = row + 32 * full;
    uint32_t m = mask32_scalar(p, (int)(cols % 32));
    bits[full] = m;
    while (m) {
      vals[k++] = p[__builtin_ctz(m)];
      m &= m - 1;
    }
  }
  return k;
}

int64_t pack_row_plain(const uint32_t* row, int64_t cols, uint32_t* bits, uint32_t* vals) {
  const int64_t words = (cols + 31) / 32;
  int64_t k = 0;
  for (int64_t w = 0; w < words; ++w) {
    const uint32_t* p = row + 32 * w;
    const int count = (int)((cols - 32 * w) < 32 ? (cols - 32 * w) : 32);
    uint32_t m = mask32_scalar(p, count);
    bits[w] = m;
    while (m) {
      vals[k++] = p[__builtin_ctz(m)];
      m &= m - 1;
    }
  }
  return k;
}

}  // namespace

extern "C" int n2v2r_host_pack_nonzeros(const void* h_a, int64_t rows, int64_t cols, int64_t ld, uint32_t* h_bitmap,
                                        void* h_values, int64_t capacity, int64_t* h_row_start, int threads,
                                        int64_t* nnz) {
  if (!h_a || !h_bitmap || !h_values || !h_row_start || !nnz || rows < 0 || cols < 0 || ld < cols || capacity < 0) {
    n2v2r::set_error("n2v2r_host_pack_nonzeros: bad argument");
    return N2V2R_EINVAL;
  }
  *nnz = 0;
  if (rows == 0 || cols == 0) return N2V2R_OK;
  const bool avx2 = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("popcnt");
  const CompactLut& lut = compact_lut();
  const char* env = getenv("N2V2R_PACK_PREFETCH");
  const int64_t ahead = env ? atoll(env) : kPrefetchAhead;
  const int64_t words = (cols + 31) / 32;
  const uint32_t* a = (const uint32_t*)h_a;
  uint32_t* values = (uint32_t*)h_values;
  if (threads < 1) threads = 1;
  const int64_t claims = (rows + kRowsPerClaim - 1) / kRowsPerClaim;
  if ((int64_t)threads > claims) threads = (int)claims;

  std::atomic<int64_t> next_claim{0}, cursor{0};
  std::atomic<int> overflow{0};

  auto worker = [&]() {
    std::vector<uint32_t> tmp((size_t)(kRowsPerClaim * cols) + 8);
    int64_t counts[kRowsPerClaim];
    for (;;) {
      if (overflow.load(std::memory_order_relaxed)) return;
      const int64_t c = next_claim.fetch_add(1, std::memory_order_relaxed);
      if (c >= claims) return;
      const int64_t r0 = c * kRowsPerClaim;
      const int64_t r1 = r0 + kRowsPerClaim < rows ? r0 + kRowsPerClaim : rows;
      int64_t k = 0;
      for (int64_t r = r0; r < r1; ++r) {
        const int64_t got = avx2 ? pack_row_avx2(a + r * ld, cols, h_bitmap + r * words, tmp.data() + k, lut, ahead)
                                 : pack_row_plain(a + r * ld, cols, h_bitmap + r * words, tmp.data() + k);
        counts[r - r0] = got;
        k += got;
      }
      const int64_t pos = cursor.fetch_add(k, std::memory_order_relaxed);
      if (pos + k > capacity) {
        overflow.store(1, std::memory_order_relaxed);
        return;
      }
      memcpy(values + pos, tmp.data(), (size_t)k * sizeof(uint32_t));
      int64_t at = pos;
      for (int64_t r = r0; r < r1; ++r) {
        h_row_start[r] = at;
        at += counts[r - r0];
      }
    }
  };

  std::vector<std::thread> pool;
  pool.reserve((size_t)threads - 1);
  for (int t = 1; t < threads; ++t) pool.emplace_back(worker);
  worker();
  for (auto& th : pool) th.join();
  *nnz = overflow.load() ? -1 : cursor.load();
  return N2V2R_OK;
}
