// K9 distances, K10 segmented radix-sort rank, K11 Borda.
// Replaces model.py:57-96 (__rank), model_utils.py:39-67 (compute_pairwise_distances),
// model.py:167-185 + model_utils.py:22-36 (descending sort + Borda points).
#include "common.cuh"
#include <algorithm>
#include <math.h>
#include <mutex>

namespace n2v2r {

constexpr int kMaxK = 96;

// One thread per node.  `one` = reference layer (or mean of layers), `two` = layer i.
template <int KMAX, class E>
__global__ void __launch_bounds__(128) distances_kernel(const E* __restrict__ emb, int T, int64_t n, int k,
                                                        const int* __restrict__ col_dim, const int* __restrict__ col_metric,
                                                        int n_cols, int strategy, double* __restrict__ scores) {
  const int64_t node = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (node >= n) return;
  const int comp = blockIdx.y;
  const int i = (strategy == N2V2R_ONE_VS_REST) ? comp : comp + 1;      // model.py:59-66
  double one[KMAX], two[KMAX];
#pragma unroll 1
  for (int c = 0; c < k; ++c) {
    two[c] = (double)emb[((int64_t)i * n + node) * k + c];
    double acc = 0.0;
    int cnt = 0;
    if (strategy == N2V2R_SEQUENTIAL) {                                   // model.py:74-76
      acc = (double)emb[((int64_t)(i - 1) * n + node) * k + c];
      cnt = 1;
    } else {
      for (int t = 0; t < T; ++t) {                                       // np.mean over axis 0, model.py:78-84
        const bool use = (strategy == N2V2R_ONE_VS_BEFORE) ? (t < i) : (t != i);
        if (use) { acc += (double)emb[((int64_t)t * n + node) * k + c]; ++cnt; }
      }
    }
    one[c] = (strategy == N2V2R_SEQUENTIAL) ? acc : acc / (double)cnt;
  }
  for (int c = 0; c < n_cols; ++c) {
    const int d = col_dim[c], m = col_metric[c];
    double r;
    if (m == N2V2R_EUCLIDEAN) {                                           // ||u - v||_2
      double s = 0.0;
      for (int q = 0; q < d; ++q) { const double t = one[q] - two[q]; s += t * t; }
      r = sqrt(s);
    } else {
      double mu = 0.0, mv = 0.0;
      if (m == N2V2R_CORRELATION) {                                       // centred cosine
        for (int q = 0; q < d; ++q) { mu += one[q]; mv += two[q]; }
        mu /= (double)d; mv /= (double)d;
      }
      double uv = 0.0, uu = 0.0, vv = 0.0;
      for (int q = 0; q < d; ++q) {
        const double a = one[q] - mu, b = two[q] - mv;
        uv += a * b; uu += a * a; vv += b * b;
      }
      r = 1.0 - uv / sqrt(uu * vv);                                       // NaN for a zero row, like scipy 1.18
      if (r < 0.0) r = 0.0;                                               // np.clip(dist, 0, 2) keeps NaN
      if (r > 2.0) r = 2.0;
    }
    scores[((int64_t)comp * n_cols + c) * n + node] = r;
  }
}

// ascending radix key for "descending score, NaN last"
__device__ __forceinline__ uint64_t rank_key(double s) {
  if (s != s) return ~0ull;
  if (s == 0.0) s = 0.0;                 // -0.0 ties with +0.0
  uint64_t b = (uint64_t)__double_as_longlong(s);
  b = (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);   // ascending order key
  return ~b;                             // descending; max finite key < ~0
}

// One CTA (1024 threads) sorts one column: LSD radix, 8-bit digits, stable.
__global__ void __launch_bounds__(1024) rank_sort_kernel(const double* __restrict__ scores, int64_t n_, int n_cols, int cols_per_group,
                                                         uint64_t* __restrict__ keybuf, uint32_t* __restrict__ idxbuf,
                                                         int32_t* __restrict__ positions,
                                                         unsigned long long* __restrict__ borda) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t base[256];
  __shared__ uint32_t warp_cnt[32][256];
  const int n = (int)n_;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (int col = blockIdx.x; col < n_cols; col += gridDim.x) {
    uint64_t* kb[2] = {keybuf + (size_t)col * 2 * n, keybuf + (size_t)col * 2 * n + n};
    uint32_t* ib[2] = {idxbuf + (size_t)col * 2 * n, idxbuf + (size_t)col * 2 * n + n};
    const double* sc = scores + (size_t)col * n;
    for (int i = tid; i < n; i += 1024) { kb[0][i] = rank_key(sc[i]); ib[0][i] = (uint32_t)i; }
    int cur = 0;
    for (int pass = 0; pass < 8; ++pass) {
      const int shift = pass * 8;
      if (tid < 256) hist[tid] = 0;
      __syncthreads();
      const uint64_t* kin = kb[cur];
      const uint32_t* iin = ib[cur];
      for (int i = tid; i < n; i += 1024) atomicAdd(&hist[(kin[i] >> shift) & 255u], 1u);
      __syncthreads();
      const int trivial = __syncthreads_or(tid < 256 && hist[tid] == (uint32_t)n);
      if (trivial) continue;               // every key has the same digit: nothing to move
      if (w == 0) {                        // exclusive scan of 256 bins by one warp (8 per lane)
        uint32_t v[8], s = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) { v[q] = hist[lane * 8 + q]; s += v[q]; }
        uint32_t inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        uint32_t run = inc - s;
#pragma unroll
        for (int q = 0; q < 8; ++q) { base[lane * 8 + q] = run; run += v[q]; }
      }
      __syncthreads();
      uint64_t* kout = kb[cur ^ 1];
      uint32_t* iout = ib[cur ^ 1];
      for (int c0 = 0; c0 < n; c0 += 1024) {
        const int i = c0 + tid;
        const bool valid = i < n;
        uint64_t key = 0; uint32_t id = 0; uint32_t d = 256u + (uint32_t)lane;   // invalid lanes never match each other
        if (valid) { key = kin[i]; id = iin[i]; d = (uint32_t)((key >> shift) & 255u); }
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
#pragma unroll
        for (int q = 0; q < 8; ++q) warp_cnt[w][lane + 32 * q] = 0;
        __syncwarp();
        if (valid && rank == 0) warp_cnt[w][d] = __popc(peers);
        __syncthreads();
        uint32_t run = 0;
        if (tid < 256) {
#pragma unroll 8
          for (int ww = 0; ww < 32; ++ww) { const uint32_t t = warp_cnt[ww][tid]; warp_cnt[ww][tid] = run; run += t; }
        }
        __syncthreads();
        if (valid) {
          const uint32_t dst = base[d] + warp_cnt[w][d] + rank;
          kout[dst] = key; iout[dst] = id;
        }
        __syncthreads();
        if (tid < 256) base[tid] += run;
      }
      __syncthreads();
      cur ^= 1;
    }
    const uint32_t* sorted = ib[cur];
    for (int r = tid; r < n; r += 1024) {
      const uint32_t id = sorted[r];
      if (positions) positions[(size_t)col * n + id] = r;
      atomicAdd(&borda[(size_t)(col / cols_per_group) * n + id], (unsigned long long)(n - r));   // model_utils.py:25: num_candidates - position
    }
    __syncthreads();
  }
}

// ---- K12: binary-relevance DCG / nDCG of many rankings (post_utils.py:45-128) ------------------------
// One CTA per ranking.  dcg = sum_pos rel[item(pos)] / log2(pos + 2); the ideal ranking puts the m relevant items
// that occur in the ranking first: idcg = sum_{pos < m} 1 / log2(pos + 2); ndcg = dcg / idcg, 0 when m == 0
// (post_utils.py:82-83, :113-114).  fp64, block-wide tree sums in a fixed order (deterministic).
__global__ void __launch_bounds__(256) ndcg_kernel(const int32_t* __restrict__ rankings, int64_t n_rankings, int64_t len,
                                                    const uint8_t* __restrict__ relevant, int64_t n_items,
                                                    double* __restrict__ ndcg, double* __restrict__ dcg_out) {
  __shared__ double sd[256];
  __shared__ long long sc[256];
  for (int64_t r = blockIdx.x; r < n_rankings; r += gridDim.x) {
    const int32_t* row = rankings + r * len;
    double d = 0.0;
    long long m = 0;
    for (int64_t pos = threadIdx.x; pos < len; pos += 256) {
      const int32_t item = row[pos];
      if (item >= 0 && item < n_items && relevant[item]) {
        d += 1.0 / log2((double)(pos + 2));
        ++m;
      }
    }
    sd[threadIdx.x] = d; sc[threadIdx.x] = m;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) { sd[threadIdx.x] += sd[threadIdx.x + o]; sc[threadIdx.x] += sc[threadIdx.x + o]; }
      __syncthreads();
    }
    const double dcg = sd[0];
    const long long mm = sc[0];
    __syncthreads();
    double id = 0.0;
    for (int64_t pos = threadIdx.x; pos < mm; pos += 256) id += 1.0 / log2((double)(pos + 2));
    sd[threadIdx.x] = id;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) sd[threadIdx.x] += sd[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      ndcg[r] = mm > 0 ? dcg / sd[0] : 0.0;
      if (dcg_out) dcg_out[r] = dcg;
    }
    __syncthreads();
  }
}

// ---- signed ranks: out = prior > 0 ? score : -score, per node, for every column (model_utils.py:7-19) ----
template <class T>
__global__ void __launch_bounds__(256) signed_scores_kernel(const T* __restrict__ scores, int64_t n, int64_t total,
                                                            const double* __restrict__ prior, T* __restrict__ out) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const T v = scores[e];
    out[e] = prior[e % n] > 0.0 ? v : -v;          // zero, negative and NaN priors take the negative branch, like the reference
  }
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" int n2v2r_signed_scores(void* stream, const void* d_scores, int is_int64, int64_t n, int n_cols, const double* d_prior,
                                   void* d_out) {
  N2V2R_REQUIRE(d_scores && d_prior && d_out, "signed_scores: null pointer");
  N2V2R_REQUIRE(n > 0 && n_cols >= 0, "signed_scores: bad shape");
  if (n_cols == 0) return N2V2R_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t total = n * n_cols;
  const unsigned grid = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)sm_count() * 16);
  if (is_int64) signed_scores_kernel<long long><<<grid, 256, 0, s>>>((const long long*)d_scores, n, total, d_prior, (long long*)d_out);
  else signed_scores_kernel<double><<<grid, 256, 0, s>>>((const double*)d_scores, n, total, d_prior, (double*)d_out);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_distances(void* stream, const void* d_embed, int dtype, int T, int64_t n, int k,
                               const int32_t* h_col_dim, const int32_t* h_col_metric, int n_cols,
                               int strategy, double* d_scores) {
  N2V2R_REQUIRE(d_embed && d_scores && h_col_dim && h_col_metric, "distances: null pointer");
  N2V2R_REQUIRE(T >= 1 && n > 0 && k >= 1 && k <= kMaxK && n_cols >= 1, "distances: bad shape (k <= %d)", kMaxK);
  N2V2R_REQUIRE(strategy >= 0 && strategy <= 2, "distances: unknown strategy %d", strategy);
  N2V2R_REQUIRE(dtype == N2V2R_F32 || dtype == N2V2R_F64, "distances: unknown dtype %d", dtype);
  for (int c = 0; c < n_cols; ++c) {
    N2V2R_REQUIRE(h_col_dim[c] >= 1 && h_col_dim[c] <= k, "distances: column %d dim %d outside [1,%d]", c, h_col_dim[c], k);
    N2V2R_REQUIRE(h_col_metric[c] >= 0 && h_col_metric[c] <= 2, "distances: unsupported metric code %d", h_col_metric[c]);
  }
  const int n_comp = (strategy == N2V2R_ONE_VS_REST) ? T : T - 1;
  if (n_comp <= 0) return N2V2R_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int* d_meta = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_meta, sizeof(int) * 2 * n_cols, s));
  N2V2R_CUDA(cudaMemcpyAsync(d_meta, h_col_dim, sizeof(int) * n_cols, cudaMemcpyHostToDevice, s));
  N2V2R_CUDA(cudaMemcpyAsync(d_meta + n_cols, h_col_metric, sizeof(int) * n_cols, cudaMemcpyHostToDevice, s));
  dim3 grid((unsigned)((n + 127) / 128), (unsigned)n_comp);
  const int* cd = d_meta; const int* cm = d_meta + n_cols;
  if (dtype == N2V2R_F32) {
    const float* e = (const float*)d_embed;
    if (k <= 32) distances_kernel<32, float><<<grid, 128, 0, s>>>(e, T, n, k, cd, cm, n_cols, strategy, d_scores);
    else distances_kernel<kMaxK, float><<<grid, 128, 0, s>>>(e, T, n, k, cd, cm, n_cols, strategy, d_scores);
  } else {
    const double* e = (const double*)d_embed;
    if (k <= 32) distances_kernel<32, double><<<grid, 128, 0, s>>>(e, T, n, k, cd, cm, n_cols, strategy, d_scores);
    else distances_kernel<kMaxK, double><<<grid, 128, 0, s>>>(e, T, n, k, cd, cm, n_cols, strategy, d_scores);
  }
  N2V2R_CHECK_LAUNCH();
  N2V2R_CUDA(cudaFreeAsync(d_meta, s));
  return N2V2R_OK;
}

extern "C" size_t n2v2r_rank_workspace_bytes(int64_t n, int n_cols) {
  return align_up((size_t)n_cols * 2 * n * sizeof(uint64_t)) + align_up((size_t)n_cols * 2 * n * sizeof(uint32_t));
}

extern "C" int n2v2r_rank_borda(void* stream, const double* d_scores, int64_t n, int n_cols, int n_groups,
                                int32_t* d_positions, int64_t* d_borda, void* d_ws, size_t ws_bytes) {
  N2V2R_REQUIRE(d_scores && d_borda && d_ws, "rank_borda: null pointer");
  N2V2R_REQUIRE(n > 0 && n < (1ll << 31) && n_cols >= 1 && n_groups >= 1 && n_cols % n_groups == 0,
                "rank_borda: bad shape (n_cols must be a multiple of n_groups)");
  if (ws_bytes < n2v2r_rank_workspace_bytes(n, n_cols)) { set_error("rank_borda: workspace too small"); return N2V2R_ENOMEM; }
  cudaStream_t s = (cudaStream_t)stream;
  Arena a(d_ws, ws_bytes);
  uint64_t* keys = a.take<uint64_t>((size_t)n_cols * 2 * n);
  uint32_t* idx = a.take<uint32_t>((size_t)n_cols * 2 * n);
  N2V2R_CUDA(cudaMemsetAsync(d_borda, 0, sizeof(int64_t) * n * n_groups, s));
  const int grid = n_cols < sm_count() * 2 ? n_cols : sm_count() * 2;
  rank_sort_kernel<<<grid, 1024, 0, s>>>(d_scores, n, n_cols, n_cols / n_groups, keys, idx, d_positions, (unsigned long long*)d_borda);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

// ---- rank + Borda without any allocation by anyone (no counterpart in the reference) ----
// The sort buffers of n2v2r_rank_borda (two uint64 key arrays and two uint32 index arrays per column) live in a file-scope device
// array instead of a caller workspace, so a caller that cannot allocate device memory (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc) only passes its inputs and outputs.  Every GPU that loads the library holds the array
// (24 MiB) until the process exits.  All calls share it: two calls that overlap on the GPU (different streams, different
// threads) race.  Nothing in it has to start at zero -- rank_sort_kernel writes every key / index word before it reads it -- so
// the contents a previous call left behind are harmless; the only thing cleared is the caller's Borda table.
constexpr size_t kRankNoAllocMaxElems = (size_t)N2V2R_RANK_NOALLOC_MAX_ELEMS;                  // n * n_cols
constexpr size_t kRankNoAllocKeyBytes = kRankNoAllocMaxElems * 2 * sizeof(uint64_t);
constexpr size_t kRankNoAllocBytes = kRankNoAllocKeyBytes + kRankNoAllocMaxElems * 2 * sizeof(uint32_t);
static __device__ __align__(256) unsigned char g_rank_scratch[kRankNoAllocBytes];

// In host code `g_rank_scratch` is only a handle of the symbol: its device address is looked up once per device (with that
// device current) and cached.
static int rank_scratch_address(unsigned char** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static unsigned char* addr[kMaxDevices] = {};
  int dev = -1;
  N2V2R_CUDA(cudaGetDevice(&dev));
  N2V2R_REQUIRE(dev >= 0 && dev < kMaxDevices, "rank_borda_noalloc: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    N2V2R_CUDA(cudaGetSymbolAddress(&p, g_rank_scratch));
    N2V2R_REQUIRE(p != nullptr, "rank_borda_noalloc: the scratch symbol has no device address");
    addr[dev] = (unsigned char*)p;
  }
  *out = addr[dev];
  return N2V2R_OK;
}

extern "C" int n2v2r_rank_borda_noalloc(void* stream, const double* d_scores, int64_t n, int n_cols, int n_groups,
                                        int32_t* d_positions, int64_t* d_borda) {
  N2V2R_REQUIRE(d_scores && d_borda, "rank_borda_noalloc: null pointer");
  N2V2R_REQUIRE(n > 0 && n < (1ll << 31) && n_cols >= 1 && n_groups >= 1 && n_cols % n_groups == 0,
                "rank_borda_noalloc: bad shape (n_cols must be a multiple of n_groups)");
  N2V2R_REQUIRE((size_t)n <= kRankNoAllocMaxElems && (size_t)n * (size_t)n_cols <= kRankNoAllocMaxElems,
                "rank_borda_noalloc: n * n_cols = %lld exceeds the built-in scratch of %lld score entries (N2V2R_RANK_NOALLOC_MAX_ELEMS); "
                "use n2v2r_rank_borda with a caller workspace", (long long)n * n_cols, (long long)kRankNoAllocMaxElems);
  unsigned char* scratch = nullptr;
  int rc = rank_scratch_address(&scratch);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  uint64_t* keys = (uint64_t*)scratch;
  uint32_t* idx = (uint32_t*)(scratch + kRankNoAllocKeyBytes);
  N2V2R_CUDA(cudaMemsetAsync(d_borda, 0, sizeof(int64_t) * n * n_groups, s));     // the caller's output; the scratch needs no clearing
  const int grid = n_cols < sm_count() * 2 ? n_cols : sm_count() * 2;
  rank_sort_kernel<<<grid, 1024, 0, s>>>(d_scores, n, n_cols, n_cols / n_groups, keys, idx, d_positions, (unsigned long long*)d_borda);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_ndcg(void* stream, const int32_t* d_rankings, int64_t n_rankings, int64_t len,
                          const uint8_t* d_relevant, int64_t n_items, double* d_ndcg, double* d_dcg) {
  N2V2R_REQUIRE(d_rankings && d_relevant && d_ndcg, "ndcg: null pointer");
  N2V2R_REQUIRE(n_rankings >= 0 && len >= 0 && n_items >= 0 && n_items < (1ll << 31), "ndcg: bad shape");
  if (n_rankings == 0) return N2V2R_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = (int)(n_rankings < (int64_t)sm_count() * 8 ? n_rankings : (int64_t)sm_count() * 8);
  ndcg_kernel<<<grid, 256, 0, s>>>(d_rankings, n_rankings, len, d_relevant, n_items, d_ndcg, d_dcg);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
