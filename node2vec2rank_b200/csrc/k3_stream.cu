// K3/K4: streaming skinny products  Y (+)= A X  and  Z = A^T X  over bf16-plane layers.
// Replaces the ARPACK operator matvec/rmatvec (scipy _svds.py:442-457) reached from
// spectral_embedding.UASE at model.py:54-55.
//
// Two implementations behind one entry (stream_apply):
//   * N2V2R_ALGO_TCGEN05 (default): TMA-fed tcgen05.mma kernel, k3_tcgen05.cu;
//   * N2V2R_ALGO_SIMT: plain fp32-FMA tile kernel below, used for cross-checking the MMA
//     path on the GPU and for shapes the MMA kernel does not take.
// A^T products use the transposed plane copy when the layer is not symmetric.
#include "k5_orth.cuh"

namespace n2v2r {

int stream_apply_tcgen05(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld,
                         const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate);  // k3_tcgen05.cu
int stream_apply_rect(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld, int64_t rows_out,
                      int64_t k_len, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate);   // k3_tcgen05.cu
int stream_apply_e4m3(cudaStream_t s, const StreamPlan& plan, const void* plane8, int64_t ld8, const float* x, void* xs_ws,
                      float* parts, float* out, int accumulate);                                        // k3_fp8.cu
int stream_apply_tiles(cudaStream_t s, const StreamPlan& plan, const uint32_t* tile_ptr, const void* entries, int n_planes,
                       int tile_rows, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate);   // k3_sparse.cu
int stream_apply_rows(cudaStream_t s, int64_t n, int l, const uint32_t* list_ptr, const void* lists, int rb, int W, int list_cap,
                      const float* x, float* out, int accumulate);                                        // k3_rows.cu

template <int LC>
__global__ void __launch_bounds__(256) simt_apply_kernel(const __nv_bfloat16* __restrict__ planes, int n_planes, int64_t n,
                                                         int64_t ld, const float* __restrict__ X, float* __restrict__ out,
                                                         int accumulate) {
  constexpr int L = LC * 16;
  __shared__ float As[32][65];
  __shared__ float Xs[32][L];
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  const int64_t m0 = (int64_t)blockIdx.x * 64;
  const size_t pstride = (size_t)n * ld;
  float acc[4][LC];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int c = 0; c < LC; ++c) acc[i][c] = 0.f;
  const int lr = tid >> 5, lc = tid & 31;
  for (int64_t k0 = 0; k0 < n; k0 += 32) {
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const int m = lr + 8 * rr;
      float v = 0.f;
      if (m0 + m < n && k0 + lc < ld) {
        const size_t off = (size_t)(m0 + m) * ld + k0 + lc;
        for (int p = n_planes - 1; p >= 0; --p) v += __bfloat162float(planes[p * pstride + off]);
      }
      As[lc][m] = v;
    }
    for (int e = tid; e < 32 * L; e += 256) {
      const int kk = e / L, c = e - kk * L;
      Xs[kk][c] = (k0 + kk < n) ? X[(k0 + kk) * L + c] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < 32; ++kk) {
      float a[4], x[LC];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int c = 0; c < LC; ++c) x[c] = Xs[kk][tx + 16 * c];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < LC; ++c) acc[i][c] = fmaf(a[i], x[c], acc[i][c]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m < n) {
#pragma unroll
      for (int c = 0; c < LC; ++c) {
        float* o = out + m * L + tx + 16 * c;
        *o = accumulate ? (*o + acc[i][c]) : acc[i][c];
      }
    }
  }
}

int make_stream_plan(int64_t n, int l, int algo, int k_chunk, StreamPlan* plan) {
  N2V2R_REQUIRE(n > 0 && l >= 16 && l <= kMaxL && l % 16 == 0, "block width l must be 16, 32, 48, 64 or 80 (got %d)", l);
  plan->n = n;
  plan->n_pad = round_up(n, 64);
  plan->l = l;
  plan->algo = algo;
  plan->k_chunk_user = k_chunk > 0 ? (int)round_up(k_chunk, 64) : 0;
  if (k_chunk <= 0) k_chunk = 4096;
  k_chunk = (int)round_up(k_chunk, 64);
  plan->k_chunk = k_chunk;
  plan->n_splits = (int)((plan->n_pad + k_chunk - 1) / k_chunk);
  if (algo == N2V2R_ALGO_TCGEN05) {
    // + room for the e4m3 operand (K padded to 128, column maxima and scales) and one more K split (k3_fp8.cu)
    plan->xs_bytes = align_up(sizeof(__nv_bfloat16) * 3 * (size_t)l * plan->n_pad) + 65536;
    plan->part_bytes = align_up(sizeof(float) * (size_t)(plan->n_splits + 1) * round_up(n, 256) * l);
  } else {
    plan->xs_bytes = 0;
    plan->part_bytes = 0;
  }
  return N2V2R_OK;
}

int stream_apply_slab(cudaStream_t s, const StreamPlan& plan, const n2v2r_layer_t& L, int transpose, const float* x,
                      __nv_bfloat16* xs, float* parts, float* out, int accumulate, double* bytes_moved) {
  N2V2R_REQUIRE(L.storage == N2V2R_STORE_PLANES && plan.algo == N2V2R_ALGO_TCGEN05, "row slabs: bf16 planes on the tcgen05 path only");
  N2V2R_REQUIRE(L.rows > 0 && L.row_offset >= 0 && L.row_offset + L.rows <= plan.n && L.d_planes && L.d_planes_t && L.ld_t >= L.rows,
                "row slab: need rows inside the layer, the slab and its transpose");
  if (bytes_moved) *bytes_moved += (double)L.rows * (double)plan.n * 2.0 * L.n_planes + 4.0 * plan.l * (double)(plan.n + L.rows);
  if (transpose) return stream_apply_rect(s, plan, L.d_planes_t, L.n_planes, L.ld_t, plan.n, L.rows, x, xs, parts, out, accumulate);
  return stream_apply_rect(s, plan, L.d_planes, L.n_planes, L.ld, L.rows, plan.n, x, xs, parts, out, accumulate);
}

int stream_apply(cudaStream_t s, const StreamPlan& plan, const n2v2r_layer_t& L, int transpose, const float* x,
                 __nv_bfloat16* xs, float* parts, float* out, int accumulate, double* bytes_moved) {
  N2V2R_REQUIRE(L.rows == 0, "row-slab layers go through stream_apply_slab");
  if (L.storage == N2V2R_STORE_TILES) {
    const uint32_t* tp = (transpose && L.d_tile_ptr_t) ? L.d_tile_ptr_t : L.d_tile_ptr;
    const void* en = (transpose && L.d_entries_t) ? L.d_entries_t : L.d_entries;
    N2V2R_REQUIRE(plan.algo == N2V2R_ALGO_TCGEN05, "compressed-tile layers run on the tcgen05 path only");
    if (bytes_moved)   // the entry lists once + tile offsets + skinny operand read + skinny result written
      *bytes_moved += 8.0 * (double)((transpose && L.d_entries_t) ? L.n_entries_t : L.n_entries) +
                      4.0 * (double)((plan.n + L.tile_rows - 1) / L.tile_rows) * (double)(plan.n_pad / 64) + 4.0 * plan.l * 2.0 * (double)plan.n;
    return stream_apply_tiles(s, plan, tp, en, L.n_planes, L.tile_rows, x, xs, parts, out, accumulate);
  }
  if (L.storage == N2V2R_STORE_ROWS) {
    const bool tr = transpose && L.d_tile_ptr_t;
    if (bytes_moved)   // the lists once + list offsets + skinny operand read + skinny result written
      *bytes_moved += 8.0 * (double)(tr ? L.n_entries_t : L.n_entries) + 4.0 * plan.l * 2.0 * (double)plan.n;
    return stream_apply_rows(s, plan.n, plan.l, tr ? L.d_tile_ptr_t : L.d_tile_ptr, tr ? L.d_entries_t : L.d_entries, L.tile_rows,
                             L.panel_cols, tr ? L.list_cap_t : L.list_cap, x, out, accumulate);
  }
  if (L.storage == N2V2R_STORE_E4M3) {
    const void* p8 = (transpose && L.d_planes_t) ? L.d_planes_t : L.d_planes;
    N2V2R_REQUIRE(plan.algo == N2V2R_ALGO_TCGEN05, "e4m3 layers run on the tcgen05 path only");
    if (bytes_moved) *bytes_moved += (double)plan.n * (double)plan.n + 4.0 * plan.l * 2.0 * (double)plan.n;
    return stream_apply_e4m3(s, plan, p8, L.ld, x, xs, parts, out, accumulate);
  }
  const void* planes = (transpose && L.d_planes_t) ? L.d_planes_t : L.d_planes;
  N2V2R_REQUIRE(planes && L.n_planes >= 1 && L.n_planes <= 3 && L.ld >= plan.n && L.ld % 64 == 0,
                "layer: need 1..3 planes and ld a multiple of 64 >= n");
  if (bytes_moved)   // matrix read once + skinny operand read + skinny result written
    *bytes_moved += (double)plan.n * (double)plan.n * 2.0 * L.n_planes + 4.0 * plan.l * 2.0 * (double)plan.n;
  if (plan.algo == N2V2R_ALGO_TCGEN05)
    return stream_apply_tcgen05(s, plan, planes, L.n_planes, L.ld, x, xs, parts, out, accumulate);
  const unsigned grid = (unsigned)((plan.n + 63) / 64);
  const __nv_bfloat16* P = (const __nv_bfloat16*)planes;
  switch (plan.l / 16) {
    case 1: simt_apply_kernel<1><<<grid, 256, 0, s>>>(P, L.n_planes, plan.n, L.ld, x, out, accumulate); break;
    case 2: simt_apply_kernel<2><<<grid, 256, 0, s>>>(P, L.n_planes, plan.n, L.ld, x, out, accumulate); break;
    case 3: simt_apply_kernel<3><<<grid, 256, 0, s>>>(P, L.n_planes, plan.n, L.ld, x, out, accumulate); break;
    case 4: simt_apply_kernel<4><<<grid, 256, 0, s>>>(P, L.n_planes, plan.n, L.ld, x, out, accumulate); break;
    case 5: simt_apply_kernel<5><<<grid, 256, 0, s>>>(P, L.n_planes, plan.n, L.ld, x, out, accumulate); break;
    default: set_error("unsupported l"); return N2V2R_EINVAL;
  }
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" size_t n2v2r_apply_workspace_bytes(int64_t n, int l, int k_chunk) {
  StreamPlan p;
  if (make_stream_plan(n, l, N2V2R_ALGO_TCGEN05, k_chunk, &p) != N2V2R_OK) return 0;
  return p.xs_bytes + p.part_bytes + 256;
}

extern "C" int n2v2r_layer_apply(void* stream, const n2v2r_layer_t* h_layer, int64_t n, int l, int transpose,
                                 const float* d_x, float* d_out, int algo, int k_chunk, void* d_ws, size_t ws_bytes) {
  N2V2R_REQUIRE(h_layer && d_x && d_out, "layer_apply: null pointer");
  StreamPlan p;
  int rc = make_stream_plan(n, l, algo, k_chunk, &p);
  if (rc) return rc;
  Arena a(d_ws, ws_bytes);
  __nv_bfloat16* xs = a.take<__nv_bfloat16>(p.xs_bytes / sizeof(__nv_bfloat16));
  float* parts = a.take<float>(p.part_bytes / sizeof(float));
  if (algo == N2V2R_ALGO_TCGEN05 && (!d_ws || !a.ok())) { set_error("layer_apply: workspace too small"); return N2V2R_ENOMEM; }
  return stream_apply((cudaStream_t)stream, p, *h_layer, transpose, d_x, xs, parts, d_out, 0, nullptr);
}
