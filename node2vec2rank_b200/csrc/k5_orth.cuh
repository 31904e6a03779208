// Host-side launchers of the small dense kernels (k5_orth.cu) and the streaming products
// (k3_stream.cu), shared with the solver (uase_solver.cu).
#pragma once
#include "common.cuh"
#include <vector>

namespace n2v2r {

constexpr int kMaxL = 80;
constexpr int kMaxParts = 512;

// out (l x l fp64) [+]= X^T Y, deterministic two-stage reduction; partials: kMaxParts*l*l doubles
int launch_xty(cudaStream_t s, const float* X, const float* Y, int64_t n, int l, double* partials, int max_parts,
               double* out, int accumulate);
// mode 0 RR, 1 CHOL, 2 FINAL -- see k5_orth.cu
int launch_smallmat(cudaStream_t s, int mode, int l, const double* H, const double* S, double* W, double* lam, double* C,
                    double* C2, int* flag);
// Out[n][lo] = X[n][l] C[l][lo]  (fp64 accumulate)
int launch_apply_small(cudaStream_t s, const float* X, int64_t n, int l, const double* C, int ldc, int lo, float* Out, int64_t ld_out);
// out[j] = ||Yw[:, j] - lam[j] Qw[:, j]||^2 ; partials: max_parts * l doubles
int launch_resid(cudaStream_t s, const float* Yw, const float* Qw, int64_t n, int l, const double* lam, double* partials,
                 int max_parts, double* out);
// X -= U[:, :p] G[:p, :]  (deflation against the p locked Ritz vectors)
int launch_sub_proj(cudaStream_t s, float* X, const float* U, int64_t n, int l, int p, const double* G);
int launch_sum_partials(cudaStream_t s, const double* partials, int n_parts, int count, double* out);
int launch_split_transpose(cudaStream_t s, const float* X, int64_t n, int l, int64_t n_pad, __nv_bfloat16* Xs);
int launch_random_block(cudaStream_t s, float* X, int64_t n, int l, uint64_t seed);
int launch_axpbypcz(cudaStream_t s, float* out, const float* x, double a, const float* y, double b, const float* z, double c,
                    int64_t count);

// ---- streaming skinny products (k3_stream.cu) ----------------------------------------------
struct StreamPlan {
  int64_t n, n_pad;
  int l, algo, k_chunk, n_splits;
  int k_chunk_user;                // K chunk fixed by the caller (0 = kernel picks)
  size_t xs_bytes, part_bytes;     // split-transposed operand, K-split partial sums
};
void set_mma_event_sink(std::vector<cudaEvent_t>* sink);   // k3_tcgen05.cu: time every MMA launch (nullptr = off)
int make_stream_plan(int64_t n, int l, int algo, int k_chunk, StreamPlan* plan);
// out[n][l] (= or +=) A X (transpose=0) / A^T X (transpose=1).  x: fp32 [n][l].
// xs / parts: scratch sized by the plan (may be null for the SIMT algo).
int stream_apply(cudaStream_t s, const StreamPlan& plan, const n2v2r_layer_t& L, int transpose, const float* x,
                 __nv_bfloat16* xs, float* parts, float* out, int accumulate, double* bytes_moved);
// row slab of a layer (L.rows > 0): transpose = 0: out[0..rows) (+)= slab x (x: n x l);  transpose = 1: out (n x l) (+)= slab^T x
// (x: rows x l).  The caller offsets the pointers into the full n x l blocks.
int stream_apply_slab(cudaStream_t s, const StreamPlan& plan, const n2v2r_layer_t& L, int transpose, const float* x,
                      __nv_bfloat16* xs, float* parts, float* out, int accumulate, double* bytes_moved);

}  // namespace n2v2r
