// UASE solver: randomised block subspace iteration with Rayleigh-Ritz on M = A A^T,
// A = [A_1 | ... | A_T].  Replaces spectral_embedding.UASE (model.py:54-55), i.e. scipy
// svds/ARPACK (_svds.py:428-535) + the sqrt(sigma) scaling of the right vectors.
//
// Per iteration (Q n x l orthonormal, replicated on every rank):
//   Z_t = A_t^T Q  (local layers)          -- streaming product, K4
//   Y   = sum_t A_t Z_t  (+ all-reduce)     -- streaming product, K3      => Y = M Q
//   H = Q^T Y, S = Y^T Y                    -- K5 (fp64 accumulate)
//   (W, Lam) = eig(H); C = W chol(W^T S W)^-1 -- K7 Jacobi + K6 Cholesky, one CTA
//   residuals ||Y w_j - lam_j Q w_j||, Q <- Y C, second CholeskyQR pass
// Stop: for every swept d, sqrt(sum_{j<=d} r_j^2) / (lam_d - lam_{d+1}) <= tol  (Davis-Kahan
// bound on the sine of the subspace angle), or the bound stagnates (fp32 floor), or max_iters.
// Final: G = sum_t Z_t^T Z_t (all-reduce) = W Sigma^2 W^T; embedding_t = Z_t W Sigma^-1/2.
#include "k5_orth.cuh"
#include <math.h>
#include <string.h>
#include <vector>
#include <algorithm>

namespace n2v2r {

struct SolverBufs {
  float *Q, *Y, *Qn;
  float* Z;                 // T_local * n * l
  __nv_bfloat16* xs;
  float* parts;
  double *partials, *rpart;
  double *H, *S, *W, *lam, *C, *C2, *G, *resid;
  int* flag;
};

static size_t carve(const n2v2r_uase_params_t* p, const StreamPlan& plan, void* ws, size_t bytes, SolverBufs* b) {
  Arena a(ws, bytes);
  const size_t nl = (size_t)p->n * p->l, ll = (size_t)p->l * p->l;
  b->Q = a.take<float>(nl);
  b->Y = a.take<float>(nl);
  b->Qn = a.take<float>(nl);
  b->Z = a.take<float>(nl * (size_t)p->T_local);
  b->xs = a.take<__nv_bfloat16>(plan.xs_bytes / sizeof(__nv_bfloat16));
  b->parts = a.take<float>(plan.part_bytes / sizeof(float));
  b->partials = a.take<double>((size_t)kMaxParts * ll);
  b->rpart = a.take<double>((size_t)sm_count() * 4 * p->l + 1024);
  b->H = a.take<double>(ll); b->S = a.take<double>(ll); b->W = a.take<double>(ll);
  b->C = a.take<double>(ll); b->C2 = a.take<double>(ll); b->G = a.take<double>(ll);
  b->lam = a.take<double>(p->l); b->resid = a.take<double>(p->l);
  b->flag = a.take<int>(4);
  return a.off;
}

static int check_params(const n2v2r_uase_params_t* p) {
  N2V2R_REQUIRE(p, "uase: null params");
  N2V2R_REQUIRE(p->n > 0 && p->T_local >= 1 && p->T_global >= p->T_local && p->layer_offset >= 0, "uase: bad layer counts");
  N2V2R_REQUIRE(p->l >= 16 && p->l <= kMaxL && p->l % 16 == 0, "uase: l must be 16, 32, 48, 64 or 80 (got %d)", p->l);
  N2V2R_REQUIRE(p->k >= 1 && p->k <= p->l, "uase: need 1 <= k <= l (k=%d, l=%d)", p->k, p->l);
  N2V2R_REQUIRE(p->k < p->n, "uase: k = %d must be smaller than the number of nodes %lld", p->k, (long long)p->n);  // svds: k < min(shape)
  N2V2R_REQUIRE(p->max_iters >= 1, "uase: max_iters must be >= 1");
  return N2V2R_OK;
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" size_t n2v2r_uase_workspace_bytes(const n2v2r_uase_params_t* p) {
  if (check_params(p) != N2V2R_OK) return 0;
  StreamPlan plan;
  if (make_stream_plan(p->n, p->l, p->algo, p->k_chunk, &plan) != N2V2R_OK) return 0;
  SolverBufs b;
  return carve(p, plan, nullptr, 0, &b) + 1024;
}

#define RC(call) do { int _rc = (call); if (_rc != N2V2R_OK) return _rc; } while (0)

extern "C" int n2v2r_uase(void* stream, const n2v2r_uase_params_t* p, const n2v2r_layer_t* L, void* d_ws, size_t ws_bytes,
                          n2v2r_allreduce_fn allreduce, void* user, float* d_embed, float* d_left, double* h_sigma,
                          n2v2r_uase_stats_t* st) {
  RC(check_params(p));
  N2V2R_REQUIRE(L && d_ws && d_embed && h_sigma, "uase: null pointer");
  N2V2R_REQUIRE(allreduce || p->T_local == p->T_global, "uase: sharded layers need an all-reduce callback");
  cudaStream_t s = (cudaStream_t)stream;
  StreamPlan plan;
  RC(make_stream_plan(p->n, p->l, p->algo, p->k_chunk, &plan));
  SolverBufs b;
  if (carve(p, plan, d_ws, ws_bytes, &b) > ws_bytes) { set_error("uase: workspace too small"); return N2V2R_ENOMEM; }
  const int64_t n = p->n;
  const int l = p->l, k = p->k, T = p->T_local;
  const size_t nl = (size_t)n * l;
  n2v2r_uase_stats_t stats;
  memset(&stats, 0, sizeof(stats));
  cudaEvent_t ev_all0, ev_all1;
  N2V2R_CUDA(cudaEventCreate(&ev_all0));
  N2V2R_CUDA(cudaEventCreate(&ev_all1));
  std::vector<cudaEvent_t> evs;
  N2V2R_CUDA(cudaEventRecord(ev_all0, s));

  // CholeskyQR2 of the buffer X -> orthonormal, result in `dst` (tmp is scratch)
  auto cholqr_pass = [&](const float* X, float* dst) -> int {
    RC(launch_xty(s, X, X, n, l, b.partials, kMaxParts, b.G, 0));
    RC(launch_smallmat(s, 1, l, b.G, nullptr, nullptr, nullptr, b.C, nullptr, b.flag));
    RC(launch_apply_small(s, X, n, l, b.C, l, l, dst, l, nullptr, nullptr, nullptr, nullptr, nullptr));
    stats.kernel_launches += 4;
    return N2V2R_OK;
  };

  RC(launch_random_block(s, b.Y, n, l, p->seed));
  RC(cholqr_pass(b.Y, b.Qn));
  RC(cholqr_pass(b.Qn, b.Q));
  stats.kernel_launches += 1;

  std::vector<double> lam(l), res(l), hist;
  double bound = INFINITY, best = INFINITY;
  int since_best = 0;
  int it = 0;
  bool converged = false, stagnated = false;
  while (true) {
    ++it;
    cudaEvent_t e0, e1;
    N2V2R_CUDA(cudaEventCreate(&e0));
    N2V2R_CUDA(cudaEventCreate(&e1));
    N2V2R_CUDA(cudaEventRecord(e0, s));
    for (int t = 0; t < T; ++t)                                   // Z_t = A_t^T Q
      RC(stream_apply(s, plan, L[t], 1, b.Q, b.xs, b.parts, b.Z + t * nl, 0, &stats.stream_bytes));
    for (int t = 0; t < T; ++t)                                   // Y = sum_t A_t Z_t
      RC(stream_apply(s, plan, L[t], 0, b.Z + t * nl, b.xs, b.parts, b.Y, t > 0, &stats.stream_bytes));
    N2V2R_CUDA(cudaEventRecord(e1, s));
    evs.push_back(e0); evs.push_back(e1);
    stats.layer_passes += 2 * T;
    stats.kernel_launches += 2 * T * (p->algo == N2V2R_ALGO_TCGEN05 ? 3 : 1);
    if (allreduce && allreduce(user, b.Y, (int64_t)nl, N2V2R_F32, stream) != 0) {
      set_error("uase: all-reduce callback failed (Y)");
      return N2V2R_ECOMM;
    }
    // Rayleigh-Ritz on span(Q) with Y = M Q
    RC(launch_xty(s, b.Q, b.Y, n, l, b.partials, kMaxParts, b.H, 0));
    RC(launch_xty(s, b.Y, b.Y, n, l, b.partials, kMaxParts, b.S, 0));
    RC(launch_smallmat(s, 0, l, b.H, b.S, b.W, b.lam, b.C, nullptr, b.flag));
    int n_parts = 0;
    RC(launch_apply_small(s, b.Y, n, l, b.C, l, l, b.Qn, l, b.Q, b.W, b.lam, b.rpart, &n_parts));
    RC(launch_sum_partials(s, b.rpart, n_parts, l, b.resid));
    stats.kernel_launches += 7;
    N2V2R_CUDA(cudaMemcpyAsync(lam.data(), b.lam, sizeof(double) * l, cudaMemcpyDeviceToHost, s));
    N2V2R_CUDA(cudaMemcpyAsync(res.data(), b.resid, sizeof(double) * l, cudaMemcpyDeviceToHost, s));
    N2V2R_CUDA(cudaStreamSynchronize(s));
    // stopping quantity
    bound = 0.0;
    auto angle_bound = [&](int d) -> double {
      double r2 = 0.0;
      for (int j = 0; j < d && j < l; ++j) r2 += res[j];
      const double gap = (d < l) ? lam[d - 1] - lam[d] : lam[d - 1];
      if (!(gap > 0.0)) return INFINITY;
      return sqrt(r2) / gap;
    };
    if (p->h_check_dims && p->n_check > 0) {
      for (int c = 0; c < p->n_check; ++c) {
        const int d = p->h_check_dims[c];
        if (d >= 1 && d <= k) bound = std::max(bound, angle_bound(d));
      }
    } else {
      bound = angle_bound(k);
    }
    if (bound < best * 0.9) { best = bound; since_best = 0; } else { ++since_best; }
    converged = bound <= p->tol;
    stagnated = since_best >= 12 && it >= p->min_iters;
    if ((converged && it >= p->min_iters) || stagnated || it >= p->max_iters) break;
    RC(cholqr_pass(b.Qn, b.Q));                                   // second CholeskyQR pass -> next Q
  }
  // final Rayleigh-Ritz through the right factor: G = sum_t Z_t^T Z_t = Q^T M Q
  for (int t = 0; t < T; ++t) RC(launch_xty(s, b.Z + t * nl, b.Z + t * nl, n, l, b.partials, kMaxParts, b.G, t > 0));
  if (allreduce && allreduce(user, b.G, (int64_t)l * l, N2V2R_F64, stream) != 0) {
    set_error("uase: all-reduce callback failed (G)");
    return N2V2R_ECOMM;
  }
  RC(launch_smallmat(s, 2, l, b.G, nullptr, b.W, b.lam, b.C, b.C2, b.flag));
  for (int t = 0; t < T; ++t)
    RC(launch_apply_small(s, b.Z + t * nl, n, l, b.C, l, k, d_embed + (size_t)t * n * k, k, nullptr, nullptr, nullptr, nullptr, nullptr));
  if (d_left) RC(launch_apply_small(s, b.Q, n, l, b.C2, l, k, d_left, k, nullptr, nullptr, nullptr, nullptr, nullptr));
  stats.kernel_launches += 2 * T + 1 + T + (d_left ? 1 : 0);
  N2V2R_CUDA(cudaMemcpyAsync(lam.data(), b.lam, sizeof(double) * l, cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaEventRecord(ev_all1, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  for (int j = 0; j < k; ++j) h_sigma[j] = sqrt(std::max(lam[j], 0.0));
  float ms = 0.f;
  for (size_t e = 0; e + 1 < evs.size(); e += 2) {
    N2V2R_CUDA(cudaEventElapsedTime(&ms, evs[e], evs[e + 1]));
    stats.stream_ms += ms;
    cudaEventDestroy(evs[e]); cudaEventDestroy(evs[e + 1]);
  }
  N2V2R_CUDA(cudaEventElapsedTime(&ms, ev_all0, ev_all1));
  stats.total_ms = ms;
  cudaEventDestroy(ev_all0); cudaEventDestroy(ev_all1);
  stats.iterations = it;
  stats.converged = converged ? 1 : 0;
  stats.max_angle_bound = bound;
  for (int j = 0; j < l && j < 96; ++j) {
    stats.ritz[j] = lam[j];
    stats.resid[j] = lam[j] > 0 ? sqrt(std::max(res[j], 0.0)) / lam[j] : 0.0;
  }
  if (st) *st = stats;
  if (!converged && !stagnated) {
    set_error("uase: reached max_iters=%d with subspace-angle bound %.3e > tol %.3e", p->max_iters, bound, p->tol);
    return N2V2R_ENOTCONV;
  }
  return N2V2R_OK;
}
