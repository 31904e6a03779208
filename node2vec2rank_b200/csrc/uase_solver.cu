// UASE solver: Chebyshev-filtered block subspace iteration with Rayleigh-Ritz, locking and
// deflation on M = A A^T, A = [A_1 | ... | A_T].  Replaces spectral_embedding.UASE
// (model.py:54-55), i.e. scipy svds/ARPACK (_svds.py:428-535) + the sqrt(sigma) scaling.
//
// Outer iteration (Q n x l orthonormal, replicated on every rank):
//   Z_t = A_t^T Q (local layers), Y = sum_t A_t Z_t (+ all-reduce)   -- streaming products K4/K3
//   H = Q^T Y -> (W, Lam) = eig(H)                                   -- K5 + K7 (one-CTA Jacobi, fp64)
//   Ritz vectors Qw = Q W, Yw = Y W, residuals r_j = ||Yw_j - lam_j Qw_j||
//   host: lock the leading Ritz pairs at the residual floor, test convergence, pick the degree m
//   degree-m Chebyshev filter (three-term recurrence, interval [0, lam_l] damped, scaled at the
//   largest unlocked Ritz value), every step deflated against the locked vectors:
//       X_{j+1} = (2 s_{j+1}/e)(M X_j - c X_j) - s_j s_{j+1} X_{j-1}
//   Q <- CholeskyQR2([locked | filtered])                            -- K5 + K6
// The first filter (Ritz values of a random block: no information about outliers, so no usable growth cap) has degree
// <= 2.  Outliers (lam_j >= 64 lam_k) are locked early, as soon as their angle estimate is below lam_k / (2 lam_j); every
// locked column re-enters the next basis after the power step of the Rayleigh-Ritz application (M q_j / lam_j), so it
// keeps converging at lam_l / lam_j per outer iteration while the filter works on the rest.
// Stop: Davis-Kahan bound sqrt(sum_{j<=d} r_j^2)/(lam_d - lam_{d+1}) <= tol at every swept d, or
// every wanted residual r_j <= max(8 eps32 lam_j, 0.02 eps32 lam_1) (the measured fp32 floor is ~eps32 lam_j),
// or the residuals stagnate, or max_iters applications.
// Final: embedding_t = Z_t W Sigma^-1/2 (= V_t Sigma^1/2) from the last Rayleigh-Ritz application.
#include "k5_orth.cuh"
#include <float.h>
#include <math.h>
#include <string.h>
#include <vector>
#include <algorithm>

namespace n2v2r {

int comm_allreduce_f32(void* comm, float* d_buf, int64_t count, cudaStream_t s);   // comm.cu
int batch_apply(void* handle, int slot, const float* x, float* out, float* z_keep, float* zt, cudaStream_t s);   // batch.cu
void batch_leave(void* handle, int slot);

struct SolverBufs {
  float *Q, *Y, *Qn, *Qw, *X0, *X1, *X2, *Zt;
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
  b->Qw = a.take<float>(nl);
  b->X0 = a.take<float>(nl);
  b->X1 = a.take<float>(nl);
  b->X2 = a.take<float>(nl);
  b->Zt = a.take<float>(nl);
  b->Z = a.take<float>(nl * (size_t)p->T_local);
  b->xs = a.take<__nv_bfloat16>(plan.xs_bytes / sizeof(__nv_bfloat16));
  b->parts = a.take<float>(plan.part_bytes / sizeof(float));
  b->partials = a.take<double>((size_t)kMaxParts * ll);
  b->rpart = a.take<double>((size_t)kMaxParts * p->l + 1024);
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

static double cheb_T(double x, int m) { return cosh((double)m * acosh(x > 1.0 ? x : 1.0)); }

extern "C" int n2v2r_uase(void* stream, const n2v2r_uase_params_t* p, const n2v2r_layer_t* L, void* d_ws, size_t ws_bytes,
                          n2v2r_allreduce_fn allreduce, void* user, float* d_embed, float* d_left, double* h_sigma,
                          n2v2r_uase_stats_t* st) {
  struct LeaveBatch {                      // however the solve ends, the other solves of the batch stop waiting for it
    void* b; int slot;
    ~LeaveBatch() { if (b) batch_leave(b, slot); }
  } leave_batch{p ? p->batch : nullptr, p ? p->batch_slot : 0};
  RC(check_params(p));
  N2V2R_REQUIRE(L && d_ws && d_embed && h_sigma, "uase: null pointer");
  N2V2R_REQUIRE(!p->batch || p->T_local == p->T_global, "uase: seed batching runs on unsharded layers");
  N2V2R_REQUIRE(allreduce || p->T_local == p->T_global || (p->comm_mode == 1 && user),
                "uase: sharded layers need an all-reduce callback or an NCCL communicator (comm_mode = 1)");
  cudaStream_t s = (cudaStream_t)stream;
  StreamPlan plan;
  RC(make_stream_plan(p->n, p->l, p->algo, p->k_chunk, &plan));
  SolverBufs b;
  if (carve(p, plan, d_ws, ws_bytes, &b) > ws_bytes) { set_error("uase: workspace too small"); return N2V2R_ENOMEM; }
  const int64_t n = p->n;
  const int l = p->l, k = p->k, T = p->T_local;
  const size_t nl = (size_t)n * l;
  const int m_cap = p->cheb_degree <= 0 ? 32 : p->cheb_degree;
  // residual target: 8 eps32 lam_j (measured fp32 floor ~ eps32 lam_j); tuning comes through the params (0 = default)
  const double floor_mult = p->floor_mult > 0.0 ? p->floor_mult : 0.02;
  const double rel_floor = p->rel_floor > 0.0 ? p->rel_floor : 8.0 * (double)FLT_EPSILON;
  const bool trace = p->trace != 0;
  // The Ritz values of the random start block say nothing about outliers (C3: lam_1 is estimated 500x too small), so the
  // growth cap below cannot protect the first filter: its degree is limited to 2.
  const int first_cap = p->first_degree > 0 ? p->first_degree : 2;
  const bool early_lock = p->no_early_lock == 0;
  const double need_extra = 1.0;
  const double rho_max = 1e4;            // cap on T_m(x_unlocked)/T_m(x_k): keeps every fp32 column dominated by its own Ritz vector
  n2v2r_uase_stats_t stats;
  memset(&stats, 0, sizeof(stats));
  cudaEvent_t ev_all0, ev_all1;
  N2V2R_CUDA(cudaEventCreate(&ev_all0));
  N2V2R_CUDA(cudaEventCreate(&ev_all1));
  std::vector<cudaEvent_t> evs, mma_evs;
  struct SinkGuard { SinkGuard(std::vector<cudaEvent_t>* v) { set_mma_event_sink(v); } ~SinkGuard() { set_mma_event_sink(nullptr); } }
      sink_guard(p->algo == N2V2R_ALGO_TCGEN05 ? &mma_evs : nullptr);
  N2V2R_CUDA(cudaEventRecord(ev_all0, s));
  const int per_pass = (p->algo == N2V2R_ALGO_TCGEN05 ? 3 : 1);
  bool slabs = false;
  for (int t = 0; t < T; ++t) slabs = slabs || L[t].rows > 0;
  for (int t = 0; t < T; ++t) N2V2R_REQUIRE(!slabs || L[t].rows > 0, "uase: row-slab and whole layers cannot be mixed");
  N2V2R_REQUIRE(!slabs || (p->comm_mode == 1 && user), "uase: row slabs need the library-owned NCCL communicator (comm_mode = 1)");
  N2V2R_REQUIRE(!slabs || !p->batch, "uase: row slabs and seed batching cannot be combined");

  // one pass of CholeskyQR: dst = X chol(X^T X)^-1
  auto cholqr_pass = [&](const float* X, float* dst) -> int {
    RC(launch_xty(s, X, X, n, l, b.partials, kMaxParts, b.G, 0));
    RC(launch_smallmat(s, 1, l, b.G, nullptr, nullptr, nullptr, b.C, nullptr, b.flag));
    RC(launch_apply_small(s, X, n, l, b.C, l, l, dst, l));
    stats.kernel_launches += 4;
    return N2V2R_OK;
  };
  // out = M x = sum_t A_t (A_t^T x) (+ all-reduce).  keep_z: Z_t kept per layer (Rayleigh-Ritz application)
  auto apply_M = [&](const float* x, float* out, bool keep_z) -> int {
    cudaEvent_t e0, e1;
    N2V2R_CUDA(cudaEventCreate(&e0));
    N2V2R_CUDA(cudaEventCreate(&e1));
    N2V2R_CUDA(cudaEventRecord(e0, s));
    if (p->batch) {                          // served together with the other solves of the batch (batch.cu)
      RC(batch_apply(p->batch, p->batch_slot, x, out, keep_z ? b.Z : nullptr, b.Zt, s));
    } else if (slabs) {
      // row slabs (fewer layers than GPUs): this rank holds rows [r0, r0 + m) of each of its layers.  Z_t = sum over the
      // slabs of slab^T x[slab rows] (all-reduce inside the layer's group); out[slab rows] += slab Z_t; the global all-reduce
      // below adds the layers and assembles the slabs.
      N2V2R_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * nl, s));
      for (int t = 0; t < T; ++t) {
        float* z = keep_z ? b.Z + t * nl : b.Zt;
        RC(stream_apply_slab(s, plan, L[t], 1, x + (size_t)L[t].row_offset * l, b.xs, b.parts, z, 0, &stats.stream_bytes));
        if (p->group_comm) RC(comm_allreduce_f32(p->group_comm, z, (int64_t)nl, s));
        RC(stream_apply_slab(s, plan, L[t], 0, z, b.xs, b.parts, out + (size_t)L[t].row_offset * l, 1, &stats.stream_bytes));
      }
    } else {
      for (int t = 0; t < T; ++t) {
        float* z = keep_z ? b.Z + t * nl : b.Zt;
        RC(stream_apply(s, plan, L[t], 1, x, b.xs, b.parts, z, 0, &stats.stream_bytes));       // Z_t = A_t^T x
        RC(stream_apply(s, plan, L[t], 0, z, b.xs, b.parts, out, t > 0, &stats.stream_bytes));  // out += A_t Z_t
      }
    }
    N2V2R_CUDA(cudaEventRecord(e1, s));
    evs.push_back(e0); evs.push_back(e1);
    stats.layer_passes += 2 * T;
    stats.kernel_launches += 2 * T * per_pass;
    stats.iterations += 1;
    if (p->comm_mode == 1 && user && (p->T_local != p->T_global || slabs)) {
      RC(comm_allreduce_f32(user, out, (int64_t)nl, s));
    } else if (allreduce && allreduce(user, out, (int64_t)nl, N2V2R_F32, stream) != 0) {
      set_error("uase: all-reduce callback failed (Y)");
      return N2V2R_ECOMM;
    }
    return N2V2R_OK;
  };
  auto deflate = [&](float* X, int locked) -> int {
    if (locked <= 0) return N2V2R_OK;
    RC(launch_xty(s, b.Qw, X, n, l, b.partials, kMaxParts, b.G, 0));     // rows < locked of Qw^T X are used
    RC(launch_sub_proj(s, X, b.Qw, n, l, locked, b.G));
    stats.kernel_launches += 3;
    return N2V2R_OK;
  };

  RC(launch_random_block(s, b.Y, n, l, p->seed));
  RC(cholqr_pass(b.Y, b.Qn));
  RC(cholqr_pass(b.Qn, b.Q));
  stats.kernel_launches += 1;

  std::vector<double> lam(l), res(l);
  double bound = INFINITY;
  int locked = 0, outer = 0, stalls = 0;
  bool converged = false, stagnated = false, last_filtered = false, bound_met = false, floor_met = false;
  std::vector<double> dim_bounds(N2V2R_MAX_CHECK, 0.0);
  double best_worst = INFINITY, prev_worst = INFINITY, predicted_gain = 1.0;
  while (true) {
    ++outer;
    RC(apply_M(b.Q, b.Y, true));
    // Rayleigh-Ritz on span(Q): H = Q^T M Q
    RC(launch_xty(s, b.Q, b.Y, n, l, b.partials, kMaxParts, b.H, 0));
    RC(launch_smallmat(s, 0, l, b.H, nullptr, b.W, b.lam, nullptr, nullptr, b.flag));
    RC(launch_apply_small(s, b.Y, n, l, b.W, l, l, b.X1, l));                                        // X1 = Yw = Y W
    RC(launch_apply_small(s, b.Q, n, l, b.W, l, l, b.Qw, l));                                        // Qw = Q W
    RC(launch_resid(s, b.X1, b.Qw, n, l, b.lam, b.rpart, kMaxParts, b.resid));                       // ||Yw - Qw Lam||^2 per column
    stats.kernel_launches += 6;
    N2V2R_CUDA(cudaMemcpyAsync(lam.data(), b.lam, sizeof(double) * l, cudaMemcpyDeviceToHost, s));
    N2V2R_CUDA(cudaMemcpyAsync(res.data(), b.resid, sizeof(double) * l, cudaMemcpyDeviceToHost, s));
    N2V2R_CUDA(cudaStreamSynchronize(s));
    for (int j = 0; j < l; ++j) res[j] = sqrt(std::max(res[j], 0.0));    // residual norms
    // ---- host decisions --------------------------------------------------------------------
    auto angle_bound = [&](int d) -> double {
      double r2 = 0.0;
      for (int j = 0; j < d && j < l; ++j) r2 += res[j] * res[j];
      const double gap = (d < l) ? lam[d - 1] - lam[d] : lam[d - 1];
      if (!(gap > 0.0)) return INFINITY;
      return sqrt(r2) / gap;
    };
    bound = 0.0;
    if (p->h_check_dims && p->n_check > 0) {
      for (int c = 0; c < p->n_check; ++c) {
        const int d = p->h_check_dims[c];
        if (d >= 1 && d <= k) {
          const double bd = angle_bound(d);
          bound = std::max(bound, bd);
          if (c < N2V2R_MAX_CHECK) dim_bounds[c] = bd;
        }
      }
    } else {
      bound = angle_bound(k);
      dim_bounds[0] = bound;
    }
    const double floor_ = floor_mult * (double)FLT_EPSILON * std::max(lam[0], 0.0);
    auto target = [&](int j) { return std::max(floor_, rel_floor * lam[j]); };
    auto at_floor = [&](int j) { return res[j] <= target(j); };
    // Outliers are locked early: a Ritz pair with lam_j >= 64 lam_k whose angle estimate res_j / gap_j is below
    // lam_k / (2 lam_j) is deflated from the filter like a converged one (per filter step its leakage x_j * angle then stays
    // below the growth of the wanted directions, which is what the growth cap needs) and keeps converging through the
    // Rayleigh-Ritz application of every outer iteration at a factor lam_l / lam_j <= 1/64.  Convergence is still only
    // declared on the residual targets (all_floor below).
    auto early = [&](int j) {
      if (!early_lock || outer < 2 || j + 1 >= l || !(lam[j] >= 64.0 * lam[k - 1]) || !(lam[k - 1] > 0.0)) return false;
      const double gap = lam[j] - lam[j + 1];
      return gap > 0.0 && res[j] / gap <= 0.5 * lam[k - 1] / lam[j];
    };
    locked = 0;
    while (locked < k && (at_floor(locked) || early(locked))) ++locked;
    bool all_floor = true;
    for (int j = 0; j < k; ++j) all_floor = all_floor && at_floor(j);
    // stagnation = the fp32 residual floor of this matrix is above the target.  After a filtered iteration:
    //  (A) close to the target (worst <= 16) and the drop was < 1/30 of what the filter should have given, or
    //  (B) worst <= 64 and the drop was < 2x, or (C) two filtered iterations in a row without a 30 % drop.
    double worst = 0.0;
    for (int j = 0; j < k; ++j) worst = std::max(worst, res[j] / target(j));
    // A plain step (degree 1: the growth cap forbids a filter because lam_1 / lam_k is already large) is judged the
    // same way when it should have gained at least 4x (its factor is lam_k / lam_l); slower plain steps are not judged.
    if (last_filtered || (outer > 1 && predicted_gain >= 4.0)) {
      const double drop = prev_worst / std::max(worst, 1e-300);
      if (worst <= 16.0 && drop < predicted_gain / 30.0) stagnated = true;
      if (worst <= 64.0 && drop < 2.0) stagnated = true;
      if (worst > 0.7 * best_worst) ++stalls; else stalls = 0;
      if (stalls >= 2) stagnated = true;
    }
    best_worst = std::min(best_worst, worst);
    prev_worst = worst;
    bound_met = bound <= p->tol;
    floor_met = all_floor;
    converged = bound_met || all_floor || stagnated;
    last_filtered = false;
    if ((converged && stats.iterations >= p->min_iters) || stats.iterations >= p->max_iters) break;
    // ---- filter degree ----------------------------------------------------------------------
    const double hi = std::max(lam[l - 1], 0.0);           // damp [0, lam_l]
    const double c = 0.5 * hi, e = 0.5 * hi;
    const double lam_u = lam[locked < k ? locked : k - 1];   // largest unlocked Ritz value
    int m = 1;
    if (m_cap > 1 && e > 0.0 && lam[k - 1] > hi) {
      const double xu = (lam_u - c) / e, xk = (lam[k - 1] - c) / e;
      while (m < m_cap && cheb_T(xu, m + 1) / cheb_T(xk, m + 1) <= rho_max) ++m;
      const double need = ceil(log(4.0 * std::max(worst, 1.0)) / acosh(xk)) + need_extra;   // degree that would reach the targets
      if (m > 1) m = std::max(2, std::min(m, (int)std::min(need, 1e6)));
      m = std::min(m, p->max_iters - stats.iterations);
      if (outer == 1 && first_cap > 0) m = std::min(m, first_cap);
    }
    if (trace)
      fprintf(stderr, "[uase] outer %d apps %d locked %d bound %.3e worst %.3e lam1 %.4e lamk %.4e lam_l %.4e -> degree %d\n", outer,
              stats.iterations, locked, bound, worst, lam[0], lam[k - 1], lam[l - 1], m);
    if (m <= 1) {                                            // plain subspace step: Q <- orth(M Qw)
      predicted_gain = hi > 0.0 ? lam[k - 1] / hi : 1.0;
      RC(cholqr_pass(b.X1, b.Qn));
      RC(cholqr_pass(b.Qn, b.Q));
      continue;
    }
    last_filtered = true;
    predicted_gain = cheb_T((lam[k - 1] - c) / e, m);
    // X0 = Qw, X1 = (s1/e)(Yw - c Qw), then the three-term recurrence
    const double s1 = e / (lam_u - c);
    double sg = s1;
    N2V2R_CUDA(cudaMemcpyAsync(b.X0, b.Qw, sizeof(float) * nl, cudaMemcpyDeviceToDevice, s));
    if (locked > 0) {
      // the locked columns leave the filter, but not the iteration: they re-enter the next basis after the power step of this
      // Rayleigh-Ritz application, Yw_j / lam_j = M q_j / lam_j (stashed in Qn before Yw is overwritten)
      std::vector<double> dg((size_t)l * locked, 0.0);
      for (int j = 0; j < locked; ++j) dg[(size_t)j * locked + j] = lam[j] > 0.0 ? 1.0 / lam[j] : 1.0;
      N2V2R_CUDA(cudaMemcpyAsync(b.C2, dg.data(), sizeof(double) * dg.size(), cudaMemcpyHostToDevice, s));
      N2V2R_CUDA(cudaStreamSynchronize(s));          // dg is a pageable temporary
      RC(launch_apply_small(s, b.X1, n, l, b.C2, locked, locked, b.Qn, l));
      stats.kernel_launches += 1;
    }
    RC(launch_axpbypcz(s, b.X1, b.X1, s1 / e, b.Qw, -c * s1 / e, nullptr, 0.0, (int64_t)nl));
    stats.kernel_launches += 2;
    float *x0 = b.X0, *x1 = b.X1, *x2 = b.X2;
    for (int j = 2; j <= m; ++j) {
      RC(deflate(x1, locked));
      RC(apply_M(x1, b.Y, false));
      const double s2 = 1.0 / (2.0 / s1 - sg);
      RC(launch_axpbypcz(s, x2, b.Y, 2.0 * s2 / e, x1, -2.0 * s2 * c / e, x0, -sg * s2, (int64_t)nl));
      stats.kernel_launches += 1;
      float* t = x0; x0 = x1; x1 = x2; x2 = t;
      sg = s2;
    }
    if (locked > 0) {
      RC(deflate(x1, locked));
      N2V2R_CUDA(cudaMemcpy2DAsync(x1, sizeof(float) * l, b.Qn, sizeof(float) * l, sizeof(float) * locked, (size_t)n,
                                   cudaMemcpyDeviceToDevice, s));
    }
    RC(cholqr_pass(x1, b.Qn));
    RC(cholqr_pass(b.Qn, b.Q));
  }
  if (trace)
    fprintf(stderr, "[uase] stop: outer %d apps %d locked %d bound %.3e converged %d stagnated %d\n", outer, stats.iterations, locked,
            bound, (int)converged, (int)stagnated);
  // embeddings from the last Rayleigh-Ritz application: Z_t W Sigma^-1/2 ; left = Q W Sigma^1/2
  RC(launch_smallmat(s, 2, l, b.H, nullptr, b.W, b.lam, b.C, b.C2, b.flag));
  for (int t = 0; t < T; ++t)
    RC(launch_apply_small(s, b.Z + t * nl, n, l, b.C, l, k, d_embed + (size_t)t * n * k, k));
  if (d_left) RC(launch_apply_small(s, b.Q, n, l, b.C2, l, k, d_left, k));
  stats.kernel_launches += 1 + T + (d_left ? 1 : 0);
  N2V2R_CUDA(cudaEventRecord(ev_all1, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  for (int j = 0; j < k; ++j) h_sigma[j] = sqrt(std::max(lam[j], 0.0));
  float ms = 0.f;
  for (size_t e = 0; e + 1 < evs.size(); e += 2) {
    N2V2R_CUDA(cudaEventElapsedTime(&ms, evs[e], evs[e + 1]));
    stats.stream_ms += ms;
    cudaEventDestroy(evs[e]); cudaEventDestroy(evs[e + 1]);
  }
  for (size_t e = 0; e + 1 < mma_evs.size(); e += 2) {
    N2V2R_CUDA(cudaEventElapsedTime(&ms, mma_evs[e], mma_evs[e + 1]));
    stats.mma_ms += ms;
    cudaEventDestroy(mma_evs[e]); cudaEventDestroy(mma_evs[e + 1]);
  }
  stats.mma_launches = (int)(mma_evs.size() / 2);
  N2V2R_CUDA(cudaEventElapsedTime(&ms, ev_all0, ev_all1));
  stats.total_ms = ms;
  cudaEventDestroy(ev_all0); cudaEventDestroy(ev_all1);
  stats.outer_iterations = outer;
  stats.locked = locked;
  stats.converged = converged ? ((bound_met || floor_met) ? 1 : 2) : 0;   // 2 = residuals stagnated above the target
  stats.bound_met = bound_met ? 1 : 0;
  stats.max_angle_bound = bound;
  for (int c = 0; c < N2V2R_MAX_CHECK; ++c) stats.angle_bounds[c] = dim_bounds[c];
  for (int j = 0; j < l && j < 96; ++j) {
    stats.ritz[j] = lam[j];
    stats.resid[j] = lam[0] > 0 ? res[j] / lam[0] : 0.0;     // residual norms relative to lam_1
  }
  if (st) *st = stats;
  if (!converged) {
    set_error("uase: reached max_iters=%d operator applications with subspace-angle bound %.3e > tol %.3e (%d of %d Ritz pairs at the residual floor)",
              p->max_iters, bound, p->tol, locked, k);
    return N2V2R_ENOTCONV;
  }
  return N2V2R_OK;
}
