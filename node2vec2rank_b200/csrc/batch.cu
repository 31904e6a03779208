// Seed batching for stability runs (README.md:54, :173: "running it multiple times and aggregating"; BASELINE configs[4]):
// several UASE solves on the SAME resident layers, each from its own start block, run concurrently -- one host thread and
// one CUDA stream per solve -- and meet in this object whenever they need the operator M X = sum_t A_t (A_t^T X).  The
// requests that have arrived are served together: every layer is streamed ONCE per product for up to three operands
// (stream_apply_batched: the operands' accumulators sit side by side in TMEM), instead of once per solve.  Each solve keeps
// its own adaptive control flow (filter degrees, locking, stopping), and receives exactly the bits the unbatched solve gets.
#include "k5_orth.cuh"
#include <condition_variable>
#include <mutex>
#include <vector>

namespace n2v2r {

int stream_apply_batched(cudaStream_t s, const StreamPlan& plan, const void* planes, int n_planes, int64_t ld, int nb,
                         const float* const* x, __nv_bfloat16* xs, float* parts, float* const* out, int accumulate);   // k3_tcgen05.cu

constexpr int kMaxBatch = 16;

struct BatchRequest {
  const float* x; float* out; float* z_keep; float* zt; cudaStream_t s;
};

struct Batch {
  std::mutex mu;
  std::condition_variable cv;
  int participants = 0, active = 0, submitted = 0;
  uint64_t generation = 0;
  bool has[kMaxBatch] = {};
  BatchRequest req[kMaxBatch];
  cudaEvent_t ready[kMaxBatch] = {};
  cudaEvent_t done[2] = {};
  cudaStream_t bstream = nullptr;
  StreamPlan plan;
  std::vector<n2v2r_layer_t> layers;
  __nv_bfloat16* xs = nullptr;
  float* parts = nullptr;
  int max_nb = 1;
  int error = N2V2R_OK;
  long launches = 0, layer_streams = 0, products = 0;
  std::vector<cudaEvent_t> mma_events;
};

static size_t batch_carve(const StreamPlan& plan, int nb, void* ws, __nv_bfloat16** xs, float** parts) {
  Arena a(ws, (size_t)-1);
  *xs = a.take<__nv_bfloat16>((size_t)nb * 3 * plan.l * plan.n_pad + 32768);
  *parts = a.take<float>((size_t)(plan.n_splits + 1) * round_up(plan.n, 256) * plan.l * nb);
  return a.off;
}

static int max_blocks(int l) { int nb = 512 / (3 * l); return nb > 3 ? 3 : (nb < 1 ? 1 : nb); }

// serve every submitted request (called with the lock held): layer by layer, groups of up to max_nb operands per launch
static int batch_launch(Batch* B) {
  int idx[kMaxBatch], m = 0;
  for (int i = 0; i < B->participants; ++i)
    if (B->has[i]) idx[m++] = i;
  if (m == 0) return N2V2R_OK;
  for (int j = 0; j < m; ++j) {
    N2V2R_CUDA(cudaEventRecord(B->ready[idx[j]], B->req[idx[j]].s));
    N2V2R_CUDA(cudaStreamWaitEvent(B->bstream, B->ready[idx[j]], 0));
  }
  set_mma_event_sink(&B->mma_events);
  const size_t nl = (size_t)B->plan.n * B->plan.l;
  int rc = N2V2R_OK;
  for (size_t t = 0; t < B->layers.size() && rc == N2V2R_OK; ++t) {
    const n2v2r_layer_t& L = B->layers[t];
    const bool batchable = L.storage == N2V2R_STORE_PLANES && B->plan.algo == N2V2R_ALGO_TCGEN05;
    for (int g0 = 0; g0 < m && rc == N2V2R_OK; g0 += B->max_nb) {
      const int nb = (m - g0 < B->max_nb) ? m - g0 : B->max_nb;
      const float* xin[3]; float* zout[3]; float* yout[3];
      for (int j = 0; j < nb; ++j) {
        const BatchRequest& r = B->req[idx[g0 + j]];
        xin[j] = r.x;
        zout[j] = r.z_keep ? r.z_keep + t * nl : r.zt;
        yout[j] = r.out;
      }
      if (batchable && nb > 1) {
        const void* pt = L.d_planes_t ? L.d_planes_t : L.d_planes;
        rc = stream_apply_batched(B->bstream, B->plan, pt, L.n_planes, L.ld, nb, xin, B->xs, B->parts, zout, 0);            // Z_t = A_t^T x
        if (rc == N2V2R_OK) {
          const float* zin[3] = {zout[0], zout[1], zout[2]};
          rc = stream_apply_batched(B->bstream, B->plan, L.d_planes, L.n_planes, L.ld, nb, zin, B->xs, B->parts, yout, t > 0);   // out += A_t Z_t
        }
        B->launches += 2;
      } else {
        for (int j = 0; j < nb && rc == N2V2R_OK; ++j) {
          rc = stream_apply(B->bstream, B->plan, L, 1, xin[j], B->xs, B->parts, zout[j], 0, nullptr);
          if (rc == N2V2R_OK) rc = stream_apply(B->bstream, B->plan, L, 0, zout[j], B->xs, B->parts, yout[j], t > 0, nullptr);
          B->launches += 2;
        }
      }
      B->layer_streams += 2;
    }
  }
  set_mma_event_sink(nullptr);
  B->products += m;
  if (rc != N2V2R_OK) return rc;
  N2V2R_CUDA(cudaEventRecord(B->done[B->generation & 1], B->bstream));
  return N2V2R_OK;
}

static void batch_release(Batch* B) {        // lock held: close the generation and wake the waiters
  for (int i = 0; i < B->participants; ++i) B->has[i] = false;
  B->submitted = 0;
  ++B->generation;
  B->cv.notify_all();
}

// out = M x for participant `slot`; returns when the launches have been queued and `s` has been made to wait for them
int batch_apply(void* handle, int slot, const float* x, float* out, float* z_keep, float* zt, cudaStream_t s) {
  Batch* B = (Batch*)handle;
  N2V2R_REQUIRE(B && slot >= 0 && slot < B->participants, "batch: bad slot");
  std::unique_lock<std::mutex> lk(B->mu);
  if (B->error != N2V2R_OK) return B->error;
  B->req[slot] = BatchRequest{x, out, z_keep, zt, s};
  B->has[slot] = true;
  ++B->submitted;
  const uint64_t gen = B->generation;
  if (B->submitted >= B->active) {
    const int rc = batch_launch(B);
    if (rc != N2V2R_OK) B->error = rc;
    batch_release(B);
  } else {
    B->cv.wait(lk, [&] { return B->generation != gen; });
  }
  if (B->error != N2V2R_OK) return B->error;
  N2V2R_CUDA(cudaStreamWaitEvent(s, B->done[gen & 1], 0));
  return N2V2R_OK;
}

// a solve has finished (or failed): the others no longer wait for it
void batch_leave(void* handle, int slot) {
  Batch* B = (Batch*)handle;
  if (!B || slot < 0 || slot >= B->participants) return;
  std::unique_lock<std::mutex> lk(B->mu);
  --B->active;
  if (B->active > 0 && B->submitted >= B->active) {
    const int rc = batch_launch(B);
    if (rc != N2V2R_OK) B->error = rc;
    batch_release(B);
  }
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" size_t n2v2r_batch_workspace_bytes(const n2v2r_uase_params_t* p, int participants) {
  if (!p || participants < 1) return 0;
  StreamPlan plan;
  if (make_stream_plan(p->n, p->l, p->algo, p->k_chunk, &plan) != N2V2R_OK) return 0;
  __nv_bfloat16* xs; float* parts;
  return batch_carve(plan, max_blocks(p->l), nullptr, &xs, &parts) + 1024;
}

extern "C" int n2v2r_batch_create(const n2v2r_uase_params_t* p, const n2v2r_layer_t* h_layers, int participants, void* d_ws,
                                  size_t ws_bytes, void** handle) {
  N2V2R_REQUIRE(p && h_layers && handle && d_ws && participants >= 1 && participants <= kMaxBatch, "batch_create: bad arguments (1..%d participants)", kMaxBatch);
  N2V2R_REQUIRE(p->T_local == p->T_global, "batch_create: seed batching runs on unsharded layers");
  Batch* B = new Batch();
  int rc = make_stream_plan(p->n, p->l, p->algo, p->k_chunk, &B->plan);
  if (rc) { delete B; return rc; }
  B->max_nb = max_blocks(p->l);
  if (batch_carve(B->plan, B->max_nb, d_ws, &B->xs, &B->parts) > ws_bytes) { delete B; set_error("batch_create: workspace too small"); return N2V2R_ENOMEM; }
  B->participants = B->active = participants;
  B->layers.assign(h_layers, h_layers + p->T_local);
  cudaError_t e = cudaStreamCreateWithFlags(&B->bstream, cudaStreamNonBlocking);
  for (int i = 0; i < participants && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&B->ready[i], cudaEventDisableTiming);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&B->done[i], cudaEventDisableTiming);
  if (e != cudaSuccess) { delete B; return cuda_fail(e, "batch_create", __FILE__, __LINE__); }
  *handle = B;
  return N2V2R_OK;
}

// stats: [0] product requests served, [1] layer streams (passes over a layer), [2] MMA-kernel milliseconds, [3] MMA launches
extern "C" int n2v2r_batch_destroy(void* handle, double* h_stats4) {
  Batch* B = (Batch*)handle;
  if (!B) return N2V2R_OK;
  if (B->bstream) cudaStreamSynchronize(B->bstream);
  double ms_total = 0.0;
  for (size_t e = 0; e + 1 < B->mma_events.size(); e += 2) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, B->mma_events[e], B->mma_events[e + 1]) == cudaSuccess) ms_total += ms;
    cudaEventDestroy(B->mma_events[e]); cudaEventDestroy(B->mma_events[e + 1]);
  }
  if (h_stats4) {
    h_stats4[0] = (double)B->products; h_stats4[1] = (double)B->layer_streams; h_stats4[2] = ms_total;
    h_stats4[3] = (double)(B->mma_events.size() / 2);
  }
  for (int i = 0; i < kMaxBatch; ++i) if (B->ready[i]) cudaEventDestroy(B->ready[i]);
  for (int i = 0; i < 2; ++i) if (B->done[i]) cudaEventDestroy(B->done[i]);
  if (B->bstream) cudaStreamDestroy(B->bstream);
  const int err = B->error;
  delete B;
  return err;
}
