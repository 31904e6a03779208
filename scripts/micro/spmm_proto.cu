// Micro-benchmark / prototype: fp32 row-list product Y = A X for a sparse layer (C3: n = 50 000, 2.64 % dense, l = 48) on the
// CUDA cores.  Groups of 4 lanes own two rows each and keep their accumulators in registers for the whole pass; the operand X
// streams through shared memory in panels of W rows (one bulk copy per panel, ring of S stages) and every entry gathers its
// X row from there (l x 4 bytes of shared-memory reads per 8-byte entry: that gather is what bounds the kernel).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I node2vec2rank_b200/csrc -o spmm_proto scripts/micro/spmm_proto.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <algorithm>
#include <cstring>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

struct Args {
  int n, l, W, S, n_panels, rows_per_cta, pf;
  const uint32_t* row_ptr;
  const uint2* entries;
  const float* x;
  float* y;
};

__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// One row's cursor: the 8-entry chunk (64-byte aligned in the entry array) the row is in, held two entries per lane, the chunk
// after it (loaded one chunk ahead) and the position inside the current one.  It lives in registers across panels; entries
// past the row's end read as column 0xFFFFFFFF (= stop).
struct Cursor {
  uint32_t k, end, c0, v0, c1, v1, n0, w0, n1, w1;
  int pos;
};
__device__ __forceinline__ void load_next(Cursor& cu, const uint2* entries, int t, int pf) {
  const uint32_t idx = 8u * (cu.k + 1) + 2u * t;
  const uint4 q = ldg128(entries + idx);
  cu.n0 = idx < cu.end ? q.x : 0xFFFFFFFFu; cu.w0 = q.y;
  cu.n1 = idx + 1 < cu.end ? q.z : 0xFFFFFFFFu; cu.w1 = q.w;
  if (pf && t == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(entries + idx + pf));
}

template <int L16, int MODE>   // MODE 0: product; 1: no gather (entries only); 2: gather always from X row 0 of the panel (no conflicts)
__global__ void __launch_bounds__(736, 1) spmm_rows_kernel(const Args a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem);
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t row_bytes = (uint32_t)a.l * 4u;
  const uint32_t stage_bytes = (uint32_t)a.W * row_bytes;
  const int row_base = blockIdx.x * a.rows_per_cta;
  const int rows_here = max(0, min(a.rows_per_cta, a.n - row_base));
  const int n_groups = (rows_here + 1) >> 1;
  const int n_cwarps = (n_groups + 7) >> 3;
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, n_cwarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 22) {
    if (lane == 0 && n_cwarps > 0) {
      int st = 0; uint32_t ph = 0;
      for (int p = 0; p < a.n_panels; ++p) {
        mbar_wait(bar_empty + 8 * st, ph ^ 1u);
        const int rows = min(a.W, a.n - p * a.W);
        mbar_expect_tx(bar_full + 8 * st, rows * row_bytes);
        bulk_g2s(ring + st * stage_bytes, a.x + (size_t)p * a.W * a.l, rows * row_bytes, bar_full + 8 * st);
        if (++st == a.S) { st = 0; ph ^= 1u; }
      }
    }
    return;
  }
  if (warp >= n_cwarps) return;
  const int g = threadIdx.x >> 2, t = lane & 3;
  const uint32_t gmask = 0xFu << (lane & 28);
  const int r0 = row_base + 2 * g;
  Cursor cu[2];
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const bool ok = r0 + rr < row_base + rows_here;
    const uint32_t start = ok ? a.row_ptr[r0 + rr] : 0;
    cu[rr].end = ok ? a.row_ptr[r0 + rr + 1] : 0;
    cu[rr].k = (start >> 3) - 1;
    cu[rr].pos = (int)(start & 7u);
    load_next(cu[rr], a.entries, t, 0);
    cu[rr].c0 = cu[rr].n0; cu[rr].v0 = cu[rr].w0; cu[rr].c1 = cu[rr].n1; cu[rr].v1 = cu[rr].w1;
    ++cu[rr].k;
    load_next(cu[rr], a.entries, t, a.pf);
  }
  float acc[2][4 * L16];
#pragma unroll
  for (int i = 0; i < 4 * L16; ++i) { acc[0][i] = 0.f; acc[1][i] = 0.f; }
  int st = 0; uint32_t ph = 0;
  for (int p = 0; p < a.n_panels; ++p) {
    mbar_wait(bar_full + 8 * st, ph);
    const uint32_t hi = (uint32_t)min(a.n, (p + 1) * a.W);
    const uint32_t xbase = ring + st * stage_bytes - (uint32_t)(p * a.W) * row_bytes + t * 16;
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      Cursor& c_ = cu[rr];
      bool active = true;
      while (true) {                                     // the warp stays in lock step: one entry per group and iteration
        const int src = (lane & 28) | (c_.pos >> 1);
        const uint32_t c = __shfl_sync(0xffffffffu, (c_.pos & 1) ? c_.c1 : c_.c0, src);
        const float v = __uint_as_float(__shfl_sync(0xffffffffu, (c_.pos & 1) ? c_.v1 : c_.v0, src));
        active = active && c < hi;
        if (!__any_sync(0xffffffffu, active)) break;
        if (active) {
          if (MODE != 1) {
            const uint32_t ad = MODE == 2 ? (ring + st * stage_bytes + t * 16) : (xbase + c * row_bytes);
#pragma unroll
            for (int s = 0; s < L16; ++s) {
              const float4 x = lds128(ad + s * 64);
              acc[rr][4 * s + 0] = fmaf(v, x.x, acc[rr][4 * s + 0]);
              acc[rr][4 * s + 1] = fmaf(v, x.y, acc[rr][4 * s + 1]);
              acc[rr][4 * s + 2] = fmaf(v, x.z, acc[rr][4 * s + 2]);
              acc[rr][4 * s + 3] = fmaf(v, x.w, acc[rr][4 * s + 3]);
            }
          } else {
            acc[rr][0] += v * (float)c;
          }
          if (++c_.pos == 8) {
            c_.c0 = c_.n0; c_.v0 = c_.w0; c_.c1 = c_.n1; c_.v1 = c_.w1;
            c_.pos = 0;
            ++c_.k;
            load_next(c_, a.entries, t, a.pf);
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * st);
    if (++st == a.S) { st = 0; ph ^= 1u; }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = r0 + rr;
    if (r < row_base + rows_here) {
#pragma unroll
      for (int s = 0; s < L16; ++s)
        *reinterpret_cast<float4*>(a.y + (size_t)r * a.l + 4 * (t + 4 * s)) = make_float4(acc[rr][4 * s], acc[rr][4 * s + 1], acc[rr][4 * s + 2], acc[rr][4 * s + 3]);
    }
  }
}

// ---- variant B: the entries cross HBM as per-(row block, panel) lists, sorted by row, with a u16 row-pointer header; the
// producer brings one list and one X panel per stage with bulk copies, the groups walk their rows' entries in shared memory.
struct ArgsB {
  int n, l, W, S, n_panels, rb, list_cap;     // list_cap: bytes reserved per stage for a list
  const uint32_t* list_ptr;                   // [n_blocks * n_panels + 1], in 16-byte units
  const uint4* lists;
  const float* x;
  float* y;
};
__device__ __forceinline__ uint2 lds64(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t a) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
  return v;
}

template <int L16, int MODE>
__global__ void __launch_bounds__(736, 1) spmm_lists_kernel(const ArgsB a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = smem_u32(smem);
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t row_bytes = (uint32_t)a.l * 4u;
  const uint32_t x_bytes = (uint32_t)a.W * row_bytes;
  const uint32_t stage_bytes = x_bytes + (uint32_t)a.list_cap;
  const int row_base = blockIdx.x * a.rb;
  const int rows_here = max(0, min(a.rb, a.n - row_base));
  const int n_groups = (rows_here + 1) >> 1;
  const int n_cwarps = (n_groups + 7) >> 3;
  const uint32_t hdr_bytes = (uint32_t)((2 * (a.rb + 1) + 15) & ~15);
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, n_cwarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 22) {
    if (lane == 0 && n_cwarps > 0) {
      int st = 0; uint32_t ph = 0;
      const uint32_t* lp = a.list_ptr + (size_t)blockIdx.x * a.n_panels;
      for (int p = 0; p < a.n_panels; ++p) {
        mbar_wait(bar_empty + 8 * st, ph ^ 1u);
        const int rows = min(a.W, a.n - p * a.W);
        const uint32_t q0 = lp[p], q1 = lp[p + 1];
        mbar_expect_tx(bar_full + 8 * st, rows * row_bytes + (q1 - q0) * 16u);
        bulk_g2s(ring + st * stage_bytes, a.x + (size_t)p * a.W * a.l, rows * row_bytes, bar_full + 8 * st);
        bulk_g2s(ring + st * stage_bytes + x_bytes, a.lists + q0, (q1 - q0) * 16u, bar_full + 8 * st);
        if (++st == a.S) { st = 0; ph ^= 1u; }
      }
    }
    return;
  }
  if (warp >= n_cwarps) return;
  const int g = threadIdx.x >> 2, t = lane & 3;
  const int r0 = row_base + 2 * g;
  float acc[2][4 * L16];
#pragma unroll
  for (int i = 0; i < 4 * L16; ++i) { acc[0][i] = 0.f; acc[1][i] = 0.f; }
  int st = 0; uint32_t ph = 0;
  for (int p = 0; p < a.n_panels; ++p) {
    mbar_wait(bar_full + 8 * st, ph);
    const uint32_t xbase = ring + st * stage_bytes + t * 16;
    const uint32_t hdr = ring + st * stage_bytes + x_bytes;
    const uint32_t ent = hdr + hdr_bytes;
    uint32_t pp[3];
    pp[0] = pp[1] = pp[2] = 0;
    if (g < n_groups) { pp[0] = lds16(hdr + 4 * g); pp[1] = lds16(hdr + 4 * g + 2); pp[2] = 2 * g + 2 <= a.rb ? lds16(hdr + 4 * g + 4) : pp[1]; }
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const uint32_t e1 = pp[rr + 1];
      for (uint32_t e = pp[rr]; __any_sync(0xffffffffu, e < e1); ++e) {
        if (e < e1) {
          const uint2 q = lds64(ent + 8 * e);
          const float v = __uint_as_float(q.y);
          if (MODE != 1) {
            const uint32_t ad = MODE == 2 ? xbase : (xbase + q.x * row_bytes);
#pragma unroll
            for (int s = 0; s < L16; ++s) {
              const float4 x = lds128(ad + s * 64);
              acc[rr][4 * s + 0] = fmaf(v, x.x, acc[rr][4 * s + 0]);
              acc[rr][4 * s + 1] = fmaf(v, x.y, acc[rr][4 * s + 1]);
              acc[rr][4 * s + 2] = fmaf(v, x.z, acc[rr][4 * s + 2]);
              acc[rr][4 * s + 3] = fmaf(v, x.w, acc[rr][4 * s + 3]);
            }
          } else {
            acc[rr][0] += v * (float)q.x;
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * st);
    if (++st == a.S) { st = 0; ph ^= 1u; }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = r0 + rr;
    if (r < row_base + rows_here) {
#pragma unroll
      for (int s = 0; s < L16; ++s)
        *reinterpret_cast<float4*>(a.y + (size_t)r * a.l + 4 * (t + 4 * s)) = make_float4(acc[rr][4 * s], acc[rr][4 * s + 1], acc[rr][4 * s + 2], acc[rr][4 * s + 3]);
    }
  }
}

template <int L16, int MODE>
static float runB(const ArgsB& a, int grid, size_t smem, int reps) {
  cudaFuncSetAttribute(spmm_lists_kernel<L16, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; ++i) spmm_lists_kernel<L16, MODE><<<grid, 736, smem>>>(a);
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) spmm_lists_kernel<L16, MODE><<<grid, 736, smem>>>(a);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms / reps;
}

__global__ void ref_kernel(int n, int l, const uint32_t* row_ptr, const uint2* entries, const float* x, double* y) {
  const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (i >= (long)n * l) return;
  const int r = (int)(i / l), j = (int)(i % l);
  double s = 0.0;
  for (uint32_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) s += (double)__uint_as_float(entries[e].y) * (double)x[(size_t)entries[e].x * l + j];
  y[i] = s;
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static inline double urand() {
  rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
  return (double)((rng_state >> 11) + 1) / 9007199254740993.0;
}

template <int L16, int MODE>
static float run(const Args& a, int grid, size_t smem, int reps) {
  cudaFuncSetAttribute(spmm_rows_kernel<L16, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; ++i) spmm_rows_kernel<L16, MODE><<<grid, 736, smem>>>(a);
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) spmm_rows_kernel<L16, MODE><<<grid, 736, smem>>>(a);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms / reps;
}

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 50000;
  const double density = argc > 2 ? atof(argv[2]) : 0.0264;
  const int W = argc > 3 ? atoi(argv[3]) : 256;
  const int S = argc > 4 ? atoi(argv[4]) : 4;
  const int pf = argc > 5 ? atoi(argv[5]) : 256;
  const int l = 48;
  std::vector<uint32_t> rp(n + 1, 0);
  std::vector<uint2> ent;
  ent.reserve((size_t)(1.05 * density * n * (double)n));
  const double lg = log(1.0 - density);
  for (int r = 0; r < n; ++r) {
    long c = -1;
    while (true) {
      c += 1 + (long)floor(log(urand()) / lg);
      if (c >= n) break;
      const float v = (float)urand();
      uint2 q; q.x = (uint32_t)c; memcpy(&q.y, &v, 4);
      ent.push_back(q);
    }
    rp[r + 1] = (uint32_t)ent.size();
  }
  printf("n=%d entries=%zu (%.4f %%) W=%d S=%d pf=%d\n", n, ent.size(), 100.0 * ent.size() / ((double)n * n), W, S, pf);
  std::vector<float> hx((size_t)n * l);
  for (auto& v : hx) v = (float)(urand() - 0.5);
  uint32_t* d_rp; uint2* d_ent; float *d_x, *d_y; double* d_ref;
  cudaMalloc(&d_rp, (n + 1) * 4); cudaMalloc(&d_ent, (ent.size() + 1024) * 8); cudaMemset(d_ent, 0, (ent.size() + 1024) * 8); cudaMalloc(&d_x, hx.size() * 4); cudaMalloc(&d_y, hx.size() * 4);
  cudaMalloc(&d_ref, hx.size() * 8);
  cudaMemcpy(d_rp, rp.data(), (n + 1) * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_ent, ent.data(), ent.size() * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(d_x, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  Args a;
  a.pf = pf; a.n = n; a.l = l; a.W = W; a.S = S; a.n_panels = (n + W - 1) / W;
  a.rows_per_cta = ((n + sms - 1) / sms + 1) & ~1;
  a.row_ptr = d_rp; a.entries = d_ent; a.x = d_x; a.y = d_y;
  if (a.rows_per_cta > 352) { printf("n too large for one wave\n"); return 1; }
  const int grid = (n + a.rows_per_cta - 1) / a.rows_per_cta;
  const size_t smem = (size_t)S * W * l * 4;
  printf("grid=%d rows_per_cta=%d smem=%zu\n", grid, a.rows_per_cta, smem);
  const float ms0 = run<3, 0>(a, grid, smem, 20);
  cudaError_t err = cudaDeviceSynchronize();
  printf("product: %.4f ms  (%s)\n", ms0, cudaGetErrorString(err));
  ref_kernel<<<(unsigned)(((long)n * l + 255) / 256), 256>>>(n, l, d_rp, d_ent, d_x, d_ref);
  std::vector<float> hy(hx.size());
  std::vector<double> href(hx.size());
  cudaMemcpy(hy.data(), d_y, hy.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(href.data(), d_ref, href.size() * 8, cudaMemcpyDeviceToHost);
  double maxerr = 0, maxref = 0;
  for (size_t i = 0; i < hy.size(); ++i) { maxerr = fmax(maxerr, fabs(hy[i] - href[i])); maxref = fmax(maxref, fabs(href[i])); }
  printf("max abs err %.3e (max |ref| %.3e)\n", maxerr, maxref);
  const float ms1 = run<3, 1>(a, grid, smem, 20);
  printf("entries only (no gather): %.4f ms\n", ms1);
  const float ms2 = run<3, 2>(a, grid, smem, 20);
  printf("gather from one row (no bank conflicts, no address math): %.4f ms\n", ms2);
  {
    // variant B: per-(block, panel) lists
    const int rb = a.rows_per_cta, n_blocks = grid, n_panels = a.n_panels;
    const size_t hdr_u16 = (size_t)((2 * (rb + 1) + 15) & ~15) / 2;
    std::vector<uint32_t> lp((size_t)n_blocks * n_panels + 1, 0);
    std::vector<uint32_t> cur(rp.begin(), rp.end() - 1);      // per row: next entry not yet placed
    std::vector<uint8_t> buf;
    buf.reserve(ent.size() * 8 + (size_t)n_blocks * n_panels * (hdr_u16 * 2 + 16));
    size_t max_list = 0;
    for (int b = 0; b < n_blocks; ++b) {
      for (int p = 0; p < n_panels; ++p) {
        const size_t off = buf.size();
        buf.resize(off + hdr_u16 * 2);
        uint32_t cnt = 0;
        const uint32_t hi = (uint32_t)std::min(n, (p + 1) * W);
        for (int r = 0; r <= rb; ++r) {
          ((uint16_t*)(buf.data() + off))[r] = (uint16_t)cnt;
          const int row = b * rb + r;
          if (r == rb || row >= n) continue;
          while (cur[row] < rp[row + 1] && ent[cur[row]].x < hi) {
            uint2 q = ent[cur[row]++];
            q.x -= (uint32_t)(p * W);
            const size_t o = buf.size();
            buf.resize(o + 8);
            memcpy(buf.data() + o, &q, 8);
            ++cnt;
          }
        }
        while (buf.size() % 16) buf.push_back(0);
        max_list = std::max(max_list, buf.size() - off);
        lp[(size_t)b * n_panels + p + 1] = (uint32_t)(buf.size() / 16);
      }
    }
    printf("lists: %.1f MB (entries %.1f MB), longest list %zu bytes\n", buf.size() / 1e6, ent.size() * 8 / 1e6, max_list);
    uint32_t* d_lp; uint4* d_lists;
    cudaMalloc(&d_lp, lp.size() * 4); cudaMalloc(&d_lists, buf.size());
    cudaMemcpy(d_lp, lp.data(), lp.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_lists, buf.data(), buf.size(), cudaMemcpyHostToDevice);
    ArgsB ab;
    ab.n = n; ab.l = l; ab.W = W; ab.n_panels = n_panels; ab.rb = rb; ab.list_cap = (int)((max_list + 127) & ~(size_t)127);
    ab.S = (int)((227 * 1024 - 2048) / ((size_t)W * l * 4 + ab.list_cap));
    if (ab.S > 8) ab.S = 8;
    ab.list_ptr = d_lp; ab.lists = d_lists; ab.x = d_x; ab.y = d_y;
    const size_t smemB = (size_t)ab.S * ((size_t)W * l * 4 + ab.list_cap);
    cudaMemset(d_y, 0, hx.size() * 4);
    const float b0 = runB<3, 0>(ab, grid, smemB, 20);
    err = cudaDeviceSynchronize();
    cudaMemcpy(hy.data(), d_y, hy.size() * 4, cudaMemcpyDeviceToHost);
    maxerr = 0;
    for (size_t i = 0; i < hy.size(); ++i) maxerr = fmax(maxerr, fabs(hy[i] - href[i]));
    printf("B (lists in smem, S=%d): product %.4f ms (%s), max abs err %.3e\n", ab.S, b0, cudaGetErrorString(err), maxerr);
    printf("B entries only: %.4f ms\n", runB<3, 1>(ab, grid, smemB, 20));
    printf("B gather from one row: %.4f ms\n", runB<3, 2>(ab, grid, smemB, 20));
    int khzb = 0;
    cudaDeviceGetAttribute(&khzb, cudaDevAttrClockRate, 0);
    printf("B: %.2f SM-clk per entry\n", b0 * 1e-3 * khzb * 1e3 * sms / (double)ent.size());
  }
  double clk = 0; int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  clk = ms0 * 1e-3 * khz * 1e3 * sms / (double)ent.size();
  printf("%.2f SM-clk per entry at %d MHz (ideal %.2f)\n", clk, khz / 1000, l * 4 / 128.0);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
