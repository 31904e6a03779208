// K1d / K3 for ROW-LIST layers: the storage and the product of a SPARSE layer on the CUDA cores.
//
// After the threshold / top-percent cut of preprocessing_utils.py:131-139 a C3 layer is 2.6 % dense.  The compressed-tile
// kernels (k3_sparse*.cu) expand such a layer into dense bf16 tiles and pay the dense tile MMA work: 38 x the arithmetic the
// non-zeros need, which the tensor cores' 31 x advantage only just pays back.  Here the product is done the sparse way, in
// exact fp32: Y[r, :] += a_rc X[c, :] for every stored entry, on the FMA pipes.  What bounds it is not HBM and not the
// arithmetic but the GATHER of the X rows (l x 4 bytes of shared-memory reads per 8-byte entry), so the layout is built
// around that gather:
//   * rows are cut into blocks of `rb` rows (one CTA each: rb ~ n / #SMs), columns into panels of `W`;
//   * the non-zeros of (block b, panel p) form one LIST: a u16 row-pointer header (rb + 1 offsets in PAIRS of entries, padded to
//     16 bytes; every row is padded to whole pairs so that a group reads two entries with one 16-byte load) and
//     the 8-byte entries {column inside the panel, fp32 value} sorted by row (inside a row: one column parity first, see
//     rows_fill_kernel), the list padded to 16 bytes;
//   * list_ptr[b n_panels + p] is the offset of the list in 16-byte units.
// One producer thread brings, per stage of a ring, the X panel (W x l fp32: one bulk copy of the contiguous rows) and the list
// of the CTA's block (one bulk copy); groups of 4 lanes own RPG consecutive rows each, keep their l accumulators per row in
// registers for the WHOLE pass (no partial sums, no reduction kernel, no operand split), read their rows' entries from the
// list in shared memory (broadcast) and gather the X row as l/16 16-byte loads per lane.  A warp walks row j of its 8 groups in
// step (a quarter warp whose two groups are both done costs no shared-memory cycle).  Walking the RPG rows of a group as ONE
// sequence with the accumulator chosen by predication was measured and lost (0.83 ms against 0.69 ms at C3): the extra FMAs
// cost more issue slots than the idle steps they save.
// Also measured and rejected: loading the next pair of entries while the current one is multiplied and issuing both gathers
// of a pair before the first FMA (software pipelining): 0.666 ms against 0.639 ms -- the kernel sits on the shared-memory
// wavefronts and the issue slots (ncu: 89 % / 52 %), not on latency, and the extra registers and moves cost more than they hide.
// Row blocks of half the size with two CTAs per SM (one row per lane group, 44 consumer warps per SM; taken when a layer is built
// with more row blocks than SMs) gain 3 % at l = 32 (0.451 against 0.463 ms) and lose at l = 48, where two CTAs' stages no longer
// fit: the builder's default stays one block per SM.
// The values are the layer's fp32 values: no bf16 plane split, P = 3 layers cost what P = 1 layers cost.
#include "sparse_common.cuh"
#include <algorithm>

namespace n2v2r {

static int g_rows_rpg = 0;       // debug: force the rows per group (0 = heuristic)
constexpr uint32_t kRowsPad = 0xFFFFFFFFu;   // column of the padding entry that completes the last 16-byte pair of an odd row
static int g_rows_lanes8 = 1;       // debug: 0 = groups of 4 lanes also for l = 64; 2 = groups of 8 lanes also for l = 32
static int g_rows_two_per_sm = 1;   // debug: 0 = never two CTAs per SM
static int g_rows_parity = 1;    // debug: 0 = the builder writes every row's entries in plain column order

constexpr int kRowsWarps = 22;                   // consumer warps of the product kernel; warp kRowsWarps is the producer
constexpr int kRowsThreads = 32 * (kRowsWarps + 1);
// rows owned by one group of 4 lanes (consecutive rows): the builder orders the entries by it, the product kernel is compiled for it
static inline int rows_per_group(int rb) { return (rb + 8 * kRowsWarps - 1) / (8 * kRowsWarps); }

// ---------------------------------------------------------------------------------------------------------------------
// builder
// ---------------------------------------------------------------------------------------------------------------------
// counts[(b n_panels + p) rb + r] = non-zeros of row b rb + r in the columns of panel p.  One CTA per row, a warp per panel.
__global__ void rows_count_kernel(const float* __restrict__ in, int64_t n, int64_t ld, int rb, int W, int n_panels,
                                  uint16_t* __restrict__ counts) {
  const int64_t row = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int b = (int)(row / rb), r = (int)(row - (int64_t)b * rb);
  const float* src = in + row * ld;
  for (int p = warp; p < n_panels; p += nw) {
    int cnt = 0;
    for (int c = lane; c < W; c += 32) {
      const int64_t col = (int64_t)p * W + c;
      const float v = col < n ? __ldg(src + col) : 0.f;
      cnt += __popc(__ballot_sync(0xffffffffu, v != 0.f));
    }
    if (lane == 0) counts[((size_t)b * n_panels + p) * rb + r] = (uint16_t)cnt;
  }
}

// units[i] = size of list i in 16-byte units (header + entries)
__global__ void rows_sizes_kernel(const uint16_t* __restrict__ counts, int64_t n_lists, int rb, uint32_t hdr_units,
                                  uint32_t* __restrict__ units, unsigned long long* __restrict__ nnz) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  uint32_t cnt = 0;
  if (i < n_lists) {
    const uint16_t* c = counts + (size_t)i * rb;
    uint32_t pairs = 0;
    for (int r = 0; r < rb; ++r) { cnt += c[r]; pairs += (c[r] + 1u) >> 1; }        // every row is padded to whole 16-byte pairs
    units[i] = hdr_units + pairs;
  }
  // non-zeros in total (one atomic per warp)
  for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(nnz, (unsigned long long)cnt);
}

// exclusive scan of units -> list_ptr (n_lists + 1), and the largest list; one CTA
__global__ void rows_scan_kernel(const uint32_t* __restrict__ units, int64_t n_lists, uint32_t* __restrict__ list_ptr,
                                 uint32_t* __restrict__ max_units) {
  __shared__ uint32_t warp_sum[32];
  __shared__ uint32_t carry_s, max_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { carry_s = 0; max_s = 0; }
  __syncthreads();
  uint32_t mx = 0;
  for (int64_t base = 0; base < n_lists; base += blockDim.x) {
    const int64_t i = base + threadIdx.x;
    const uint32_t v = i < n_lists ? units[i] : 0u;
    mx = max(mx, v);
    uint32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = lane < (int)(blockDim.x >> 5) ? warp_sum[lane] : 0u;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_sum[lane] = w;        // inclusive over warps
    }
    __syncthreads();
    const uint32_t before = carry_s + (warp ? warp_sum[warp - 1] : 0u) + incl - v;
    if (i < n_lists) list_ptr[i] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry_s = before + v;
    __syncthreads();
  }
  atomicMax(&max_s, mx);
  __syncthreads();
  if (threadIdx.x == 0) { list_ptr[n_lists] = carry_s; *max_units = max_s; }
}

// one CTA per list: header from the counts, then a warp per row writes its entries in column order
__global__ void rows_fill_kernel(const float* __restrict__ in, int64_t n, int64_t ld, int rb, int W, int n_panels,
                                 const uint16_t* __restrict__ counts, const uint32_t* __restrict__ list_ptr, uint4* __restrict__ lists,
                                 int rpg, int parity_order) {
  extern __shared__ uint32_t ptr_s[];            // rb + 1
  const int64_t li = blockIdx.x;
  const int b = (int)(li / n_panels), p = (int)(li - (int64_t)b * n_panels);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const uint16_t* c = counts + (size_t)li * rb;
  if (warp == 0) {                               // exclusive scan of rb counts by one warp
    uint32_t carry = 0;
    for (int base = 0; base < rb; base += 32) {
      const int r = base + lane;
      const uint32_t v = r < rb ? ((uint32_t)c[r] + 1u) >> 1 : 0u;          // pairs of entries
      uint32_t incl = v;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      if (r < rb) ptr_s[r] = carry + incl - v;
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) ptr_s[rb] = carry;
  }
  __syncthreads();
  uint8_t* list = (uint8_t*)(lists + list_ptr[li]);
  uint16_t* hdr = (uint16_t*)list;
  const uint32_t hdr_bytes = (uint32_t)((2 * (rb + 1) + 15) & ~15);
  for (int r = threadIdx.x; r < (int)(hdr_bytes / 2); r += blockDim.x) hdr[r] = (uint16_t)(r <= rb ? ptr_s[r] : 0u);
  uint2* ent = (uint2*)(list + hdr_bytes);
  for (int r = warp; r < rb; r += nw) {
    const int64_t row = (int64_t)b * rb + r;
    if (row >= n) break;
    const float* src = in + row * ld;
    const uint32_t at0 = 2u * ptr_s[r];
    uint32_t at = at0;
    if (lane == 0 && (c[r] & 1u)) ent[at0 + c[r]] = make_uint2(kRowsPad, 0u);      // odd row: the second half of its last pair
    // Two lane groups share every 128-byte phase of the product kernel's gather (lanes 8q .. 8q + 7 = groups 2q, 2q + 1), and
    // the X rows of two columns of the same parity start in the same half of the banks (row stride l x 4 bytes = an odd
    // number of 64-byte halves for l = 16, 48, 80): a 2-way bank conflict.  So the rows of even groups list their columns
    // as even, odd, even, odd, ... and the rows of odd groups as odd, even, odd, even, ... (ascending inside each parity
    // class; the surplus of the larger class follows): in step i the two groups of a phase then gather from opposite halves
    // as long as both rows still have columns of both parities.  (l = 32, 64: the kernel swaps the chunk order instead.)
    if (!parity_order) {
      for (int c0 = 0; c0 < W; c0 += 32) {
        const int64_t col = (int64_t)p * W + c0 + lane;
        const float v = col < n ? __ldg(src + col) : 0.f;
        const uint32_t m = __ballot_sync(0xffffffffu, v != 0.f);
        if (v != 0.f) ent[at + __popc(m & ((1u << lane) - 1u))] = make_uint2((uint32_t)(c0 + lane), __float_as_uint(v));
        at += __popc(m);
      }
      continue;
    }
    const uint32_t first = (uint32_t)((r / rpg) & 1);                  // parity of the columns that open the row
    const uint32_t mask_first = first ? 0xAAAAAAAAu : 0x55555555u;      // c0 is even: parity of the column = parity of the lane
    uint32_t n_cls[2] = {0u, 0u};
    for (int c0 = 0; c0 < W; c0 += 32) {
      const int64_t col = (int64_t)p * W + c0 + lane;
      const float v = col < n ? __ldg(src + col) : 0.f;
      const uint32_t m = __ballot_sync(0xffffffffu, v != 0.f);
      n_cls[0] += __popc(m & mask_first);
      n_cls[1] += __popc(m & ~mask_first);
    }
    const uint32_t paired = min(n_cls[0], n_cls[1]);
    uint32_t k_cls[2] = {0u, 0u};
    for (int c0 = 0; c0 < W; c0 += 32) {
      const int64_t col = (int64_t)p * W + c0 + lane;
      const float v = col < n ? __ldg(src + col) : 0.f;
      const uint32_t m = __ballot_sync(0xffffffffu, v != 0.f);
      const uint32_t cls = (((uint32_t)lane & 1u) == first) ? 0u : 1u;
      const uint32_t mc = cls ? (m & ~mask_first) : (m & mask_first);
      if (v != 0.f) {
        const uint32_t k = (cls ? k_cls[1] : k_cls[0]) + __popc(mc & ((1u << lane) - 1u));
        const uint32_t pos = k < paired ? 2u * k + cls : paired + k;
        ent[at + pos] = make_uint2((uint32_t)(c0 + lane), __float_as_uint(v));
      }
      k_cls[0] += __popc(m & mask_first);
      k_cls[1] += __popc(m & ~mask_first);
    }
  }
}

__global__ void rows_to_dense_kernel(const uint32_t* __restrict__ list_ptr, const uint4* __restrict__ lists, int64_t n, int rb, int W,
                                     int n_panels, float* __restrict__ out, int64_t ld_out) {
  const int64_t li = blockIdx.x;
  const int b = (int)(li / n_panels), p = (int)(li - (int64_t)b * n_panels);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const uint8_t* list = (const uint8_t*)(lists + list_ptr[li]);
  const uint16_t* hdr = (const uint16_t*)list;
  const uint2* ent = (const uint2*)(list + ((2 * (rb + 1) + 15) & ~15));
  for (int r = warp; r < rb; r += nw) {
    const int64_t row = (int64_t)b * rb + r;
    if (row >= n) break;
    for (uint32_t e = 2u * hdr[r] + lane; e < 2u * hdr[r + 1]; e += 32) {
      const uint2 q = ent[e];
      if (q.x != kRowsPad) out[row * ld_out + (int64_t)p * W + q.x] = __uint_as_float(q.y);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// product
// ---------------------------------------------------------------------------------------------------------------------
struct RowsArgs {
  int64_t n;
  int l, W, S, n_panels, rb, list_cap, accumulate;
  const uint32_t* list_ptr;
  const uint4* lists;
  const float* x;
  float* y;
};

__device__ __forceinline__ float4 lds128f(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

template <int L16, int RPG>
__global__ void __launch_bounds__(kRowsThreads, (RPG == 1 && L16 <= 3) ? 2 : 1) rows_product_kernel(const RowsArgs a) {
  extern __shared__ __align__(128) uint8_t smem_rows[];
  __shared__ __align__(8) uint64_t bars[16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_rows) + 127u) & ~127u;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t row_bytes = (uint32_t)a.l * 4u;
  const uint32_t x_bytes = (uint32_t)a.W * row_bytes;
  const uint32_t stage_bytes = x_bytes + (uint32_t)a.list_cap;
  const int64_t row_base = (int64_t)blockIdx.x * a.rb;
  const int rows_here = (int)max((int64_t)0, min((int64_t)a.rb, a.n - row_base));
  const int n_groups = (rows_here + RPG - 1) / RPG;
  const int n_cwarps = (n_groups + 7) >> 3;
  const uint32_t hdr_bytes = (uint32_t)((2 * (a.rb + 1) + 15) & ~15);
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, n_cwarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == kRowsWarps) {
    // ===== producer: X panel + this block's list per stage =====
    if (lane == 0 && n_cwarps > 0) {
      int st = 0; uint32_t ph = 0;
      const uint32_t* lp = a.list_ptr + (size_t)blockIdx.x * a.n_panels;
      uint32_t q0 = __ldg(lp);
      for (int p = 0; p < a.n_panels; ++p) {
        const uint32_t q1 = __ldg(lp + p + 1);
        mbar_wait(bar_empty + 8 * st, ph ^ 1u);
        const uint32_t rows = (uint32_t)min((int64_t)a.W, a.n - (int64_t)p * a.W);
        mbar_expect_tx(bar_full + 8 * st, rows * row_bytes + (q1 - q0) * 16u);
        bulk_copy_g2s(ring + st * stage_bytes, a.x + (size_t)p * a.W * a.l, rows * row_bytes, bar_full + 8 * st);
        bulk_copy_g2s(ring + st * stage_bytes + x_bytes, a.lists + q0, (q1 - q0) * 16u, bar_full + 8 * st);
        q0 = q1;
        if (++st == a.S) { st = 0; ph ^= 1u; }
      }
    }
    return;
  }
  if (warp >= n_cwarps) return;
  // ===== consumers: a group of 4 lanes owns RPG consecutive rows; lane t holds columns 4 (t + 4 s) .. + 3, s < L16 =====
  const int g = threadIdx.x >> 2, t = lane & 3;
  const bool have = g < n_groups;
  // l = 32, 64: every X row starts in the same half of the banks, so the odd group of a phase reads its 64-byte chunks in
  // swapped order (1, 0, 3, 2) and the two groups never meet in the same half; l = 16, 48, 80: see the builder's entry order
  const uint32_t swap = (L16 % 2 == 0) ? (uint32_t)(g & 1) : 0u;
  float acc[RPG][4 * L16];
#pragma unroll
  for (int j = 0; j < RPG; ++j)
#pragma unroll
    for (int i = 0; i < 4 * L16; ++i) acc[j][i] = 0.f;
  int st = 0; uint32_t ph = 0;
  for (int p = 0; p < a.n_panels; ++p) {
    mbar_wait(bar_full + 8 * st, ph);
    const uint32_t xbase = ring + st * stage_bytes + t * 16;
    const uint32_t hdr = ring + st * stage_bytes + x_bytes;
    const uint32_t ent = hdr + hdr_bytes;
    uint32_t pp[RPG + 1];
#pragma unroll
    for (int j = 0; j <= RPG; ++j) pp[j] = have ? ld_shared_u16(hdr + 2u * (uint32_t)min(RPG * g + j, a.rb)) : 0u;
#pragma unroll
    for (int j = 0; j < RPG; ++j) {
      // one row after the other (the accumulators are registers: the row index must be static); the warp walks row j of its 8
      // groups in step, a group whose row is exhausted idles until the longest row of the warp is done
      for (uint32_t e = pp[j]; __any_sync(0xffffffffu, e < pp[j + 1]); ++e) {      // e counts 16-byte pairs of entries
        if (e < pp[j + 1]) {
          const uint4 q = ld_shared_v4(ent + 16u * e);
          {
            const float v = __uint_as_float(q.y);
            const uint32_t ad = xbase + q.x * row_bytes;
#pragma unroll
            for (int s = 0; s < L16; ++s) {
              const float4 x = lds128f(ad + (((uint32_t)s ^ swap) << 6));
              acc[j][4 * s + 0] = fmaf(v, x.x, acc[j][4 * s + 0]);
              acc[j][4 * s + 1] = fmaf(v, x.y, acc[j][4 * s + 1]);
              acc[j][4 * s + 2] = fmaf(v, x.z, acc[j][4 * s + 2]);
              acc[j][4 * s + 3] = fmaf(v, x.w, acc[j][4 * s + 3]);
            }
          }
          if (q.z != kRowsPad) {
            const float v = __uint_as_float(q.w);
            const uint32_t ad = xbase + q.z * row_bytes;
#pragma unroll
            for (int s = 0; s < L16; ++s) {
              const float4 x = lds128f(ad + (((uint32_t)s ^ swap) << 6));
              acc[j][4 * s + 0] = fmaf(v, x.x, acc[j][4 * s + 0]);
              acc[j][4 * s + 1] = fmaf(v, x.y, acc[j][4 * s + 1]);
              acc[j][4 * s + 2] = fmaf(v, x.z, acc[j][4 * s + 2]);
              acc[j][4 * s + 3] = fmaf(v, x.w, acc[j][4 * s + 3]);
            }
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * st);
    if (++st == a.S) { st = 0; ph ^= 1u; }
  }
#pragma unroll
  for (int j = 0; j < RPG; ++j) {
    const int r = RPG * g + j;
    if (have && r < rows_here) {
      float* dst = a.y + (size_t)(row_base + r) * a.l + 4 * t;
#pragma unroll
      for (int s = 0; s < L16; ++s) {
        float4 v = make_float4(acc[j][4 * s], acc[j][4 * s + 1], acc[j][4 * s + 2], acc[j][4 * s + 3]);
        float* d = dst + 16 * ((uint32_t)s ^ swap);
        if (a.accumulate) {
          const float4 o = *reinterpret_cast<const float4*>(d);
          v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
        }
        *reinterpret_cast<float4*>(d) = v;
      }
    }
  }
}

// ---- the same product with groups of EIGHT lanes, for l = 64 (and, behind a debug switch, 32) -------------------------
// A row of X is then one (l = 32) or two (l = 64) whole 128-byte wavefronts, read by one group alone: no pairing of groups
// inside a wavefront, hence no half-empty wavefronts when the partner's row is shorter and no bank conflicts whatever the
// columns are.  A warp holds 4 groups, a group owns RPG <= 4 consecutive rows (4 NCH accumulators per row and lane, lane t
// holds columns 4 t .. 4 t + 3 of every 32-column chunk).  Same lists, same producer.
template <int NCH, int RPG>     // NCH = l / 32
__global__ void __launch_bounds__(kRowsThreads, 1) rows_product8_kernel(const RowsArgs a) {
  extern __shared__ __align__(128) uint8_t smem_rows[];
  __shared__ __align__(8) uint64_t bars[16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_rows) + 127u) & ~127u;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[8]);
  const uint32_t row_bytes = (uint32_t)a.l * 4u;
  const uint32_t x_bytes = (uint32_t)a.W * row_bytes;
  const uint32_t stage_bytes = x_bytes + (uint32_t)a.list_cap;
  const int64_t row_base = (int64_t)blockIdx.x * a.rb;
  const int rows_here = (int)max((int64_t)0, min((int64_t)a.rb, a.n - row_base));
  const int n_groups = (rows_here + RPG - 1) / RPG;
  const int n_cwarps = (n_groups + 3) >> 2;
  const uint32_t hdr_bytes = (uint32_t)((2 * (a.rb + 1) + 15) & ~15);
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, n_cwarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == kRowsWarps) {
    if (lane == 0 && n_cwarps > 0) {
      int st = 0; uint32_t ph = 0;
      const uint32_t* lp = a.list_ptr + (size_t)blockIdx.x * a.n_panels;
      uint32_t q0 = __ldg(lp);
      for (int p = 0; p < a.n_panels; ++p) {
        const uint32_t q1 = __ldg(lp + p + 1);
        mbar_wait(bar_empty + 8 * st, ph ^ 1u);
        const uint32_t rows = (uint32_t)min((int64_t)a.W, a.n - (int64_t)p * a.W);
        mbar_expect_tx(bar_full + 8 * st, rows * row_bytes + (q1 - q0) * 16u);
        bulk_copy_g2s(ring + st * stage_bytes, a.x + (size_t)p * a.W * a.l, rows * row_bytes, bar_full + 8 * st);
        bulk_copy_g2s(ring + st * stage_bytes + x_bytes, a.lists + q0, (q1 - q0) * 16u, bar_full + 8 * st);
        q0 = q1;
        if (++st == a.S) { st = 0; ph ^= 1u; }
      }
    }
    return;
  }
  if (warp >= n_cwarps) return;
  const int g = threadIdx.x >> 3, t = lane & 7;
  const bool have = g < n_groups;
  float acc[RPG][4 * NCH];
#pragma unroll
  for (int j = 0; j < RPG; ++j)
#pragma unroll
    for (int i = 0; i < 4 * NCH; ++i) acc[j][i] = 0.f;
  int st = 0; uint32_t ph = 0;
  for (int p = 0; p < a.n_panels; ++p) {
    mbar_wait(bar_full + 8 * st, ph);
    const uint32_t xbase = ring + st * stage_bytes + t * 16;
    const uint32_t hdr = ring + st * stage_bytes + x_bytes;
    const uint32_t ent = hdr + hdr_bytes;
    uint32_t pp[RPG + 1];
#pragma unroll
    for (int j = 0; j <= RPG; ++j) pp[j] = have ? ld_shared_u16(hdr + 2u * (uint32_t)min(RPG * g + j, a.rb)) : 0u;
#pragma unroll
    for (int j = 0; j < RPG; ++j) {
      for (uint32_t e = pp[j]; __any_sync(0xffffffffu, e < pp[j + 1]); ++e) {      // e counts 16-byte pairs of entries
        if (e < pp[j + 1]) {
          const uint4 q = ld_shared_v4(ent + 16u * e);
          {
            const float v = __uint_as_float(q.y);
            const uint32_t ad = xbase + q.x * row_bytes;
#pragma unroll
            for (int s = 0; s < NCH; ++s) {
              const float4 x = lds128f(ad + s * 128);
              acc[j][4 * s + 0] = fmaf(v, x.x, acc[j][4 * s + 0]);
              acc[j][4 * s + 1] = fmaf(v, x.y, acc[j][4 * s + 1]);
              acc[j][4 * s + 2] = fmaf(v, x.z, acc[j][4 * s + 2]);
              acc[j][4 * s + 3] = fmaf(v, x.w, acc[j][4 * s + 3]);
            }
          }
          if (q.z != kRowsPad) {
            const float v = __uint_as_float(q.w);
            const uint32_t ad = xbase + q.z * row_bytes;
#pragma unroll
            for (int s = 0; s < NCH; ++s) {
              const float4 x = lds128f(ad + s * 128);
              acc[j][4 * s + 0] = fmaf(v, x.x, acc[j][4 * s + 0]);
              acc[j][4 * s + 1] = fmaf(v, x.y, acc[j][4 * s + 1]);
              acc[j][4 * s + 2] = fmaf(v, x.z, acc[j][4 * s + 2]);
              acc[j][4 * s + 3] = fmaf(v, x.w, acc[j][4 * s + 3]);
            }
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * st);
    if (++st == a.S) { st = 0; ph ^= 1u; }
  }
#pragma unroll
  for (int j = 0; j < RPG; ++j) {
    const int r = RPG * g + j;
    if (have && r < rows_here) {
      float* dst = a.y + (size_t)(row_base + r) * a.l + 4 * t;
#pragma unroll
      for (int s = 0; s < NCH; ++s) {
        float4 v = make_float4(acc[j][4 * s], acc[j][4 * s + 1], acc[j][4 * s + 2], acc[j][4 * s + 3]);
        float* d = dst + 32 * s;
        if (a.accumulate) {
          const float4 o = *reinterpret_cast<const float4*>(d);
          v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
        }
        *reinterpret_cast<float4*>(d) = v;
      }
    }
  }
}

template <int NCH, int RPG>
static int launch_rows8(cudaStream_t s, const RowsArgs& a, int grid, size_t smem) {
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(rows_product8_kernel<NCH, RPG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024)));
    attr_set = true;
  }
  rows_product8_kernel<NCH, RPG><<<grid, kRowsThreads, smem, s>>>(a);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

template <int NCH>
static int launch_rows8_r(cudaStream_t s, const RowsArgs& a, int grid, size_t smem, int rpg) {
  switch (rpg) {
    case 1: return launch_rows8<NCH, 1>(s, a, grid, smem);
    case 2: return launch_rows8<NCH, 2>(s, a, grid, smem);
    case 3: return launch_rows8<NCH, 3>(s, a, grid, smem);
    default: return launch_rows8<NCH, 4>(s, a, grid, smem);
  }
}

template <int L16, int RPG>
static int launch_rows(cudaStream_t s, const RowsArgs& a, int grid, size_t smem) {
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(rows_product_kernel<L16, RPG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024)));
    attr_set = true;
  }
  rows_product_kernel<L16, RPG><<<grid, kRowsThreads, smem, s>>>(a);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

template <int RPG>
static int launch_rows_l(cudaStream_t s, const RowsArgs& a, int grid, size_t smem) {
  switch (a.l / 16) {
    case 1: return launch_rows<1, RPG>(s, a, grid, smem);
    case 2: return launch_rows<2, RPG>(s, a, grid, smem);
    case 3: return launch_rows<3, RPG>(s, a, grid, smem);
    case 4: return launch_rows<4, RPG>(s, a, grid, smem);
    case 5: return launch_rows<5, RPG>(s, a, grid, smem);
    default: set_error("row lists: block width l must be 16, 32, 48, 64 or 80"); return N2V2R_EINVAL;
  }
}

// d_out (n x l) = (accumulate ? d_out : 0) + A X for a layer stored as row lists
int stream_apply_rows(cudaStream_t s, int64_t n, int l, const uint32_t* list_ptr, const void* lists, int rb, int W, int list_cap,
                      const float* x, float* out, int accumulate) {
  N2V2R_REQUIRE(list_ptr && lists && x && out, "row lists: null pointer");
  N2V2R_REQUIRE(rb >= 1 && rb <= 8 * kRowsWarps * 3 && W >= 32 && W % 32 == 0 && list_cap > 0 && list_cap % 16 == 0,
                "row lists: rows per block must be 1..%d, the panel width a multiple of 32", 8 * kRowsWarps * 3);
  N2V2R_REQUIRE(l % 16 == 0 && l >= 16 && l <= 80, "row lists: block width l must be 16, 32, 48, 64 or 80");
  RowsArgs a;
  a.n = n; a.l = l; a.W = W; a.rb = rb; a.accumulate = accumulate;
  a.n_panels = (int)((n + W - 1) / W);
  a.list_cap = (list_cap + 127) & ~127;
  a.list_ptr = list_ptr; a.lists = (const uint4*)lists; a.x = x; a.y = out;
  const size_t stage = (size_t)W * l * 4 + a.list_cap;
  const int grid = (int)((n + rb - 1) / rb);
  int rpg = rows_per_group(rb);
  if (g_rows_rpg >= rpg && g_rows_rpg <= 3) rpg = g_rows_rpg;
  // more row blocks than SMs, one row per lane group and l <= 48: two CTAs share an SM (44 consumer warps instead of 22 hide
  // the latency of the two dependent shared-memory reads per entry), each with half of the shared memory
  const bool two_per_sm = g_rows_two_per_sm && rpg == 1 && l <= 48 && grid > sm_count() && 2 * stage + 128 <= 112 * 1024;
  const size_t budget = two_per_sm ? (size_t)112 * 1024 - 128 : (size_t)227 * 1024 - 1024 - 256;
  a.S = (int)std::min<size_t>(8, budget / stage);
  // a layer built for a narrower block (wider panels) still works as long as ONE stage fits: copy and compute then alternate
  N2V2R_REQUIRE(a.S >= 1, "row lists: a stage of %zu bytes (X panel of %d x %d fp32 + the longest list, %d bytes) does not fit in shared memory", stage, W, l,
                list_cap);
  const size_t smem = a.S * stage + 128;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int rc = mma_events_begin(s, &e0, &e1);
  if (rc) return rc;
  const int rpg8 = (rb + 4 * kRowsWarps - 1) / (4 * kRowsWarps);       // rows per group of 8 lanes
  // measured at C3: l = 64 0.783 ms against 0.821 ms with groups of 4 lanes; l = 32 0.503 against 0.463 ms (half as many entries
  // per warp step: the loop overhead outweighs the full wavefronts), so only l = 64 takes it unless the debug switch says 2
  if (g_rows_lanes8 && (l == 64 || (l == 32 && g_rows_lanes8 == 2)) && rpg8 <= 4 && !two_per_sm)
    rc = l == 32 ? launch_rows8_r<1>(s, a, grid, smem, rpg8) : launch_rows8_r<2>(s, a, grid, smem, rpg8);
  else if (rpg == 1) rc = launch_rows_l<1>(s, a, grid, smem);
  else if (rpg == 2) rc = launch_rows_l<2>(s, a, grid, smem);
  else rc = launch_rows_l<3>(s, a, grid, smem);
  if (rc) return rc;
  return mma_events_end(s, e0, e1);
}

}  // namespace n2v2r

using namespace n2v2r;

extern "C" void n2v2r__debug_rows_rpg(int rpg) { g_rows_rpg = rpg; }
extern "C" void n2v2r__debug_rows_parity(int on) { g_rows_parity = on; }
extern "C" void n2v2r__debug_rows_two_per_sm(int on) { g_rows_two_per_sm = on; }
extern "C" void n2v2r__debug_rows_lanes8(int on) { g_rows_lanes8 = on; }

extern "C" size_t n2v2r_rows_counts_bytes(int64_t n, int rows_per_block, int panel_cols) {
  if (n <= 0 || rows_per_block <= 0 || panel_cols <= 0) return 0;
  const int64_t nb = (n + rows_per_block - 1) / rows_per_block, np = (n + panel_cols - 1) / panel_cols;
  return (size_t)round_up(nb * np * rows_per_block * 2, 16) + (size_t)(nb * np) * 4 + 64;
}

extern "C" int n2v2r_rows_count(void* stream, const float* d_in, int64_t n, int64_t ld, int rows_per_block, int panel_cols,
                                void* d_counts, uint32_t* d_list_ptr, int64_t* h_units, int32_t* h_max_list_bytes, int64_t* h_nnz) {
  N2V2R_REQUIRE(d_in && d_counts && d_list_ptr && h_units && h_max_list_bytes && h_nnz && n > 0 && ld >= n, "rows_count: bad arguments");
  N2V2R_REQUIRE(rows_per_block >= 1 && rows_per_block <= 8 * kRowsWarps * 3 && panel_cols >= 32 && panel_cols % 32 == 0 && panel_cols <= 4096,
                "rows_count: rows per block 1..%d, panel width a multiple of 32 up to 4096", 8 * kRowsWarps * 3);
  cudaStream_t s = (cudaStream_t)stream;
  const int rb = rows_per_block, W = panel_cols;
  const int64_t nb = (n + rb - 1) / rb, np = (n + W - 1) / W, n_lists = nb * np;
  uint16_t* counts = (uint16_t*)d_counts;
  uint32_t* units = (uint32_t*)((uint8_t*)d_counts + (size_t)round_up(n_lists * rb * 2, 16));
  unsigned long long* nnz = (unsigned long long*)(units + n_lists + (n_lists & 1));
  uint32_t* max_units = (uint32_t*)(nnz + 1);
  N2V2R_CUDA(cudaMemsetAsync(d_counts, 0, (size_t)round_up(n_lists * rb * 2, 16) + (size_t)n_lists * 4 + 64, s));
  rows_count_kernel<<<(unsigned)n, 256, 0, s>>>(d_in, n, ld, rb, W, (int)np, counts);
  N2V2R_CHECK_LAUNCH();
  const uint32_t hdr_units = (uint32_t)(((2 * (rb + 1) + 15) & ~15) / 16);
  rows_sizes_kernel<<<(unsigned)((n_lists + 255) / 256), 256, 0, s>>>(counts, n_lists, rb, hdr_units, units, nnz);
  N2V2R_CHECK_LAUNCH();
  rows_scan_kernel<<<1, 1024, 0, s>>>(units, n_lists, d_list_ptr, max_units);
  N2V2R_CHECK_LAUNCH();
  uint32_t total = 0, mx = 0;
  unsigned long long nz = 0;
  N2V2R_CUDA(cudaMemcpyAsync(&total, d_list_ptr + n_lists, 4, cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaMemcpyAsync(&mx, max_units, 4, cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaMemcpyAsync(&nz, nnz, 8, cudaMemcpyDeviceToHost, s));
  N2V2R_CUDA(cudaStreamSynchronize(s));
  N2V2R_REQUIRE(mx - hdr_units <= 65535u, "rows_count: a (block, panel) list holds more than 65535 entry pairs: the layer is too dense for row lists");
  *h_units = (int64_t)total;
  *h_max_list_bytes = (int32_t)(mx * 16u);
  *h_nnz = (int64_t)nz;
  return N2V2R_OK;
}

extern "C" int n2v2r_rows_fill(void* stream, const float* d_in, int64_t n, int64_t ld, int rows_per_block, int panel_cols,
                               const void* d_counts, const uint32_t* d_list_ptr, void* d_lists) {
  N2V2R_REQUIRE(d_in && d_counts && d_list_ptr && d_lists && n > 0 && ld >= n, "rows_fill: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int rb = rows_per_block, W = panel_cols;
  const int64_t nb = (n + rb - 1) / rb, np = (n + W - 1) / W;
  rows_fill_kernel<<<(unsigned)(nb * np), 256, (size_t)(rb + 1) * 4, s>>>(d_in, n, ld, rb, W, (int)np, (const uint16_t*)d_counts, d_list_ptr,
                                                                        (uint4*)d_lists, rows_per_group(rb), g_rows_parity);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_rows_to_dense(void* stream, const uint32_t* d_list_ptr, const void* d_lists, int64_t n, int rows_per_block,
                                   int panel_cols, float* d_out, int64_t ld_out) {
  N2V2R_REQUIRE(d_list_ptr && d_lists && d_out && n > 0 && ld_out >= n, "rows_to_dense: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t nb = (n + rows_per_block - 1) / rows_per_block, np = (n + panel_cols - 1) / panel_cols;
  rows_to_dense_kernel<<<(unsigned)(nb * np), 256, 0, s>>>(d_list_ptr, (const uint4*)d_lists, n, rows_per_block, panel_cols, (int)np, d_out, ld_out);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
