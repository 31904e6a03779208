// K3/K4 for sparse layers: the layer is stored as COMPRESSED TILES and expanded in shared memory.
//
// A thresholded / top-percent layer (preprocessing_utils.py:127-135) is mostly zeros (the C3 benchmark: 2.6 %
// non-zero), but a gather-style SpMM is slower than the dense tensor-core stream.  This
// kernel keeps the dense tcgen05 product of k3_tcgen05.cu and only changes where the A tile comes from:
//
//   HBM:  per (tile_rows x 64) tile a list of 8-byte entries {pos:u16, plane0:bf16, plane1:bf16, plane2:bf16}
//         (pos = element offset of the entry inside the SWIZZLE_128B shared-memory tile, value = p0+p1+p2 exactly),
//         tiles of one row block contiguous in k order, tile_ptr[] = first entry of every tile (even, 16-byte aligned)
//   smem: the entry lists arrive by cp.async.bulk into a landing ring; four converter warps scatter them into
//         zeroed K-major SWIZZLE_128B bf16 planes (generic stores + fence.proxy.async), the single-thread
//         tcgen05.mma issuer multiplies them with the TMA-fed split operand exactly like the dense kernel, and
//         once the MMAs of a stage have retired the converters write zeros back over the positions they set
//         (the "undo" list), so a stage is never zero-filled as a whole.
//
// The product is bit-identical to the dense-plane kernel for the same K split (zero entries add exact zeros).
// Replaces the same reference code as k3_tcgen05.cu: the ARPACK operator matvec / rmatvec (scipy _svds.py:442-457)
// under spectral_embedding.UASE (model.py:54-55); the compressed form plays the role of the reference's
// csc_matrix conversion (model.py:51-53).
#include "sparse_common.cuh"
#include <algorithm>

namespace n2v2r {

// ---- CTA-pair (cta_group::2) helpers -------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(bar), "r"(cta));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(r) : "memory");
}
// wait on a local mbarrier whose arrivals come from the peer CTA (cluster-scope acquire)
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// completion of the pair's MMAs -> the mbarrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit_2cta(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((unsigned short)3) : "memory");
}
// kind::f16 instruction descriptor of the pair MMA: M = 256 (128 rows per CTA)
__device__ __forceinline__ uint32_t idesc_bf16_m256(int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((256u >> 4) << 24);
}

// =================================================================================================
// Builder: dense fp32 (n x n, leading dimension ld) -> compressed tiles of A or of A^T
// =================================================================================================
// One CTA per (128-row tile, k block) of A: counts the non-zeros of that tile of A and of the same tile of A^T
// (the mirrored 64 x 128 block), records the number of bf16 planes the values need, and whether A != A^T.
__global__ void __launch_bounds__(256) tiles_count_kernel(const float* __restrict__ A, int64_t n, int64_t ld, int n_kb,
                                                          uint32_t* __restrict__ cnt_a, uint32_t* __restrict__ cnt_t,
                                                          int* __restrict__ flags /* [0] planes, [1] asymmetric */) {
  __shared__ float tile[128][65];
  __shared__ uint32_t wsum[2][8];
  const int kb = blockIdx.x, rt = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r0 = (int64_t)rt * 128, c0 = (int64_t)kb * 64;
  uint32_t ca = 0, ct = 0;
  int planes = 1, asym = 0;
  // block of A: rows r0.., columns c0..
  for (int j = warp; j < 128; j += 8) {
    const int64_t r = r0 + j;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int64_t c = c0 + e * 32 + lane;
      const float v = (r < n && c < n) ? A[r * ld + c] : 0.f;
      tile[j][e * 32 + lane] = v;
      if (v != 0.f) {
        __nv_bfloat16 hi, mid, lo;
        split3(v, hi, mid, lo);
        if (__bfloat162float(lo) != 0.f) planes = 3;
        else if (__bfloat162float(mid) != 0.f && planes < 2) planes = 2;
      }
      ca += __popc(__ballot_sync(0xffffffffu, v != 0.f));
    }
  }
  __syncthreads();
  // mirrored block: rows c0.. (64), columns r0.. (128); element (i, j) pairs with tile[j][i]
  for (int i = warp; i < 64; i += 8) {
    const int64_t r = c0 + i;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int jj = e * 32 + lane;
      const int64_t c = r0 + jj;
      const float v = (r < n && c < n) ? A[r * ld + c] : 0.f;
      if (__float_as_uint(v) != __float_as_uint(tile[jj][i]) && !(v == 0.f && tile[jj][i] == 0.f)) asym = 1;
      ct += __popc(__ballot_sync(0xffffffffu, v != 0.f));
    }
  }
  if (lane == 0) { wsum[0][warp] = ca; wsum[1][warp] = ct; }
  planes = max(planes, __shfl_xor_sync(0xffffffffu, planes, 16));
  planes = max(planes, __shfl_xor_sync(0xffffffffu, planes, 8));
  planes = max(planes, __shfl_xor_sync(0xffffffffu, planes, 4));
  planes = max(planes, __shfl_xor_sync(0xffffffffu, planes, 2));
  planes = max(planes, __shfl_xor_sync(0xffffffffu, planes, 1));
  asym = __any_sync(0xffffffffu, asym);
  if (lane == 0) {
    if (planes > 1) atomicMax(&flags[0], planes);
    if (asym) atomicMax(&flags[1], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t sa = 0, st = 0;
    for (int w = 0; w < 8; ++w) { sa += wsum[0][w]; st += wsum[1][w]; }
    cnt_a[(size_t)rt * n_kb + kb] = sa;
    cnt_t[(size_t)rt * n_kb + kb] = st;        // the mirrored block is tile (rt, kb) of A^T
  }
}

// counts per 128-row tile -> tile_ptr of the final layout (tile_rows = 128 * mh), lists padded to an even length.
// Single CTA; out[n_tiles] = total number of entries.
__global__ void __launch_bounds__(1024) tiles_scan_kernel(const uint32_t* __restrict__ cnt128, int n_rt128, int n_kb, int mh,
                                                          uint32_t* __restrict__ tile_ptr, unsigned long long* __restrict__ total) {
  __shared__ unsigned long long part[1024];
  const int n_rt = (n_rt128 + mh - 1) / mh;
  const int64_t n_tiles = (int64_t)n_rt * n_kb;
  const int64_t per = (n_tiles + 1023) / 1024;
  const int64_t b = per * threadIdx.x, e = (b + per < n_tiles) ? b + per : n_tiles;
  auto count_of = [&](int64_t t) -> uint32_t {
    const int rt = (int)(t / n_kb), kb = (int)(t - (int64_t)rt * n_kb);
    uint32_t c = 0;
    for (int h = 0; h < mh; ++h) {
      const int r128 = rt * mh + h;
      if (r128 < n_rt128) c += cnt128[(size_t)r128 * n_kb + kb];
    }
    return (c + 3u) & ~3u;
  };
  unsigned long long s = 0;
  for (int64_t t = b; t < e; ++t) s += count_of(t);
  part[threadIdx.x] = s;
  __syncthreads();
  // inclusive scan over the 1024 partial sums (Hillis-Steele)
  for (int off = 1; off < 1024; off <<= 1) {
    unsigned long long v = threadIdx.x >= off ? part[threadIdx.x - off] : 0ull;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  unsigned long long run = threadIdx.x ? part[threadIdx.x - 1] : 0ull;
  for (int64_t t = b; t < e; ++t) {
    tile_ptr[t] = (uint32_t)run;
    run += count_of(t);
  }
  if (threadIdx.x == 1023) {
    tile_ptr[n_tiles] = (uint32_t)part[1023];
    *total = part[1023];
  }
}

// A converter thread of the product kernel handles 4 consecutive entries (one 32-byte load) and issues its j-th scatter store
// for entry 4 b + j of a block of 128 (b = its lane): so the bank-interleaved order i = 32 j + b (32 different banks per round
// j) is stored transposed, position 4 b + j.  A shorter last block of m entries (m % 4 == 0) uses the same rule with m / 4 lanes.
__device__ __forceinline__ uint32_t converter_slot(uint32_t slot, uint32_t total4) {
  const uint32_t b0 = slot & ~127u;
  const uint32_t m = (total4 - b0 < 128u) ? total4 - b0 : 128u;
  const uint32_t q = m >> 2, s_ = slot - b0;
  return b0 + 4u * (s_ % q) + s_ / q;
}

// One CTA per (tile, k block) of the TARGET matrix B (B = A, or B = A^T when TRANS): compacts the non-zeros of the tile in a
// fixed order and writes the entries.  TRANS reads the mirrored block of A (64 rows x TR columns) so every global load is
// coalesced along a row of A.
template <int TR, bool TRANS>
__global__ void __launch_bounds__(256) tiles_fill_kernel(const float* __restrict__ A, int64_t n, int64_t ld, int n_kb,
                                                         const uint32_t* __restrict__ tile_ptr, uint2* __restrict__ entries) {
  constexpr int ROWS = TRANS ? 64 : TR;        // source block: rows x cols of A
  constexpr int COLS = TRANS ? TR : 64;
  constexpr int RPW = ROWS / 8;                // source rows per warp
  constexpr int EPL = COLS / 32;               // elements per lane per source row
  __shared__ uint32_t wtot[8];
  __shared__ uint32_t bank_cnt[32];
  __shared__ uint2 stage[kSortCap];
  __shared__ uint16_t srank[kSortCap];
  const int kb = blockIdx.x, rt = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t src_r0 = TRANS ? (int64_t)kb * 64 : (int64_t)rt * TR;
  const int64_t src_c0 = TRANS ? (int64_t)rt * TR : (int64_t)kb * 64;
  float v[RPW][EPL];
  uint32_t mine = 0;
#pragma unroll
  for (int j = 0; j < RPW; ++j) {
    const int64_t r = src_r0 + warp * RPW + j;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int64_t c = src_c0 + e * 32 + lane;
      v[j][e] = (r < n && c < n) ? A[r * ld + c] : 0.f;
      mine += __popc(__ballot_sync(0xffffffffu, v[j][e] != 0.f));
    }
  }
  if (lane == 0) wtot[warp] = mine;
  if (threadIdx.x < 32) bank_cnt[threadIdx.x] = 0;
  __syncthreads();
  const size_t tile = (size_t)rt * n_kb + kb;
  const uint32_t tile_base = tile_ptr[tile];
  uint32_t base = 0, total = 0;
  for (int w = 0; w < 8; ++w) {
    if (w < warp) base += wtot[w];
    total += wtot[w];
  }
  // Lists of up to kSortCap entries are staged in shared memory and written in BANK-INTERLEAVED order: entry number j of
  // shared-memory bank b goes to slot (j, b), so that the 32 lanes of a converter warp, which scatter 32 consecutive
  // entries with one store instruction, hit 32 different banks (a random order costs ~6 conflict wavefronts per store).
  const bool sorted = total <= (uint32_t)kSortCap;
  const uint32_t total4 = (total + 3u) & ~3u;              // lists are padded to a multiple of 4 entries (32 bytes) with skip entries
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int j = 0; j < RPW; ++j) {
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const float x = v[j][e];
      const uint32_t b = __ballot_sync(0xffffffffu, x != 0.f);
      if (x != 0.f) {
        const int sr = warp * RPW + j, sc = e * 32 + lane;           // position in the source block
        const int r = TRANS ? sc : sr, c = TRANS ? sr : sc;          // position in the target tile
        __nv_bfloat16 hi, mid, lo;
        split3(x, hi, mid, lo);
        uint2 ent;
        ent.x = swz_pos(r, c) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
        ent.y = (uint32_t)__bfloat16_as_ushort(mid) | ((uint32_t)__bfloat16_as_ushort(lo) << 16);
        const uint32_t idx = base + __popc(b & lt);
        if (sorted) stage[idx] = ent;
        else entries[(size_t)tile_base + idx] = ent;
      }
      base += __popc(b);
    }
  }
  if (sorted) {
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += 256) srank[i] = (uint16_t)atomicAdd(&bank_cnt[(stage[i].x >> 1) & 31u], 1u);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += 256) {
      const uint32_t bank = (stage[i].x >> 1) & 31u, rk = srank[i];
      uint32_t slot = 0;
      for (uint32_t bb = 0; bb < 32; ++bb) {
        const uint32_t cb = bank_cnt[bb];
        slot += cb < rk ? cb : rk;                                   // complete rounds before mine
        slot += (bb < bank && cb > rk) ? 1u : 0u;                    // banks ahead of mine in my round
      }
      entries[(size_t)tile_base + converter_slot(slot, total4)] = stage[i];
    }
    for (uint32_t i = total + threadIdx.x; i < total4; i += 256) entries[(size_t)tile_base + converter_slot(i, total4)] = make_uint2(kPosSkip, 0u);
  } else {
    for (uint32_t i = total + threadIdx.x; i < total4; i += 256) entries[(size_t)tile_base + i] = make_uint2(kPosSkip, 0u);
  }
}

// test / to_dense helper: out[(rt*TR + r) * ld + kb*64 + c] = p0 + p1 + p2 (out zero-filled by the caller)
__global__ void __launch_bounds__(256) tiles_to_dense_kernel(const uint32_t* __restrict__ tile_ptr, const uint2* __restrict__ entries,
                                                             int64_t n, int n_kb, int tile_rows, float* __restrict__ out, int64_t ld) {
  const int kb = blockIdx.x, rt = blockIdx.y;
  const size_t tile = (size_t)rt * n_kb + kb;
  const uint32_t b = tile_ptr[tile], e = tile_ptr[tile + 1];
  for (uint32_t i = b + threadIdx.x; i < e; i += blockDim.x) {
    const uint2 ent = entries[i];
    const uint32_t pos = ent.x & 0xFFFFu;
    if (pos == kPosSkip) continue;
    int r, c;
    swz_inv(pos, &r, &c);
    const int64_t row = (int64_t)rt * tile_rows + r, col = (int64_t)kb * 64 + c;
    if (row < n && col < n) {
      const float p0 = __bfloat162float(__ushort_as_bfloat16((unsigned short)(ent.x >> 16)));
      const float p1 = __bfloat162float(__ushort_as_bfloat16((unsigned short)(ent.y & 0xFFFFu)));
      const float p2 = __bfloat162float(__ushort_as_bfloat16((unsigned short)(ent.y >> 16)));
      out[row * ld + col] = (p2 + p1) + p0;
    }
  }
}

// =================================================================================================
// The product kernel
// =================================================================================================
struct SparseArgs {
  int n_planes, l, mh;                 // bf16 planes issued, block width, 128-row halves per tile (tile_rows / 128)
  int a_stages, x_stages, l_slots;
  int n_tiles, n_splits, kb_per_split, n_kb;
  int64_t rows_valid, cols_valid;      // extent of the output that may be written
  int64_t out_ld, split_stride;
  uint32_t a_plane_bytes, a_stage_bytes, x_bytes;
  uint32_t off_x, off_land, off_undo;  // byte offsets from the 1024-aligned base (A stages start at 0)
  uint32_t tmem_cols;
  int debug_mode;                      // profiling only (n2v2r__debug_sparse_mode): 1 = no scatter / restore, 2 = no MMA issue, 3 = neither
  unsigned long long* debug_clk;       // profiling only: per-CTA cycle counters of the pipeline roles (16 per CTA) or null
  const uint32_t* tile_ptr;
  const uint2* entries;
  float* out;
};

// Warp roles (384 threads): 0 TMA producer of the split operand X, 1 TMEM owner + MMA issuer, 2 producer of the entry
// lists (cp.async.bulk), 3 idle, 4-7 epilogue, 8-11 converters.
//
// CTA2 = true: two CTAs of a cluster (one TPC) work on two vertically adjacent row tiles as ONE tcgen05.mma.cta_group::2 of
// M = 256: every CTA expands its own tile and holds only HALF of the split operand (the pair MMA takes N/2 operand rows from
// each CTA), so the shared-memory operand traffic per SM -- the bound of this kernel: the tensor core reads its shared-memory
// operands at ~64 B/clk -- drops from A + B to A + B/2.  CTA 0 issues the MMAs; warp 1 of CTA 1 forwards "my A stage and my
// operand half are ready" to it; completion (tcgen05.commit ... multicast) releases the stages and the accumulators in both.
template <bool CTA2>
__global__ void __launch_bounds__(384, 1) sparse_mma_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapX1,
                                                            const __grid_constant__ CUtensorMap mapX2, const SparseArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[40];
  __shared__ uint32_t tmem_base_smem;
  __shared__ uint32_t lcnt[8];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_xfull = smem_u32(&bars[0]), bar_xempty = smem_u32(&bars[4]);
  const uint32_t bar_afull = smem_u32(&bars[8]), bar_aempty = smem_u32(&bars[12]);
  const uint32_t bar_lfull = smem_u32(&bars[16]), bar_lempty = smem_u32(&bars[24]);
  const uint32_t bar_tfull = smem_u32(&bars[32]), bar_tempty = smem_u32(&bars[33]);
  const uint32_t bar_pfull = smem_u32(&bars[34]);          // CTA2, leader: the peer's A stage + operand half are ready
  const uint32_t cr = CTA2 ? cluster_ctarank() : 0u;       // rank in the CTA pair
  const int cid = CTA2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ncl = CTA2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int n_st = CTA2 ? (a.n_tiles + 1) >> 1 : a.n_tiles; // units walk (pair of) row tiles x K splits
  const int n_units = n_st * a.n_splits;
  const int nacc = 3 * a.l;

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.x_stages; ++i) { mbar_init(bar_xfull + 8 * i, 1); mbar_init(bar_xempty + 8 * i, 1); }
    for (int i = 0; i < a.a_stages; ++i) { mbar_init(bar_afull + 8 * i, 128); mbar_init(bar_aempty + 8 * i, 1); mbar_init(bar_pfull + 8 * i, 1); }
    for (int i = 0; i < a.l_slots; ++i) { mbar_init(bar_lfull + 8 * i, 1); mbar_init(bar_lempty + 8 * i, 128); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, CTA2 ? 8 : 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (CTA2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(a.tmem_cols));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(a.tmem_cols));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  // every A stage starts as zeros; from then on the converters restore the zeros themselves
  for (uint32_t i = threadIdx.x; i < a.a_stages * (a.a_stage_bytes >> 4); i += blockDim.x) st_shared_zero16(smem0 + (i << 4));
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();                            // the peer's barriers exist before anything is signalled across
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  // CTA2: operand halves of this CTA per plane term group: rows [cr * nh, (cr + 1) * nh) of the first (3 - p) l operand rows
  const int nh0 = 3 * a.l / 2, nh1 = a.l, nh2 = a.l / 2;
  const uint32_t xoff1 = (uint32_t)nh0 * 128u, xoff2 = xoff1 + (uint32_t)nh1 * 128u;

  if (warp == 0) {
    // ===== TMA producer of the split operand =====
    if (lane == 0) {
      int xs = 0; uint32_t px = 0;
      for (int u = cid; u < n_units; u += ncl) {
        const int split = u / n_st;
        const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(bar_xempty + 8 * xs, px ^ 1u);
          mbar_expect_tx(bar_xfull + 8 * xs, a.x_bytes);
          const uint32_t xb = smem0 + a.off_x + xs * a.x_bytes;
          if (CTA2) {
            tma_load_2d(&mapX, bar_xfull + 8 * xs, xb, kb * 64, (int)cr * nh0);
            if (a.n_planes > 1) tma_load_2d(&mapX1, bar_xfull + 8 * xs, xb + xoff1, kb * 64, (int)cr * nh1);
            if (a.n_planes > 2) tma_load_2d(&mapX2, bar_xfull + 8 * xs, xb + xoff2, kb * 64, (int)cr * nh2);
          } else {
            tma_load_3d(&mapX, bar_xfull + 8 * xs, xb, kb * 64, 0, 0);
          }
          if (++xs == a.x_stages) { xs = 0; px ^= 1u; }
        }
      }
    }
  } else if (warp == 1 && CTA2 && cr == 1) {
    // ===== CTA 1 of the pair: tell the issuing CTA when this CTA's A stage and operand half are ready =====
    int as = 0, xs = 0; uint32_t pa = 0, px = 0;
    for (int u = cid; u < n_units; u += ncl) {
      const int split = u / n_st;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(bar_afull + 8 * as, pa);
        mbar_wait(bar_xfull + 8 * xs, px);
        if (lane == 0) mbar_arrive_remote(bar_pfull + 8 * as, 0);
        __syncwarp();
        if (++as == a.a_stages) { as = 0; pa ^= 1u; }
        if (++xs == a.x_stages) { xs = 0; px ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int as = 0, xs = 0; uint32_t pa = 0, px = 0, uphase = 0;
    unsigned long long dbg[3] = {0, 0, 0};
    for (int u = cid; u < n_units; u += ncl) {
      const int split = u / n_st;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      const long long ct0 = clock64();
      if (CTA2) mbar_wait_cluster(bar_tempty, uphase ^ 1u);   // both CTAs' epilogues have drained the accumulators
      else mbar_wait(bar_tempty, uphase ^ 1u);
      dbg[2] += clock64() - ct0;
      tc_fence_after();
      bool first = true;
      for (int kb = kb0; kb < kb1; ++kb) {
        const long long c0 = clock64();
        mbar_wait(bar_afull + 8 * as, pa);
        const long long c1 = clock64();
        mbar_wait(bar_xfull + 8 * xs, px);
        if (CTA2) mbar_wait_cluster(bar_pfull + 8 * as, pa);
        const long long c2 = clock64();
        if (a.debug_clk && lane == 0) { dbg[0] += c1 - c0; dbg[1] += c2 - c1; }
        tc_fence_after();
        if (lane == 0) {
          const uint32_t abase0 = smem0 + as * a.a_stage_bytes;
          const uint32_t xbase = smem0 + a.off_x + xs * a.x_bytes;
          for (int h = 0; h < ((a.debug_mode & 2) ? 0 : a.mh); ++h) {
            bool acc = !first;
            for (int p = 0; p < a.n_planes; ++p) {
              const uint32_t idesc = CTA2 ? idesc_bf16_m256((3 - p) * a.l) : idesc_bf16((3 - p) * a.l);
              const uint32_t abase = abase0 + p * a.a_plane_bytes + h * 16384;
              const uint32_t xb = CTA2 ? xbase + (p == 0 ? 0u : p == 1 ? xoff1 : xoff2) : xbase;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (CTA2) umma_bf16_2cta(tmem_base + h * nacc, sw128_desc(abase + k * 32), sw128_desc(xb + k * 32), idesc, acc ? 1u : 0u);
                else umma_bf16(tmem_base + h * nacc, sw128_desc(abase + k * 32), sw128_desc(xb + k * 32), idesc, acc ? 1u : 0u);
                acc = true;
              }
            }
          }
          if (CTA2) {
            umma_commit_2cta(bar_aempty + 8 * as);     // both CTAs' converters may restore / refill this A stage
            umma_commit_2cta(bar_xempty + 8 * xs);
          } else {
            umma_commit(bar_aempty + 8 * as);          // the converters may restore / refill this A stage
            umma_commit(bar_xempty + 8 * xs);          // the operand slot is free
          }
        }
        __syncwarp();
        first = false;
        if (++as == a.a_stages) { as = 0; pa ^= 1u; }
        if (++xs == a.x_stages) { xs = 0; px ^= 1u; }
      }
      if (lane == 0) { if (CTA2) umma_commit_2cta(bar_tfull); else umma_commit(bar_tfull); }
      __syncwarp();
      uphase ^= 1u;
    }
    if (a.debug_clk && lane == 0) { for (int i = 0; i < 3; ++i) a.debug_clk[blockIdx.x * 16 + i] = dbg[i]; }
  } else if (warp == 2) {
    // ===== producer of the entry lists =====
    int ls = 0; uint32_t pl = 0;
    for (int u = cid; u < n_units; u += ncl) {
      const int split = u / n_st, st = u - split * n_st;
      const int tile = CTA2 ? 2 * st + (int)cr : st;
      const bool tile_ok = tile < a.n_tiles;            // CTA2: an odd number of row tiles leaves the last pair's second CTA empty
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      const uint32_t* tp = a.tile_ptr + (size_t)(tile_ok ? tile : 0) * a.n_kb;
      uint32_t nv0 = 0, nv1 = 0;
      if (tile_ok && kb0 + lane < kb1) { nv0 = __ldg(tp + kb0 + lane); nv1 = __ldg(tp + kb0 + lane + 1); }
      for (int kbg = kb0; kbg < kb1; kbg += 32) {
        const uint32_t v0 = nv0, v1 = nv1;
        if (tile_ok && kbg + 32 + lane < kb1) { nv0 = __ldg(tp + kbg + 32 + lane); nv1 = __ldg(tp + kbg + 32 + lane + 1); }
        const int cnt_g = min(32, kb1 - kbg);
        for (int i = 0; i < cnt_g; ++i) {
          uint32_t p0 = __shfl_sync(0xffffffffu, v0, i);
          const uint32_t p1 = __shfl_sync(0xffffffffu, v1, i);
          if (lane == 0) {
            uint32_t left = p1 - p0;
            do {
              const uint32_t c = left < (uint32_t)kLandCap ? left : (uint32_t)kLandCap;
              mbar_wait(bar_lempty + 8 * ls, pl ^ 1u);
              lcnt[ls] = c | (left == c ? 0x80000000u : 0u);
              if (c) {
                mbar_expect_tx(bar_lfull + 8 * ls, c * 8u);
                bulk_copy_g2s(smem0 + a.off_land + ls * (kLandCap * 8), a.entries + p0, c * 8u, bar_lfull + 8 * ls);
              } else {
                mbar_arrive(bar_lfull + 8 * ls);
              }
              p0 += c; left -= c;
              if (++ls == a.l_slots) { ls = 0; pl ^= 1u; }
            } while (left);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== epilogue: TMEM -> registers -> global partials =====
    const int q = warp & 3;
    uint32_t uphase = 0;
    for (int u = cid; u < n_units; u += ncl) {
      const int split = u / n_st, st = u - split * n_st;
      const int tile = CTA2 ? 2 * st + (int)cr : st;
      mbar_wait(bar_tfull, uphase);
      tc_fence_after();
      for (int h = 0; h < a.mh; ++h) {
        const int64_t row = (int64_t)tile * 128 * a.mh + h * 128 + q * 32 + lane;
        float* dst = a.out + (size_t)split * a.split_stride + (size_t)row * a.out_ld;
        const bool row_ok = row < a.rows_valid && tile < a.n_tiles;
        const bool vec_ok = row_ok && (a.out_ld % 4 == 0) && (a.l <= a.cols_valid);
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + h * nacc;
        for (int c0 = 0; c0 < a.l; c0 += 16) {
          uint32_t hi[16], mid[16], lo[16];
          tmem_ld16(taddr + c0, hi);
          tmem_ld16(taddr + a.l + c0, mid);
          tmem_ld16(taddr + 2 * a.l + c0, lo);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            float4 v;
            v.x = (__uint_as_float(lo[e]) + __uint_as_float(mid[e])) + __uint_as_float(hi[e]);
            v.y = (__uint_as_float(lo[e + 1]) + __uint_as_float(mid[e + 1])) + __uint_as_float(hi[e + 1]);
            v.z = (__uint_as_float(lo[e + 2]) + __uint_as_float(mid[e + 2])) + __uint_as_float(hi[e + 2]);
            v.w = (__uint_as_float(lo[e + 3]) + __uint_as_float(mid[e + 3])) + __uint_as_float(hi[e + 3]);
            if (vec_ok) {
              *reinterpret_cast<float4*>(dst + c0 + e) = v;
            } else if (row_ok) {
              const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int t = 0; t < 4; ++t)
                if (c0 + e + t < a.cols_valid) dst[c0 + e + t] = vv[t];
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (CTA2 && cr == 1) mbar_arrive_remote(bar_tempty, 0); else mbar_arrive(bar_tempty); }
      uphase ^= 1u;
    }
  } else if (warp >= 8) {
    // ===== converters: entry lists -> zeroed SWIZZLE_128B planes (and back to zeros after the MMAs) =====
    const int ct = threadIdx.x - 256;                  // 0..127
    int as = 0, ls = 0; uint32_t pa = 0, pl = 0;
    uint32_t prev0 = 0, prev1 = 0, prev2 = 0, prev3 = 0;   // entries scattered into each A stage by its previous use
    unsigned long long dbg[6] = {0, 0, 0, 0, 0, 0};
    // kernel parameters the loops need, in registers (left in the constant bank they cost a dependent load per store);
    // planes that are not issued alias plane 0's address and are simply not written
    const uint32_t pb1 = a.n_planes > 1 ? a.a_plane_bytes : 0u, pb2 = a.n_planes > 2 ? 2u * a.a_plane_bytes : 0u;
    const bool has1 = a.n_planes > 1, has2 = a.n_planes > 2;
    const uint32_t stage_bytes = a.a_stage_bytes, land_off = a.off_land, undo_off = a.off_undo;
    const int n_as = a.a_stages, n_ls = a.l_slots;
    const bool skip_scatter = (a.debug_mode & 1) != 0;
    for (int u = cid; u < n_units; u += ncl) {
      const int split = u / n_st;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        const uint32_t abase = smem0 + as * stage_bytes;
        const uint32_t undo = smem0 + undo_off + as * (kUndoCap * 2);
        const uint32_t prev = as == 0 ? prev0 : as == 1 ? prev1 : as == 2 ? prev2 : prev3;
        const long long c0 = clock64();
        mbar_wait(bar_aempty + 8 * as, pa ^ 1u);        // the MMAs that read this stage have retired
        const long long c1 = clock64();
        if (skip_scatter) {
        } else if (prev > (uint32_t)kUndoCap) {
          for (uint32_t i = ct; i < (stage_bytes >> 4); i += 128) st_shared_zero16(abase + (i << 4));
        } else {
          for (uint32_t i = 4 * ct; i < prev; i += 512) {        // 4 remembered positions per thread and round: one 8-byte load
            const uint2 pp = ld_shared_v2(undo + 2 * i);
            const uint32_t ps[4] = {pp.x & 0xFFFFu, pp.x >> 16, pp.y & 0xFFFFu, pp.y >> 16};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (ps[e] != kPosSkip) {
                const uint32_t ad = abase + 2 * ps[e];
                st_shared_u16(ad, 0u);
                if (has1) st_shared_u16(ad + pb1, 0u);
                if (has2) st_shared_u16(ad + pb2, 0u);
              }
            }
          }
        }
        const long long c2 = clock64();
        asm volatile("bar.sync 1, 128;" ::: "memory");  // every restore precedes every new entry
        const long long c3 = clock64();
        long long wl = 0;
        uint32_t total = 0;
        while (true) {
          const long long w0 = clock64();
          mbar_wait(bar_lfull + 8 * ls, pl);
          wl += clock64() - w0;
          const uint32_t c = lcnt[ls];
          const uint32_t nent = c & 0x7fffffffu;         // a multiple of 4, at most kLandCap = 4 x 128
          const uint32_t land = smem0 + land_off + ls * (kLandCap * 8);
          const uint32_t j = 4 * ct;
          if (!skip_scatter && j < nent) {         // this thread's 4 entries of the chunk: one 32-byte load
            const uint4 e01 = ld_shared_v4(land + 8 * j), e23 = ld_shared_v4(land + 8 * j + 16);
            const uint32_t ex[4] = {e01.x, e01.z, e23.x, e23.z}, ey[4] = {e01.y, e01.w, e23.y, e23.w};
            if (total + j < (uint32_t)kUndoCap)
              st_shared_v2(undo + 2 * (total + j), (ex[0] & 0xFFFFu) | (ex[1] << 16), (ex[2] & 0xFFFFu) | (ex[3] << 16));
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t pos = ex[e] & 0xFFFFu;
              if (pos != kPosSkip) {
                const uint32_t ad = abase + 2 * pos;
                st_shared_u16(ad, ex[e] >> 16);
                if (has1) st_shared_u16(ad + pb1, ey[e] & 0xFFFFu);
                if (has2) st_shared_u16(ad + pb2, ey[e] >> 16);
              }
            }
          }
          mbar_arrive(bar_lempty + 8 * ls);
          total += nent;
          if (++ls == n_ls) { ls = 0; pl ^= 1u; }
          if (c & 0x80000000u) break;
        }
        if (as == 0) prev0 = total; else if (as == 1) prev1 = total; else if (as == 2) prev2 = total; else prev3 = total;
        const long long c4 = clock64();
        fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core's async proxy
        mbar_arrive(bar_afull + 8 * as);
        if (a.debug_clk) {
          const long long c5 = clock64();
          dbg[0] += c1 - c0; dbg[1] += c2 - c1; dbg[2] += c3 - c2; dbg[3] += wl; dbg[4] += (c4 - c3) - wl; dbg[5] += c5 - c4;
        }
        if (++as == n_as) { as = 0; pa ^= 1u; }
      }
    }
    if (a.debug_clk && ct == 0) { for (int i = 0; i < 6; ++i) a.debug_clk[blockIdx.x * 16 + 4 + i] = dbg[i]; }
  }
  tc_fence_before();
  __syncthreads();
  if (CTA2) cluster_sync_all();                            // nothing is signalled into a CTA that has left
  if (warp == 1) {
    tc_fence_after();
    if (CTA2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols));
  }
}

// ---- host ---------------------------------------------------------------------------------------
static int g_sparse_debug_mode = 0;
static int g_sparse_ts = 0;              // 1 = A tile in tensor memory (k3_sparse_ts.cu); switched by n2v2r__debug_sparse_ts
int stream_apply_tiles_ts(cudaStream_t s, const StreamPlan& plan, const uint32_t* tile_ptr, const void* entries, int n_planes,
                          int tile_rows, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate);
static int g_sparse_cta2 = 0;            // 1 = pair kernel (cta_group::2); switched by n2v2r__debug_sparse_cta2 until it is the default
static unsigned long long* g_sparse_debug_clk = nullptr;
// K split that balances (row tiles x splits) units over the persistent CTAs
static int pick_splits(int n_tiles, int n_kb, int max_splits, int grid) {
  int best = 1;
  double best_cost = 1e300;
  for (int s = 1; s <= max_splits && s <= n_kb; ++s) {
    const int kbps = (n_kb + s - 1) / s;
    const int s_eff = (n_kb + kbps - 1) / kbps;
    const long units = (long)n_tiles * s_eff;
    const long waves = (units + grid - 1) / grid;
    const double cost = (double)waves * (kbps + 3.0);      // + epilogue / pipeline fill per unit
    if (cost < best_cost * 0.999) { best_cost = cost; best = s_eff; }
  }
  return best;
}

int stream_apply_tiles(cudaStream_t s, const StreamPlan& plan, const uint32_t* tile_ptr, const void* entries, int n_planes,
                       int tile_rows, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate) {
  N2V2R_REQUIRE(xs && parts && tile_ptr && entries, "compressed-tile product needs its workspace and the tile lists");
  N2V2R_REQUIRE(tile_rows == 128 || tile_rows == 256, "compressed tiles: tile_rows must be 128 or 256");
  N2V2R_REQUIRE(n_planes >= 1 && n_planes <= 3, "compressed tiles: 1..3 planes");
  if (g_sparse_ts && (512 - (tile_rows / 128) * 3 * plan.l) / (n_planes * 32) >= tile_rows / 128 && !(tile_rows == 256 && n_planes == 3))
    return stream_apply_tiles_ts(s, plan, tile_ptr, entries, n_planes, tile_rows, x, xs, parts, out, accumulate);
  int rc = launch_split_transpose(s, x, plan.n, plan.l, plan.n_pad, xs);
  if (rc) return rc;
  SparseArgs a;
  a.n_planes = n_planes;
  a.l = plan.l;
  a.mh = tile_rows / 128;
  a.n_kb = (int)(plan.n_pad / 64);
  a.n_tiles = (int)((plan.n + tile_rows - 1) / tile_rows);
  const int grid_max = sm_count();
  const bool cta2 = g_sparse_cta2 != 0 && grid_max >= 2;
  a.a_plane_bytes = 16384u * a.mh;
  a.a_stage_bytes = a.a_plane_bytes * n_planes;
  // operand stage: the whole split operand [3][l][64], or (pair kernel) this CTA's halves of the first 3l / 2l / l rows
  a.x_bytes = cta2 ? 128u * (uint32_t)(3 * plan.l / 2 + (n_planes > 1 ? plan.l : 0) + (n_planes > 2 ? plan.l / 2 : 0)) : 3u * plan.l * 128u;
  {
    int tc = 32;
    while (tc < a.mh * 3 * plan.l) tc <<= 1;
    N2V2R_REQUIRE(tc <= 512, "compressed tiles: accumulators do not fit in TMEM");
    a.tmem_cols = (uint32_t)tc;
  }
  const int units_grid = cta2 ? grid_max / 2 : grid_max;
  const int n_st = cta2 ? (a.n_tiles + 1) / 2 : a.n_tiles;
  if (plan.k_chunk_user > 0) {                     // the caller fixed the K split (A/B runs against the dense kernel)
    a.kb_per_split = plan.k_chunk / 64;
  } else {
    const int sp = pick_splits(n_st, a.n_kb, plan.n_splits, units_grid);
    a.kb_per_split = (a.n_kb + sp - 1) / sp;
  }
  a.n_splits = (a.n_kb + a.kb_per_split - 1) / a.kb_per_split;
  N2V2R_REQUIRE(a.n_splits <= plan.n_splits, "compressed tiles: K split exceeds the workspace plan");
  // shared-memory plan: two A stages, eight landing slots (fewer when tight) and their undo lists first; what is left goes
  // to operand stages (up to 4), then to further A stages (up to 4)
  const uint32_t budget = 227u * 1024u - 512u - 1024u;
  auto need = [&](int as, int xs, int ls) {
    return (uint32_t)as * a.a_stage_bytes + (uint32_t)xs * a.x_bytes + (uint32_t)ls * (uint32_t)(kLandCap * 8) + (uint32_t)as * (uint32_t)(kUndoCap * 2);
  };
  a.a_stages = 2; a.x_stages = 2; a.l_slots = 8;
  while (a.l_slots > 3 && need(2, 2, a.l_slots) > budget) --a.l_slots;
  N2V2R_REQUIRE(need(2, 2, a.l_slots) <= budget, "compressed tiles: stages do not fit in shared memory (l=%d, planes=%d, tile_rows=%d)",
                plan.l, n_planes, tile_rows);
  while (a.x_stages < 4 && need(a.a_stages, a.x_stages + 1, a.l_slots) <= budget) ++a.x_stages;
  while (a.a_stages < 4 && need(a.a_stages + 1, a.x_stages, a.l_slots) <= budget) ++a.a_stages;
  a.off_x = a.a_stages * a.a_stage_bytes;                       // multiples of 1024 (operand rows come in multiples of 8)
  a.off_land = a.off_x + a.x_stages * a.x_bytes;
  a.off_undo = a.off_land + a.l_slots * (uint32_t)(kLandCap * 8);
  const size_t smem = (size_t)a.off_undo + (size_t)a.a_stages * (kUndoCap * 2) + 1024;
  const int64_t rows_pad = round_up(plan.n, 256);
  a.rows_valid = rows_pad;
  a.cols_valid = plan.l;
  a.out_ld = plan.l;
  a.split_stride = rows_pad * plan.l;
  a.debug_mode = g_sparse_debug_mode;
  a.debug_clk = g_sparse_debug_clk;
  a.tile_ptr = tile_ptr;
  a.entries = (const uint2*)entries;
  a.out = parts;
  CUtensorMap mX, mX1, mX2;
  if (cta2) {       // the split operand as one [3l x n_pad] matrix; boxes = the halves of its first 3l / 2l / l rows
    const uint64_t dims[2] = {(uint64_t)plan.n_pad, (uint64_t)(3 * plan.l)};
    const uint64_t str[1] = {(uint64_t)plan.n_pad * 2};
    const uint32_t b0[2] = {64, (uint32_t)(3 * plan.l / 2)}, b1[2] = {64, (uint32_t)plan.l}, b2[2] = {64, (uint32_t)(plan.l / 2)};
    rc = make_map(&mX, xs, 2, dims, str, b0);
    if (rc == N2V2R_OK) rc = make_map(&mX1, xs, 2, dims, str, b1);
    if (rc == N2V2R_OK) rc = make_map(&mX2, xs, 2, dims, str, b2);
    if (rc) return rc;
  } else {
    const uint64_t dims[3] = {(uint64_t)plan.n_pad, (uint64_t)plan.l, 3};
    const uint64_t str[2] = {(uint64_t)plan.n_pad * 2, (uint64_t)plan.l * plan.n_pad * 2};
    const uint32_t box[3] = {64, (uint32_t)plan.l, 3};
    rc = make_map(&mX, xs, 3, dims, str, box);
    if (rc) return rc;
    mX1 = mX; mX2 = mX;
  }
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(sparse_mma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 512)));
    N2V2R_CUDA(cudaFuncSetAttribute(sparse_mma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 512)));
    attr_set = true;
  }
  const int n_units = n_st * a.n_splits;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  rc = mma_events_begin(s, &e0, &e1);
  if (rc) return rc;
  if (cta2) {
    const int pairs = n_units < units_grid ? n_units : units_grid;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    N2V2R_CUDA(cudaLaunchKernelEx(&cfg, sparse_mma_kernel<true>, mX, mX1, mX2, a));
  } else {
    const int grid = n_units < grid_max ? n_units : grid_max;
    sparse_mma_kernel<false><<<grid, 384, smem, s>>>(mX, mX1, mX2, a);
  }
  N2V2R_CHECK_LAUNCH();
  rc = mma_events_end(s, e0, e1);
  if (rc) return rc;
  return launch_reduce_parts(s, parts, a.n_splits, rows_pad, plan.n, plan.l, out, accumulate);
}

}  // namespace n2v2r

using namespace n2v2r;

// profiling hook (not part of the ABI header): skip parts of the sparse product kernel to time the rest
extern "C" void n2v2r__debug_sparse_mode(int mode) { g_sparse_debug_mode = mode; }
extern "C" void n2v2r__debug_sparse_clk(void* d_buf) { g_sparse_debug_clk = (unsigned long long*)d_buf; }
extern "C" void n2v2r__debug_sparse_cta2(int on) { g_sparse_cta2 = on; }
extern "C" void n2v2r__debug_sparse_ts(int on) { g_sparse_ts = on; }

// ---- C ABI: building the compressed tiles --------------------------------------------------------
extern "C" int n2v2r_tiles_count(void* stream, const float* d_in, int64_t n, int64_t ld, uint32_t* d_counts, uint32_t* d_counts_t,
                                 int* h_planes, int* h_symmetric) {
  N2V2R_REQUIRE(d_in && d_counts && d_counts_t && n > 0 && ld >= n, "tiles_count: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_kb = (int)(round_up(n, 64) / 64), n_rt = (int)((n + 127) / 128);
  int* d_flags = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_flags, 2 * sizeof(int), s));
  N2V2R_CUDA(cudaMemsetAsync(d_flags, 0, 2 * sizeof(int), s));
  tiles_count_kernel<<<dim3((unsigned)n_kb, (unsigned)n_rt), 256, 0, s>>>(d_in, n, ld, n_kb, d_counts, d_counts_t, d_flags);
  cudaError_t le = cudaGetLastError();
  int flags[2] = {0, 0};
  if (le == cudaSuccess) le = cudaMemcpyAsync(flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, s);
  if (le == cudaSuccess) le = cudaStreamSynchronize(s);
  cudaFreeAsync(d_flags, s);
  if (le != cudaSuccess) return cuda_fail(le, "tiles_count", __FILE__, __LINE__);
  if (h_planes) *h_planes = flags[0] < 1 ? 1 : flags[0];
  if (h_symmetric) *h_symmetric = flags[1] ? 0 : 1;
  return N2V2R_OK;
}

extern "C" int n2v2r_tiles_scan(void* stream, const uint32_t* d_counts, int64_t n, int tile_rows, uint32_t* d_tile_ptr,
                                int64_t* h_entries) {
  N2V2R_REQUIRE(d_counts && d_tile_ptr && h_entries && n > 0 && (tile_rows == 128 || tile_rows == 256), "tiles_scan: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_kb = (int)(round_up(n, 64) / 64), n_rt128 = (int)((n + 127) / 128);
  unsigned long long* d_total = nullptr;
  N2V2R_CUDA(cudaMallocAsync((void**)&d_total, sizeof(unsigned long long), s));
  tiles_scan_kernel<<<1, 1024, 0, s>>>(d_counts, n_rt128, n_kb, tile_rows / 128, d_tile_ptr, d_total);
  cudaError_t le = cudaGetLastError();
  unsigned long long total = 0;
  if (le == cudaSuccess) le = cudaMemcpyAsync(&total, d_total, sizeof(total), cudaMemcpyDeviceToHost, s);
  if (le == cudaSuccess) le = cudaStreamSynchronize(s);
  cudaFreeAsync(d_total, s);
  if (le != cudaSuccess) return cuda_fail(le, "tiles_scan", __FILE__, __LINE__);
  if (total >= 0xFFFFFFF0ull) { set_error("tiles_scan: %llu entries do not fit 32-bit tile offsets", total); return N2V2R_EUNSUPPORTED; }
  *h_entries = (int64_t)total;
  return N2V2R_OK;
}

extern "C" int n2v2r_tiles_fill(void* stream, const float* d_in, int64_t n, int64_t ld, int transpose, int tile_rows,
                                const uint32_t* d_tile_ptr, void* d_entries) {
  N2V2R_REQUIRE(d_in && d_tile_ptr && d_entries && n > 0 && ld >= n && (tile_rows == 128 || tile_rows == 256), "tiles_fill: bad arguments");
  N2V2R_REQUIRE(((uintptr_t)d_entries & 15) == 0, "tiles_fill: the entry array must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_kb = (int)(round_up(n, 64) / 64), n_rt = (int)((n + tile_rows - 1) / tile_rows);
  const dim3 grid((unsigned)n_kb, (unsigned)n_rt);
  uint2* ent = (uint2*)d_entries;
  if (tile_rows == 256) {
    if (transpose) tiles_fill_kernel<256, true><<<grid, 256, 0, s>>>(d_in, n, ld, n_kb, d_tile_ptr, ent);
    else tiles_fill_kernel<256, false><<<grid, 256, 0, s>>>(d_in, n, ld, n_kb, d_tile_ptr, ent);
  } else {
    if (transpose) tiles_fill_kernel<128, true><<<grid, 256, 0, s>>>(d_in, n, ld, n_kb, d_tile_ptr, ent);
    else tiles_fill_kernel<128, false><<<grid, 256, 0, s>>>(d_in, n, ld, n_kb, d_tile_ptr, ent);
  }
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

extern "C" int n2v2r_tiles_to_dense(void* stream, const uint32_t* d_tile_ptr, const void* d_entries, int64_t n, int tile_rows,
                                    float* d_out, int64_t ld_out) {
  N2V2R_REQUIRE(d_tile_ptr && d_out && n > 0 && ld_out >= n && (tile_rows == 128 || tile_rows == 256), "tiles_to_dense: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_kb = (int)(round_up(n, 64) / 64), n_rt = (int)((n + tile_rows - 1) / tile_rows);
  tiles_to_dense_kernel<<<dim3((unsigned)n_kb, (unsigned)n_rt), 256, 0, s>>>(d_tile_ptr, (const uint2*)d_entries, n, n_kb, tile_rows, d_out, ld_out);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}
