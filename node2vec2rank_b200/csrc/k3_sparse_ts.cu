// K3/K4 for compressed-tile layers with the A TILE IN TENSOR MEMORY (tcgen05.mma with the A operand from TMEM).
//
// k3_sparse.cu expands a tile in shared memory and multiplies it from there; that kernel sits on the tensor core's
// shared-memory A-operand feed (~32 B/clk/SM: every 128 x 16 bf16 slice costs ~128 clk whatever N is).
// Here the converter warps expand the tile in shared memory exactly the same way (entry lists -> zeroed SWIZZLE_128B planes,
// undo list), then every converter thread reads ITS OWN ROW of the tile back (8 x 16 bytes per plane, conflict-free in the
// swizzled layout), writes it into tensor memory with tcgen05.st (lane = row, one 32-bit column = two consecutive k) and
// restores the zeros at once; the MMAs take A from TMEM and only the split operand from shared memory.  The shared-memory
// tile is free again as soon as it has been read, so what rings is the TMEM copy of A: slots of ONE 128-row half each (P x 32
// columns) next to the accumulators, released half by half as the MMAs retire.  The chain scatter -> barrier -> read back ->
// tcgen05.st -> barrier -> undo of one k block is latency, not bandwidth, so TWO converter teams of 128 threads (warps 8-11 and
// 12-15) work on alternate k blocks, each with its own staging tile, undo list and ring of landing buffers.
// Same entry format, same K split and MMA order as k3_sparse.cu: the product has the same bits.
#include "sparse_common.cuh"
#include <algorithm>

namespace n2v2r {

__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,"
      "%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
        "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
        "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void team_bar(int team) { asm volatile("bar.sync %0, 128;" ::"r"(team + 1) : "memory"); }
__device__ __forceinline__ void st_shared_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_shared_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct SparseTsArgs {
  int l, x_stages, l_slots, ta_slots;       // l_slots: landing buffers PER TEAM; ta_slots: TMEM A slots of one 128-row half each
  int n_tiles, n_splits, kb_per_split, n_kb;
  int64_t rows_valid, out_ld, split_stride;
  uint32_t a_plane_bytes, a_buf_bytes, x_bytes;
  uint32_t off_x, off_land, off_undo;      // byte offsets from the 1024-aligned base (the two staging tiles start at 0)
  const uint32_t* tile_ptr;
  const uint2* entries;
  float* out;
};

// MH = 128-row halves per tile (tile_rows / 128), P = bf16 plane terms.  Warp roles as in sparse_mma_kernel.
template <int MH, int P>
__global__ void __launch_bounds__(512, 1) sparse_ts_kernel(const __grid_constant__ CUtensorMap mapX, const SparseTsArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[40];
  __shared__ uint32_t tmem_base_smem;
  __shared__ uint32_t lcnt[8];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_xfull = smem_u32(&bars[0]), bar_xempty = smem_u32(&bars[8]);
  const uint32_t bar_tafull = smem_u32(&bars[16]), bar_taempty = smem_u32(&bars[20]);    // up to 4 slots
  const uint32_t bar_lfull = smem_u32(&bars[24]), bar_lempty = smem_u32(&bars[32]);
  __shared__ __align__(8) uint64_t tbars[2];
  const uint32_t bar_tfull = smem_u32(&tbars[0]), bar_tempty = smem_u32(&tbars[1]);
  const int n_units = a.n_tiles * a.n_splits;
  const int nacc = 3 * a.l;
  constexpr int kSlotCols = P * 32;                      // TMEM columns of one A slot: one half, per plane 64 k = 32 columns
  constexpr int kTeam = 128;                             // threads of one converter team
  const uint32_t a_col0 = (uint32_t)(MH * nacc);         // the A ring follows the accumulators

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.x_stages; ++i) { mbar_init(bar_xfull + 8 * i, 1); mbar_init(bar_xempty + 8 * i, 1); }
    for (int i = 0; i < a.ta_slots; ++i) { mbar_init(bar_tafull + 8 * i, 128); mbar_init(bar_taempty + 8 * i, 1); }
    for (int i = 0; i < 2 * a.l_slots; ++i) { mbar_init(bar_lfull + 8 * i, 1); mbar_init(bar_lempty + 8 * i, kTeam); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (uint32_t i = threadIdx.x; i < 2u * (a.a_buf_bytes >> 4); i += blockDim.x) st_shared_zero16(smem0 + (i << 4));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===== TMA producer of the split operand =====
    if (lane == 0) {
      int xs = 0; uint32_t px = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int split = u / a.n_tiles;
        const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(bar_xempty + 8 * xs, px ^ 1u);
          mbar_expect_tx(bar_xfull + 8 * xs, a.x_bytes);
          tma_load_3d(&mapX, bar_xfull + 8 * xs, smem0 + a.off_x + xs * a.x_bytes, kb * 64, 0, 0);
          if (++xs == a.x_stages) { xs = 0; px ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: A from tensor memory, the split operand from shared memory =====
    int ta = 0, xs = 0; uint32_t pta = 0, px = 0, uphase = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      mbar_wait(bar_tempty, uphase ^ 1u);
      tc_fence_after();
      bool first = true;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(bar_xfull + 8 * xs, px);
        const uint32_t xbase = smem0 + a.off_x + xs * a.x_bytes;
#pragma unroll
        for (int h = 0; h < MH; ++h) {
          mbar_wait(bar_tafull + 8 * ta, pta);           // this half's rows are in its TMEM slot
          tc_fence_after();
          if (lane == 0) {
            const uint32_t abase = tmem_base + a_col0 + (uint32_t)(ta * kSlotCols);
            bool acc = !first;
#pragma unroll
            for (int p = 0; p < P; ++p) {
              const uint32_t idesc = idesc_bf16((3 - p) * a.l);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                umma_bf16_ts(tmem_base + h * nacc, abase + (uint32_t)(p * 32 + k * 8), sw128_desc(xbase + k * 32), idesc, acc ? 1u : 0u);
                acc = true;
              }
            }
            umma_commit(bar_taempty + 8 * ta);         // this TMEM slot may be overwritten
            if (h == MH - 1) umma_commit(bar_xempty + 8 * xs);
          }
          __syncwarp();
          if (++ta == a.ta_slots) { ta = 0; pta ^= 1u; }
        }
        first = false;
        if (++xs == a.x_stages) { xs = 0; px ^= 1u; }
      }
      if (lane == 0) umma_commit(bar_tfull);
      __syncwarp();
      uphase ^= 1u;
    }
  } else if (warp == 2) {
    // ===== producer of the entry lists: k block i of this CTA goes to the landing ring of team i & 1 =====
    int ls0 = 0, ls1 = 0; uint32_t pl0 = 0, pl1 = 0;
    int tm = 0;                                          // team of the next k block
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles, tile = u - split * a.n_tiles;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      const uint32_t* tp = a.tile_ptr + (size_t)tile * a.n_kb;
      uint32_t nv0 = 0, nv1 = 0;
      if (kb0 + lane < kb1) { nv0 = __ldg(tp + kb0 + lane); nv1 = __ldg(tp + kb0 + lane + 1); }
      for (int kbg = kb0; kbg < kb1; kbg += 32) {
        const uint32_t v0 = nv0, v1 = nv1;
        if (kbg + 32 + lane < kb1) { nv0 = __ldg(tp + kbg + 32 + lane); nv1 = __ldg(tp + kbg + 32 + lane + 1); }
        const int cnt_g = min(32, kb1 - kbg);
        for (int i = 0; i < cnt_g; ++i) {
          uint32_t p0 = __shfl_sync(0xffffffffu, v0, i);
          const uint32_t p1 = __shfl_sync(0xffffffffu, v1, i);
          if (lane == 0) {
            uint32_t left = p1 - p0;
            int lsv = tm ? ls1 : ls0; uint32_t plv = tm ? pl1 : pl0;
            do {
              const uint32_t c = left < (uint32_t)kLandCap ? left : (uint32_t)kLandCap;
              const int slot = tm * a.l_slots + lsv;
              mbar_wait(bar_lempty + 8 * slot, plv ^ 1u);
              lcnt[slot] = c | (left == c ? 0x80000000u : 0u);
              if (c) {
                mbar_expect_tx(bar_lfull + 8 * slot, c * 8u);
                bulk_copy_g2s(smem0 + a.off_land + slot * (kLandCap * 8), a.entries + p0, c * 8u, bar_lfull + 8 * slot);
              } else {
                mbar_arrive(bar_lfull + 8 * slot);
              }
              p0 += c; left -= c;
              if (++lsv == a.l_slots) { lsv = 0; plv ^= 1u; }
            } while (left);
            if (tm) { ls1 = lsv; pl1 = plv; } else { ls0 = lsv; pl0 = plv; }
          }
          tm ^= 1;
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== epilogue: TMEM -> registers -> global partials =====
    const int q = warp & 3;
    uint32_t uphase = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles, tile = u - split * a.n_tiles;
      mbar_wait(bar_tfull, uphase);
      tc_fence_after();
#pragma unroll
      for (int h = 0; h < MH; ++h) {
        const int64_t row = (int64_t)tile * 128 * MH + h * 128 + q * 32 + lane;
        float* dst = a.out + (size_t)split * a.split_stride + (size_t)row * a.out_ld;
        const bool row_ok = row < a.rows_valid;
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
            if (row_ok) *reinterpret_cast<float4*>(dst + c0 + e) = v;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty);
      uphase ^= 1u;
    }
  } else if (warp >= 8) {
    // ===== converters: entry lists -> zeroed SWIZZLE_128B staging tile -> own rows -> tensor memory =====
    const int ct = threadIdx.x - 256;                  // 0 .. 255
    const int tm = ct >> 7;                            // team: takes the k blocks i of this CTA with (i & 1) == tm
    const int rr = ct & 127;                           // row inside a half = TMEM lane
    const int q = warp & 3;                            // TMEM lane quarter of this warp
    int ls = 0; uint32_t pl = 0;
    long ikb = 0;                                      // running k block index of this CTA
    const uint32_t pb = a.a_plane_bytes, buf_bytes = a.a_buf_bytes;
    const uint32_t abase = smem0 + tm * buf_bytes;
    const uint32_t undo = smem0 + a.off_undo + tm * (kUndoCap * 2);
    const uint32_t land0 = smem0 + a.off_land + tm * a.l_slots * (kLandCap * 8);
    const uint32_t lfull0 = bar_lfull + 8 * tm * a.l_slots, lempty0 = bar_lempty + 8 * tm * a.l_slots;
    const uint32_t* lc = lcnt + tm * a.l_slots;
    const int n_ls = a.l_slots, n_ta = a.ta_slots;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int split = u / a.n_tiles;
      const int kb0 = split * a.kb_per_split, kb1 = min(a.n_kb, kb0 + a.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb, ++ikb) {
        if ((int)(ikb & 1) != tm) continue;
        // ---- 1. scatter this k block's entries into the (zero) staging tile, remembering the positions
        uint32_t total = 0;
        while (true) {
          mbar_wait(lfull0 + 8 * ls, pl);
          const uint32_t c = lc[ls];
          const uint32_t nent = c & 0x7fffffffu;
          const uint32_t land = land0 + ls * (kLandCap * 8);
          const uint32_t j = 4 * rr;
          if (j < nent) {
            uint32_t ex[4], ey[4];
            const uint4 e01 = ld_shared_v4(land + 8 * j), e23 = ld_shared_v4(land + 8 * j + 16);
            ex[0] = e01.x; ey[0] = e01.y; ex[1] = e01.z; ey[1] = e01.w;
            ex[2] = e23.x; ey[2] = e23.y; ex[3] = e23.z; ey[3] = e23.w;
            if (total + j < (uint32_t)kUndoCap)
              st_shared_v2(undo + 2 * (total + j), (ex[0] & 0xFFFFu) | (ex[1] << 16), (ex[2] & 0xFFFFu) | (ex[3] << 16));
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t pos = ex[e] & 0xFFFFu;
              if (pos != kPosSkip) {
                const uint32_t ad = abase + 2 * pos;
                st_shared_u16(ad, ex[e] >> 16);
                if (P > 1) st_shared_u16(ad + pb, ey[e] & 0xFFFFu);
                if (P > 2) st_shared_u16(ad + 2 * pb, ey[e] >> 16);
              }
            }
          }
          mbar_arrive(lempty0 + 8 * ls);
          total += nent;
          if (++ls == n_ls) { ls = 0; pl ^= 1u; }
          if (c & 0x80000000u) break;
        }
        team_bar(tm);                                   // the tile is complete
        // ---- 2. my row of every half and plane (8 x 16 bytes per plane; chunk j of row r sits at (j ^ (r & 7)) in the swizzled
        //         layout) into the TMEM slot of that half: lane = row, column c = k 2c and 2c + 1
#pragma unroll
        for (int h = 0; h < MH; ++h) {
          uint32_t r[P * 32];
          const int row = h * 128 + rr;
          const uint32_t rbase = abase + (uint32_t)((row >> 3) << 10) + (uint32_t)((row & 7) << 7);
#pragma unroll
          for (int p = 0; p < P; ++p) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 v = ld_shared_v4(rbase + p * pb + (uint32_t)((j ^ (row & 7)) << 4));
              uint32_t* d = r + p * 32 + 4 * j;
              d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
            }
          }
          const long it = ikb * MH + h;
          const int ta = (int)(it % n_ta);
          const uint32_t pta = (uint32_t)((it / n_ta) & 1);
          mbar_wait(bar_taempty + 8 * ta, pta ^ 1u);    // the MMAs that read this TMEM slot have retired
          tc_fence_after();
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + a_col0 + (uint32_t)(ta * kSlotCols);
#pragma unroll
          for (int p = 0; p < P; ++p) tmem_st32(tbase + p * 32, r + p * 32);
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(bar_tafull + 8 * ta);
        }
        team_bar(tm);                                   // every row has been read: the zeros can come back
        // ---- 3. restore the zeros
        if (total > (uint32_t)kUndoCap) {
          for (uint32_t i = rr; i < (buf_bytes >> 4); i += kTeam) st_shared_zero16(abase + (i << 4));
        } else {
          for (uint32_t i = 4 * rr; i < total; i += 512) {
            const uint2 pp = ld_shared_v2(undo + 2 * i);
            const uint32_t ps[4] = {pp.x & 0xFFFFu, pp.x >> 16, pp.y & 0xFFFFu, pp.y >> 16};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (ps[e] != kPosSkip) {
                const uint32_t ad = abase + 2 * ps[e];
                st_shared_u16(ad, 0u);
                if (P > 1) st_shared_u16(ad + pb, 0u);
                if (P > 2) st_shared_u16(ad + 2 * pb, 0u);
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

static int pick_splits_ts(int n_tiles, int n_kb, int max_splits, int grid) {
  int best = 1;
  double best_cost = 1e300;
  for (int s = 1; s <= max_splits && s <= n_kb; ++s) {
    const int kbps = (n_kb + s - 1) / s;
    const int s_eff = (n_kb + kbps - 1) / kbps;
    const long units = (long)n_tiles * s_eff;
    const long waves = (units + grid - 1) / grid;
    const double cost = (double)waves * (kbps + 3.0);
    if (cost < best_cost * 0.999) { best_cost = cost; best = s_eff; }
  }
  return best;
}

template <int MH, int P>
static int launch_ts(cudaStream_t s, const CUtensorMap& mX, const SparseTsArgs& a, int grid, size_t smem) {
  static bool attr_set = false;
  if (!attr_set) {
    N2V2R_CUDA(cudaFuncSetAttribute(sparse_ts_kernel<MH, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024)));
    attr_set = true;
  }
  sparse_ts_kernel<MH, P><<<grid, 512, smem, s>>>(mX, a);
  N2V2R_CHECK_LAUNCH();
  return N2V2R_OK;
}

int stream_apply_tiles_ts(cudaStream_t s, const StreamPlan& plan, const uint32_t* tile_ptr, const void* entries, int n_planes,
                          int tile_rows, const float* x, __nv_bfloat16* xs, float* parts, float* out, int accumulate) {
  N2V2R_REQUIRE(xs && parts && tile_ptr && entries, "compressed-tile product needs its workspace and the tile lists");
  N2V2R_REQUIRE(tile_rows == 128 || tile_rows == 256, "compressed tiles: tile_rows must be 128 or 256");
  N2V2R_REQUIRE(n_planes >= 1 && n_planes <= 3 && !(tile_rows == 256 && n_planes == 3), "compressed tiles: 1..3 planes (3 with 128-row tiles)");
  int rc = launch_split_transpose(s, x, plan.n, plan.l, plan.n_pad, xs);
  if (rc) return rc;
  SparseTsArgs a;
  const int mh = tile_rows / 128;
  a.l = plan.l;
  a.n_kb = (int)(plan.n_pad / 64);
  a.n_tiles = (int)((plan.n + tile_rows - 1) / tile_rows);
  const int grid_max = sm_count();
  if (plan.k_chunk_user > 0) {
    a.kb_per_split = plan.k_chunk / 64;
  } else {
    const int sp = pick_splits_ts(a.n_tiles, a.n_kb, plan.n_splits, grid_max);
    a.kb_per_split = (a.n_kb + sp - 1) / sp;
  }
  a.n_splits = (a.n_kb + a.kb_per_split - 1) / a.kb_per_split;
  N2V2R_REQUIRE(a.n_splits <= plan.n_splits, "compressed tiles: K split exceeds the workspace plan");
  a.ta_slots = (512 - mh * 3 * plan.l) / (n_planes * 32);
  if (a.ta_slots > 4) a.ta_slots = 4;
  N2V2R_REQUIRE(a.ta_slots >= mh, "compressed tiles (A in TMEM): accumulators and the A slots do not fit in tensor memory (l=%d)", plan.l);
  a.a_plane_bytes = 16384u * mh;
  a.a_buf_bytes = a.a_plane_bytes * n_planes;
  a.x_bytes = 3u * plan.l * 128u;
  const uint32_t budget = 227u * 1024u - 1024u - 1024u;
  a.l_slots = 4;                                          // per team
  const uint32_t fixed = 2u * a.a_buf_bytes + 2u * (uint32_t)(kUndoCap * 2);
  while (a.l_slots > 2 && fixed + 2u * a.l_slots * (uint32_t)(kLandCap * 8) + 2 * a.x_bytes > budget) --a.l_slots;
  N2V2R_REQUIRE(fixed + 2u * a.l_slots * (uint32_t)(kLandCap * 8) + 2 * a.x_bytes <= budget, "compressed tiles (A in TMEM): stages do not fit in shared memory");
  a.x_stages = (int)std::min<uint32_t>(8u, (budget - fixed - 2u * a.l_slots * (uint32_t)(kLandCap * 8)) / a.x_bytes);
  a.off_x = 2u * a.a_buf_bytes;
  a.off_land = a.off_x + a.x_stages * a.x_bytes;
  a.off_undo = a.off_land + 2u * a.l_slots * (uint32_t)(kLandCap * 8);
  const size_t smem = (size_t)a.off_undo + 2u * (kUndoCap * 2) + 1024;
  const int64_t rows_pad = round_up(plan.n, 256);
  a.rows_valid = rows_pad;
  a.out_ld = plan.l;
  a.split_stride = rows_pad * plan.l;
  a.tile_ptr = tile_ptr;
  a.entries = (const uint2*)entries;
  a.out = parts;
  CUtensorMap mX;
  {
    const uint64_t dims[3] = {(uint64_t)plan.n_pad, (uint64_t)plan.l, 3};
    const uint64_t str[2] = {(uint64_t)plan.n_pad * 2, (uint64_t)plan.l * plan.n_pad * 2};
    const uint32_t box[3] = {64, (uint32_t)plan.l, 3};
    rc = make_map(&mX, xs, 3, dims, str, box);
    if (rc) return rc;
  }
  const int n_units = a.n_tiles * a.n_splits;
  const int grid = n_units < grid_max ? n_units : grid_max;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  rc = mma_events_begin(s, &e0, &e1);
  if (rc) return rc;
  if (mh == 2 && n_planes == 1) rc = launch_ts<2, 1>(s, mX, a, grid, smem);
  else if (mh == 2 && n_planes == 2) rc = launch_ts<2, 2>(s, mX, a, grid, smem);
  else if (mh == 1 && n_planes == 1) rc = launch_ts<1, 1>(s, mX, a, grid, smem);
  else if (mh == 1 && n_planes == 2) rc = launch_ts<1, 2>(s, mX, a, grid, smem);
  else rc = launch_ts<1, 3>(s, mX, a, grid, smem);
  if (rc) return rc;
  rc = mma_events_end(s, e0, e1);
  if (rc) return rc;
  return launch_reduce_parts(s, parts, a.n_splits, rows_pad, plan.n, plan.l, out, accumulate);
}

}  // namespace n2v2r
