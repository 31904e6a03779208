// Shared pieces of the compressed-tile kernels (k3_sparse.cu: operands from shared memory; k3_sparse_ts.cu: A tile in tensor memory).
#pragma once
#include "k5_orth.cuh"
#include "tc_ptx.cuh"

namespace n2v2r {

constexpr int kLandCap = 512;        // entries per landing slot (4 KB)
constexpr int kUndoCap = 2048;       // positions remembered per A stage; a denser tile pair is cleared as a whole
constexpr int kSortCap = 2048;        // tile lists up to this length are written in bank-interleaved order (builder)
constexpr uint32_t kPosSkip = 0xFFFFu;   // padding entry (tile lists are padded to a multiple of 4 entries = 32 bytes)

// element offset of (row r, column c) inside one K-major SWIZZLE_128B plane of a (tile_rows x 64) bf16 tile
__host__ __device__ __forceinline__ uint32_t swz_pos(int r, int c) {
  return (uint32_t)(((r >> 3) << 9) + ((r & 7) << 6) + ((((c >> 3) ^ (r & 7))) << 3) + (c & 7));
}
__host__ __device__ __forceinline__ void swz_inv(uint32_t pos, int* r, int* c) {
  const int rr = (pos >> 6) & 7;
  *r = (int)((pos >> 9) << 3) + rr;
  *c = (int)(((((pos >> 3) & 7) ^ rr) << 3) + (pos & 7));
}

__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void st_shared_u16(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint32_t ld_shared_u16(uint32_t addr) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 ld_shared_v2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void st_shared_zero16(uint32_t addr) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(addr), "r"(0u) : "memory");
}


}  // namespace n2v2r
