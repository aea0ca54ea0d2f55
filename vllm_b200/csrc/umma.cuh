// tcgen05 / TMEM primitives for sm_100a, hand-written PTX (no CUTLASS).
// Bit layouts follow the PTX ISA "tcgen05 matrix / instruction descriptor" tables.
#pragma once
#include "common.cuh"

namespace b200 {
namespace umma {

// ---- shared-memory matrix descriptor, K-major operand, 128-byte swizzle ------
// A tile row (one M/N index) is one 128-byte swizzle span; 8 rows form a 1024-byte
// atom; atoms are stacked every SBO = 1024 bytes.  LBO is unused for swizzled
// K-major operands (encoded as 1).  Bits 46-47 = descriptor version 1 (sm_100),
// bits 61-63 = layout type (2 = SWIZZLE_128B).
__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);  // start address, 16-byte units
  d |= (uint64_t)1 << 16;                       // leading byte offset (ignored)
  d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset
  d |= (uint64_t)1 << 46;                       // version
  d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
  return d;
}
// advance the start address by `bytes` inside the swizzle span (K step)
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (bytes >> 4); }

// ---- instruction descriptors ----------------------------------------------------
// kind::f8f6f4 / f16: [4,6) c_format (1=f32) [7,10) a_format [10,13) b_format
// [15] a_major [16] b_major (0 = K-major) [17,23) N>>3 [24,29) M>>4
constexpr uint32_t kFmtE4M3 = 0, kFmtE5M2 = 1;
__host__ __device__ constexpr uint32_t make_idesc_f8(int M, int N, uint32_t afmt = kFmtE4M3,
                                                     uint32_t bfmt = kFmtE4M3) {
  return (1u << 4) | (afmt << 7) | (bfmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// kind::mxf4nvf4 block-scaled: [4,6) b_sf_id [7,10) a_format (E2M1 = 1) [10,13) b_format
// [17,23) N>>3 [23] scale format (0 = UE4M3) [24,29) M>>4 [29,31) a_sf_id [31] K size (0 = K64)
__host__ __device__ constexpr uint32_t make_idesc_nvf4(int M, int N, uint32_t a_sf_id, uint32_t b_sf_id) {
  return (b_sf_id << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (0u << 23) |
         ((uint32_t)(M >> 4) << 24) | (a_sf_id << 29);
}

// ---- TMEM management ---------------------------------------------------------------
template <int kCols> __device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
  static_assert(kCols >= 32 && kCols <= 512 && (kCols & (kCols - 1)) == 0, "TMEM columns: power of two in [32,512]");
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols> __device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- MMA issue (one thread) -------------------------------------------------------
__device__ __forceinline__ void mma_f8f6f4(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_nvf4_block16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4nvf4.block_scale.block16 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(
          tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
// mbarrier arrives once all previously issued MMAs of this thread have completed
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// smem -> TMEM copy of scale factors: 32 rows x 128 bits, replicated to the 4 lane quarters
__device__ __forceinline__ void cp_32x128b_warpx4(uint32_t tmem_addr, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tmem_addr), "l"(smem_desc) : "memory");
}

// ---- TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns ----------
__device__ __forceinline__ void ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// wait::ld that carries a data dependency on the loaded registers, so neither the compiler nor a
// software-pipelined loop can consume them before the asynchronous load has landed
__device__ __forceinline__ void ld_wait_dep(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t.reg .b32 R;\n\t"
      "elect.sync R|P1, 0xFFFFFFFF;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

}  // namespace umma
}  // namespace b200
