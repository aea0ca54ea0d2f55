// Cross-GPU synchronisation primitives shared by the peer-memory kernels (peer_allreduce.cu, fused_quant.cu).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kArThreads = 512;
constexpr int kArMaxWorld = 8;

// Flags are epoch counters, never reset: a signal is ONE remote store (fire and forget, ~1 us over NVLink)
// instead of a compare-and-swap round trip.  CTA b keeps its epoch in its own pad; slot layout per CTA (32 words):
// [phase 0: 8 peer flags][phase 1: 8 peer flags][epoch][unused].
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
constexpr int kArSlotWords = 32;
// all ranks' CTA `blockIdx.x` meet: thread t < world signals rank t and waits for rank t's signal.  The
// bar.sync before the release store makes every thread's earlier (peer) stores part of what is released.
__device__ __forceinline__ void ar_barrier(uint32_t* const* pads, int rank, int world, int phase, uint32_t epoch) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int slot = blockIdx.x * kArSlotWords + phase * kArMaxWorld;
    st_release_sys(pads[threadIdx.x] + slot + rank, epoch);  // my flag in the peer's pad
    const uint32_t* mine = pads[rank] + slot + threadIdx.x;  // the peer's flag in my pad
    uint32_t spins = 0;
    while (ld_acquire_sys(mine) != epoch) {
      if (++spins > (1u << 24)) {
        printf("b200: peer all-reduce barrier timed out (block %d waits for rank %d, phase %d)\n", blockIdx.x,
               threadIdx.x, phase);
        __trap();
      }
    }
  }
  __syncthreads();
}

struct ArArgs {
  void* bufs[kArMaxWorld];
  uint32_t* pads[kArMaxWorld];
  int rank, world;
  int64_t numel;  // 16-bit elements, multiple of 8 * world
};

}  // namespace b200
