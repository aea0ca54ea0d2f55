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
// Spin budget of one barrier wait, host-configurable (b200_peer_set_barrier_timeout): 0 = wait for ever.  The
// default (~2^28 polls, minutes) only exists so that a protocol bug or a dead peer ends in a trap with a message
// instead of a GPU that has to be reset; legitimately skewed ranks (one still capturing a graph) fit well inside it.
unsigned long long ar_spin_limit();  // host: current budget, copied into ArArgs::spin_limit at launch
// all ranks' CTA `blockIdx.x` meet: thread t < world signals rank t and waits for rank t's signal.  The
// bar.sync before the release store makes every thread's earlier (peer) stores part of what is released.
// Epochs only grow, and the wait is a wraparound-safe ">= epoch": the one-shot kernels have no end barrier, so
// a fast peer may already have stored epoch n+1 into this slot before a stalled waiter polls for n.
__device__ __forceinline__ void ar_barrier(uint32_t* const* pads, int rank, int world, int phase, uint32_t epoch,
                                           unsigned long long limit) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int slot = blockIdx.x * kArSlotWords + phase * kArMaxWorld;
    st_release_sys(pads[threadIdx.x] + slot + rank, epoch);  // my flag in the peer's pad
    const uint32_t* mine = pads[rank] + slot + threadIdx.x;  // the peer's flag in my pad
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (limit && ++spins > limit) {
        printf("b200: peer all-reduce barrier timed out (block %d waits for rank %d, phase %d, epoch %u)\n",
               blockIdx.x, threadIdx.x, phase, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}

// In-switch reduction (NVLS): one load returns the SUM over all ranks' copies of 8 consecutive 16-bit values,
// accumulated in fp32 inside the NVSwitch and rounded once to T.  `mc` is the multicast address.
template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce_add(const void* mc);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_add<__nv_bfloat16>(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_add<__half>(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}

// Multicast store: one store lands in EVERY rank's copy (the switch replicates it).
template <typename T> __device__ __forceinline__ void multimem_st(void* mc, const uint4& v);
template <> __device__ __forceinline__ void multimem_st<__nv_bfloat16>(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
template <> __device__ __forceinline__ void multimem_st<__half>(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

struct ArArgs;
__device__ __forceinline__ int ar_rank(const ArArgs& a);

struct ArArgs {
  void* bufs[kArMaxWorld];
  uint32_t* pads[kArMaxWorld];
  int rank, world;
  int64_t numel;  // 16-bit elements, multiple of 8 * world
  unsigned long long spin_limit;  // barrier spin budget, 0 = unbounded
  // Single-GPU emulation of `world` ranks (tests on a 1-GPU box): ONE cooperative launch with gridDim.y = world,
  // rank = blockIdx.y, every "rank's" buffer / signal pad on the same device; per-rank outputs are `emu_stride`
  // elements apart.  0 in production.
  const void* mc;  // NVLS multicast mapping of the buffers (torch symmetric memory multicast_ptr + offset) or null
  int emu;
  int64_t emu_stride;
};

__device__ __forceinline__ int ar_rank(const ArArgs& a) { return a.emu ? (int)blockIdx.y : a.rank; }

}  // namespace b200
