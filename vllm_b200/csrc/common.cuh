// Shared device helpers for libb200vllm (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <type_traits>

#include "../../include/b200_vllm.h"

namespace b200 {

// ---- host-side error plumbing ------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
#define B200_CHECK_ARG(cond, code, ...)   \
  do {                                    \
    if (!(cond)) {                        \
      ::b200::set_error(__VA_ARGS__);     \
      return (code);                      \
    }                                     \
  } while (0)
#define B200_CHECK_LAUNCH()                                              \
  do {                                                                   \
    cudaError_t e__ = cudaGetLastError();                                \
    if (e__ != cudaSuccess) {                                            \
      ::b200::set_error("CUDA launch failed at %s:%d: %s", __FILE__,     \
                        __LINE__, cudaGetErrorString(e__));              \
      return B200_ERR_CUDA;                                              \
    }                                                                    \
  } while (0)

int sm_count();
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device (per-context) property of a kernel: remembered
// per (kernel, device) so a process that drives several GPUs opts in on each of them.  Returns a B200_* code.
int ensure_dyn_smem(const void* kernel, int bytes);

// ---- launch helper: programmatic dependent launch (PDL) ---------------------------------------
// Every kernel of this library is launched with the programmatic-stream-serialization attribute, so the
// NEXT kernel on the stream may be scheduled while this one drains: its CTAs run their prologue (barrier
// init, TMEM allocation, descriptor prefetch, weight prefetch) and then block in pdl_wait() until all
// memory effects of the preceding grid are visible.  The rule each kernel follows: no global-memory access
// that depends on, or could race with, an earlier kernel before pdl_wait().  Kernels that do not belong to
// this library never trigger early, so mixing with them degenerates to ordinary stream order.
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel_ex(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                    int cluster_x, bool cooperative, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[3];
  int n = 0;
  if (cluster_x > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = cluster_x;
    at[n].val.clusterDim.y = 1;
    at[n].val.clusterDim.z = 1;
    ++n;
  }
  if (cooperative) {  // all CTAs co-resident or the launch fails (single-GPU emulation of cross-rank barriers)
    at[n].id = cudaLaunchAttributeCooperative;
    at[n].val.cooperative = 1;
    ++n;
  } else if (pdl_enabled()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 int cluster_x, Args&&... args) {
  return launch_kernel_ex(kern, grid, block, smem, stream, cluster_x, false, static_cast<Args&&>(args)...);
}

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- scalar conversions --------------------------------------------------
template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- shared-memory address / mbarrier / TMA PTX ---------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// 1024-byte aligned start of the dynamic shared memory.  Plain pointer arithmetic on the __shared__ array,
// so the compiler keeps the shared address space (LDS/STS); a round trip through uintptr_t makes every
// access through the result a generic LD/ST, ~2x the latency and ordered against all other memory ops.
__device__ __forceinline__ uint8_t* align_smem_1024(uint8_t* raw) {
  return raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps instead of hanging the GPU box.  The bound is WALL TIME (%globaltimer, read
// once per 2^20 polls), not a poll count: a kernel launched with programmatic dependent launch legitimately sits in
// its first wait for as long as its predecessor runs, and the predecessor may be an all-reduce waiting for a rank
// that is seconds behind (capturing a graph, starting a profiler - seen at TP8).
constexpr unsigned long long kMbarTimeoutNs = 120ull * 1000000000ull;
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
#ifdef B200_UNBOUNDED_WAIT
  while (!mbar_try_wait(bar, parity)) {
  }
#else
  uint32_t spins = 0;
  unsigned long long t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0xFFFFFu) == 0) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kMbarTimeoutNs) {
        printf("b200: mbarrier wait timed out (block %d,%d,%d thread %d parity %u, barrier at shared 0x%x)\n",
               blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x, parity, smem_u32(bar));
        __trap();
      }
    }
  }
#endif
}

__device__ __forceinline__ void prefetch_tensormap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
// L2 cache-policy words (same encodings CUTLASS uses for TMA::CacheHintSm90)
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                            uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                            int c2, int c3, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* tmap, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tmap)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---- legacy-shape tensor-core helpers (used by the HBM-bound decode kernel) -
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// two transposed 16x16 byte matrices: lane (g = lane/4, t = lane%4) receives source rows 4t..4t+3 at column g
// (r0), at column g+8 (r1), and the same of the second matrix (r2, r3); lane L supplies the address of source
// row L%16 of matrix L/16 (layout checked on the device by scripts/ldm_probe.cu).  sm_100a only.
__device__ __forceinline__ void ldmatrix_b8_x2_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// D(16x8,f32) += A(16x16) * B(16x8), bf16 or f16 operands.
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

// four floats -> four e4m3 bytes (x0 in the low byte), round-to-nearest, saturating
__device__ __forceinline__ uint32_t pack4_e4m3(float x0, float x1, float x2, float x3) {
  const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(x0, x1), __NV_SATFINITE, __NV_E4M3);
  const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(x2, x3), __NV_SATFINITE, __NV_E4M3);
  return lo | (hi << 16);
}
__device__ __forceinline__ float e4m3_to_float(uint32_t byte) {
  const __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)byte, __NV_E4M3);
  return __half2float(*reinterpret_cast<const __half*>(&h));
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- host: TMA descriptor encode through the driver entry point -----------
// Returns 0 on success. `tmap_out` points at 128 bytes (CUtensorMap).
enum TmaSwizzle { kSwizzleNone = 0, kSwizzle32 = 1, kSwizzle64 = 2, kSwizzle128 = 3 };
enum TmaDtype { kTmaU8 = 0, kTmaU16 = 1, kTmaBf16 = 2, kTmaF16 = 3, kTmaF32 = 4, kTmaU4x16Align8 = 5 };
int encode_tmap(void* tmap_out, TmaDtype dt, int rank, const void* base, const uint64_t* dims,
                const uint64_t* strides_bytes /* rank-1 */, const uint32_t* box, TmaSwizzle sw,
                bool l2_promote_256 = true);

}  // namespace b200
