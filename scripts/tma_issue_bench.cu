// Micro-benchmark: how does the number of ISSUING warps / lanes affect TMA throughput for small boxes
// (16 rows x 128 B out of a pitch-1024 matrix = one fp8 KV-cache block of one head)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/_bin/tma_issue_bench scripts/tma_issue_bench.cu -lcuda
// Every issuing warp owns STAGES stages; a stage is LANES boxes (one per active lane) on one mbarrier.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void wait(uint64_t* b, uint32_t ph) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0,1,0,p;\n\t}" : "=r"(ok) : "r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma2d(void* dst, const void* tm, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(s32(dst)), "l"((uint64_t)tm), "r"(s32(bar)), "r"(c0), "r"(c1) : "memory");
}
// boxes are numbered over the whole matrix; box b -> rows [16*perm(b) ...), column block (b % 8) * 128
__global__ void __launch_bounds__(512) k(const __grid_constant__ CUtensorMap tm, int warps, int lanes, int stages,
                                         long long boxes_total, long long nblocks) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t* bars = (uint64_t*)(smem + (size_t)warps * stages * lanes * 2048);
  if (threadIdx.x == 0) {
    for (int i = 0; i < warps * stages; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp >= warps) return;
  const long long issuers = (long long)gridDim.x * warps;
  const long long me = (long long)blockIdx.x * warps + warp;
  const long long groups = boxes_total / lanes;  // one group = one stage fill
  const long long per = groups / issuers;
  const long long g0 = me * per, g1 = g0 + per;
  uint8_t* ring = smem + (size_t)warp * stages * lanes * 2048;
  uint64_t* wb = bars + warp * stages;
  auto issue = [&](long long g, int s) {
    if (lane == 0) expect(&wb[s], lanes * 2048);
    __syncwarp();
    if (lane < lanes) {
      const long long b = g * lanes + lane;
      const long long blk = (b * 2654435761ll) % nblocks;  // scattered blocks like a block table
      tma2d(ring + ((size_t)s * lanes + lane) * 2048, &tm, &wb[s], (int)(b & 7) * 128, (int)(blk * 16));
    }
  };
  long long issued = g0;
  for (int s = 0; s < stages && issued < g1; ++s, ++issued) issue(issued, s);
  int s = 0;
  uint32_t ph = 0;
  for (long long g = g0; g < g1; ++g) {
    wait(&wb[s], ph);
    __syncwarp();
    if (issued < g1) issue(issued++, s);
    if (++s == stages) { s = 0; ph ^= 1; }
  }
}
typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t bytes = 2ull << 30;
  uint8_t* buf;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMemset(buf, 1, bytes));
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  EncFn enc = (EncFn)p;
  const int pitch = 1024;
  const uint64_t rows = bytes / pitch;
  CUtensorMap tm;
  memset(&tm, 0, sizeof(tm));
  cuuint64_t dims[2] = {(cuuint64_t)pitch, rows};
  cuuint64_t str[1] = {(cuuint64_t)pitch};
  cuuint32_t box[2] = {128, 16};
  cuuint32_t es[2] = {1, 1};
  if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  struct C { int ctas, warps, lanes, stages; } cs[] = {
      {1, 1, 1, 48}, {1, 1, 8, 6}, {1, 2, 8, 3}, {1, 4, 4, 3}, {1, 8, 2, 3}, {1, 16, 1, 3}, {1, 8, 1, 6}, {1, 4, 1, 12},
      {2, 2, 8, 3}, {2, 4, 4, 3}, {2, 8, 2, 3}, {2, 8, 1, 6}, {3, 4, 1, 4}, {1, 1, 32, 2}, {1, 2, 16, 2},
  };
  const long long boxes_total = (long long)(bytes / 2048) / 4;  // 512 MB per run
  for (auto& c : cs) {
    const int grid = sms * c.ctas;
    const size_t sm = (size_t)c.warps * c.stages * c.lanes * 2048 + 1024;
    float best = 1e9;
    for (int it = 0; it < 3; ++it) {
      CK(cudaEventRecord(e0));
      k<<<grid, c.warps * 32, sm>>>(tm, c.warps, c.lanes, c.stages, boxes_total, (long long)rows / 16);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      CK(cudaGetLastError());
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    const long long groups = boxes_total / c.lanes;
    const double gb = (double)(groups / ((long long)grid * c.warps)) * grid * c.warps * c.lanes * 2048 / 1e9;
    printf("CTAs/SM %d  issuing warps %2d x lanes %2d x stages %2d (%3d KB in flight/SM): %7.1f GB/s  (%.1f GB/s per SM)\n",
           c.ctas, c.warps, c.lanes, c.stages, (int)(c.ctas * c.warps * c.lanes * c.stages * 2), gb / (best * 1e-3), gb / (best * 1e-3) / sms);
  }
  return 0;
}
