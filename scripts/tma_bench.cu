// Micro-benchmark: how fast can one B200 stream HBM into shared memory, as a function of the copy
// shape?  Informs the weight layout of the decode GEMMs and the KV-cache layout of paged attention.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_bench scripts/tma_bench.cu -lcuda && ./tma_bench
// mode 0: cp.async.bulk.tensor.2d, box = 128 B x ROWS rows out of a [rows][pitch] matrix (K-major operand tile)
// mode 1: cp.async.bulk (1-D), CHUNK contiguous bytes (pre-tiled operand / HND page)
// Each CTA owns a contiguous slice of tiles and keeps STAGES copies in flight; one thread waits on the
// mbarrier of a stage and immediately re-issues.  Nothing is computed.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e = (x);                                                           \
    if (e != cudaSuccess) {                                                        \
      printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__);            \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c));
}
__device__ __forceinline__ void expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void wait(uint64_t* b, uint32_t ph) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0,1,0,p;\n\t}"
                 : "=r"(ok)
                 : "r"(s32(b)), "r"(ph)
                 : "memory");
}
__device__ __forceinline__ void tma2d(void* dst, const void* tm, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                   s32(dst)),
               "l"((uint64_t)tm), "r"(s32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)),
               "l"((uint64_t)src), "r"(bytes), "r"(s32(bar))
               : "memory");
}

// tiles are numbered t = rowblock * kblocks + kb (K-major walk like the GEMM mainloop)
__global__ void __launch_bounds__(128) stream_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* base,
                                                     int mode, int rows_per_box, int chunk, int stages,
                                                     long long tiles_total, int kblocks, int boxes_per_stage) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = (uint64_t*)(smem + (size_t)stages * chunk * boxes_per_stage);
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const long long per = (tiles_total + gridDim.x - 1) / gridDim.x;
  const long long t0 = per * blockIdx.x, t1 = min(tiles_total, t0 + per);
  auto issue = [&](long long t, int s) {
    expect(&bars[s], (uint32_t)chunk * boxes_per_stage);
    for (int b = 0; b < boxes_per_stage; ++b) {
      const long long tt = t * boxes_per_stage + b;
      uint8_t* dst = smem + ((size_t)s * boxes_per_stage + b) * chunk;
      if (mode == 0) {
        const long long rb = tt / kblocks;
        const int kb = (int)(tt - rb * kblocks);
        tma2d(dst, &tm, &bars[s], kb * 128, (int)(rb * rows_per_box));
      } else {
        bulk1d(dst, base + tt * (long long)chunk, chunk, &bars[s]);
      }
    }
  };
  long long issued = t0;
  for (int s = 0; s < stages && issued < t1; ++s, ++issued) issue(issued, s);
  int s = 0;
  uint32_t ph = 0;
  for (long long t = t0; t < t1; ++t) {
    wait(&bars[s], ph);
    if (issued < t1) issue(issued++, s);
    if (++s == stages) {
      s = 0;
      ph ^= 1;
    }
  }
}

typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                          const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                          CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  int dev = 0, sms = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t bytes = 2ull << 30;  // 2 GiB >> L2
  uint8_t* buf;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMemset(buf, 1, bytes));
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  EncFn enc = (EncFn)p;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));

  struct Case { int mode, rows, pitch, chunk, stages, ctas_per_sm, boxes, swz, promo; const char* name; int grid; };
  Case cases[] = {
      {0, 128, 4096, 16384, 6, 1, 1, 1, 1, "tensor 128Bx128rows pitch4096 6st x1 (GEMM v2 weights)"},
      {0, 128, 4096, 16384, 12, 1, 1, 1, 1, "tensor 128Bx128rows pitch4096 12st x1"},
      {0, 128, 4096, 16384, 6, 2, 1, 1, 1, "tensor 128Bx128rows pitch4096 6st x2 CTAs/SM"},
      {0, 128, 4096, 16384, 6, 1, 1, 1, 0, "tensor 128Bx128rows pitch4096 6st no L2 promotion"},
      {0, 128, 14336, 16384, 6, 1, 1, 1, 1, "tensor 128Bx128rows pitch14336 6st (down_proj)"},
      {0, 64, 4096, 8192, 12, 1, 1, 1, 1, "tensor 128Bx64rows pitch4096 12st"},
      {0, 32, 4096, 4096, 24, 1, 1, 1, 1, "tensor 128Bx32rows pitch4096 24st"},
      {0, 16, 1024, 2048, 16, 3, 1, 1, 1, "tensor 128Bx16rows pitch1024 16st x3 (fp8 KV NHD chunk)"},
      {0, 16, 1024, 2048, 32, 3, 1, 1, 1, "tensor 128Bx16rows pitch1024 32st x3"},
      {0, 16, 128, 2048, 16, 3, 1, 1, 1, "tensor 128Bx16rows pitch128 16st x3 (fp8 KV HND page, contiguous)"},
      {1, 0, 0, 16384, 6, 1, 1, 0, 0, "bulk1d 16KB 6st x1 (pre-tiled weights)"},
      {1, 0, 0, 16384, 12, 1, 1, 0, 0, "bulk1d 16KB 12st x1"},
      {1, 0, 0, 32768, 6, 1, 1, 0, 0, "bulk1d 32KB 6st x1"},
      {1, 0, 0, 8192, 12, 1, 1, 0, 0, "bulk1d 8KB 12st x1"},
      {1, 0, 0, 2048, 16, 3, 1, 0, 0, "bulk1d 2KB 16st x3 (HND page)"},
      {1, 0, 0, 4096, 16, 3, 1, 0, 0, "bulk1d 4KB 16st x3"},
      {1, 0, 0, 16384, 6, 2, 1, 0, 0, "bulk1d 16KB 6st x2 CTAs/SM"},
      // latency (1 copy in flight) and per-SM rate when only a few SMs stream (grid given explicitly)
      {0, 128, 4096, 16384, 1, 1, 1, 1, 1, "LAT tensor 128Bx128rows pitch4096 1st grid148", 148},
      {1, 0, 0, 16384, 1, 1, 1, 0, 0, "LAT bulk1d 16KB 1st grid148", 148},
      {0, 128, 4096, 16384, 1, 1, 1, 1, 1, "LAT tensor 128Bx128rows pitch4096 1st grid8", 8},
      {1, 0, 0, 16384, 1, 1, 1, 0, 0, "LAT bulk1d 16KB 1st grid8", 8},
      {1, 0, 0, 2048, 1, 1, 1, 0, 0, "LAT bulk1d 2KB 1st grid8", 8},
      {0, 128, 4096, 16384, 6, 1, 1, 1, 1, "tensor 128Bx128rows pitch4096 6st grid32", 32},
      {0, 128, 4096, 16384, 12, 1, 1, 1, 1, "tensor 128Bx128rows pitch4096 12st grid32", 32},
      {1, 0, 0, 16384, 6, 1, 1, 0, 0, "bulk1d 16KB 6st grid32", 32},
      {1, 0, 0, 16384, 12, 1, 1, 0, 0, "bulk1d 16KB 12st grid32", 32},
      {0, 128, 4096, 16384, 12, 1, 1, 1, 1, "tensor 128Bx128rows pitch4096 12st grid74", 74},
      {1, 0, 0, 16384, 12, 1, 1, 0, 0, "bulk1d 16KB 12st grid74", 74},
  };
  for (auto& c : cases) {
    CUtensorMap tm;
    memset(&tm, 0, sizeof(tm));
    long long tiles;
    int kblocks = 1;
    if (c.mode == 0) {
      const uint64_t rows = bytes / c.pitch;
      cuuint64_t dims[2] = {(cuuint64_t)c.pitch, rows};
      cuuint64_t str[1] = {(cuuint64_t)c.pitch};
      cuuint32_t box[2] = {128, (cuuint32_t)c.rows};
      cuuint32_t es[2] = {1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       c.swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                       c.promo ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        printf("encode failed %d for %s\n", (int)r, c.name);
        continue;
      }
      kblocks = c.pitch / 128;
      tiles = (long long)(rows / c.rows) * kblocks;
    } else {
      tiles = bytes / c.chunk;
    }
    const int grid = c.grid ? c.grid : sms * c.ctas_per_sm;
    if (c.grid) tiles = tiles < (long long)grid * 2048 ? tiles : (long long)grid * 2048;
    const size_t smem = (size_t)c.stages * c.chunk * c.boxes + 1024;
    float best = 1e9;
    for (int it = 0; it < 4; ++it) {
      CK(cudaEventRecord(e0));
      stream_kernel<<<grid, 128, smem>>>(tm, buf, c.mode, c.rows, c.chunk, c.stages, tiles, kblocks, c.boxes);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      CK(cudaGetLastError());
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    const double gb = (double)tiles * c.chunk * c.boxes / 1e9;
    const int active = grid < sms ? grid : sms;
    printf("%-70s %8.1f GB/s  (%.1f GB/s per SM, %d KB in flight/SM, %.2f us per copy per CTA)\n", c.name,
           gb / (best * 1e-3), gb / (best * 1e-3) / active, (int)(c.stages * c.chunk * c.boxes * c.ctas_per_sm / 1024),
           best * 1e3 / ((double)tiles / grid));
  }
  return 0;
}
