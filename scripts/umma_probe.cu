// Probe (sm_100a): does tcgen05.mma kind::f8f6f4 accept an MN-major (transposed) fp8 A operand with the
// 128-byte swizzle, i.e. a V tile exactly as TMA lands it ([token][d] rows of 128 B)?  D[128 d x 16] = V^T * P^T.
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -I. -o scripts/_bin/umma_probe scripts/umma_probe.cu \
//        vllm_b200/csrc/runtime.cu -lcuda && scripts/_bin/umma_probe
// Result on B200: exact for SBO = 1024 (stride between 8-token groups), LBO ignored (tile is one span wide).
#include <cuda_fp8.h>
#include <math.h>
#include <stdlib.h>

#include "vllm_b200/csrc/umma.cuh"

using namespace b200;

__global__ void __launch_bounds__(128) probe(const uint8_t* V /*[128 tok][128 d]*/, const uint8_t* P /*[16][128 tok]*/,
                                             float* out /*[128 d][16]*/, int a_mn_major, uint32_t lbo, uint32_t sbo) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* sA = smem;          // 16 KB
  uint8_t* sB = smem + 16384;  // 2 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 2048);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  // MN-major A: row = token, bytes = d (as the cache holds V).  K-major control: row = d, bytes = token.
  for (int i = tid; i < 128 * 128; i += 128) {
    const int r = i >> 7, c = i & 127;
    const uint8_t v = a_mn_major ? V[r * 128 + c] : V[c * 128 + r];
    sA[r * 128 + ((((c >> 4) ^ (r & 7)) << 4) | (c & 15))] = v;
  }
  for (int i = tid; i < 16 * 128; i += 128) {
    const int r = i >> 7, c = i & 127;
    sB[r * 128 + ((((c >> 4) ^ (r & 7)) << 4) | (c & 15))] = P[r * 128 + c];
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) umma::tmem_alloc<32>(slot);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem = *slot;
  if (tid == 0) {
    uint32_t idesc = umma::make_idesc_f8(128, 16);
    if (a_mn_major) idesc |= 1u << 15;
    for (int ks = 0; ks < 4; ++ks) {
      uint64_t da;
      if (a_mn_major) {
        uint64_t d = 0;
        d |= (uint64_t)((smem_u32(sA + ks * 4096) & 0x3FFFF) >> 4);
        d |= (uint64_t)(lbo >> 4) << 16;
        d |= (uint64_t)(sbo >> 4) << 32;
        d |= (uint64_t)1 << 46;
        d |= (uint64_t)2 << 61;
        da = d;
      } else {
        da = umma::desc_advance(umma::make_desc_k_sw128(smem_u32(sA)), ks * 32);
      }
      const uint64_t db = umma::desc_advance(umma::make_desc_k_sw128(smem_u32(sB)), ks * 32);
      umma::mma_f8f6f4(tmem, da, db, idesc, ks != 0);
    }
    umma::commit(bar);
  }
  mbar_wait(bar, 0);
  umma::fence_after_sync();
  uint32_t r[16];
  umma::ld_32x32b_x16(tmem + ((uint32_t)(warp * 32) << 16), r);
  umma::ld_wait();
  for (int j = 0; j < 16; ++j) out[tid * 16 + j] = __uint_as_float(r[j]);
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc<32>(tmem);
}

int main() {
  static uint8_t hV[128 * 128], hP[16 * 128];
  static float fV[128 * 128], fP[16 * 128], ref[128 * 16], got[128 * 16];
  srand(1);
  for (int i = 0; i < 128 * 128; ++i) {
    __nv_fp8_e4m3 q((float)(rand() % 17 - 8) * 0.25f);
    hV[i] = *reinterpret_cast<uint8_t*>(&q);
    fV[i] = (float)q;
  }
  for (int i = 0; i < 16 * 128; ++i) {
    __nv_fp8_e4m3 q((float)(rand() % 9) * 0.125f);
    hP[i] = *reinterpret_cast<uint8_t*>(&q);
    fP[i] = (float)q;
  }
  for (int d = 0; d < 128; ++d)
    for (int h = 0; h < 16; ++h) {
      float acc = 0;
      for (int k = 0; k < 128; ++k) acc += fV[k * 128 + d] * fP[h * 128 + k];
      ref[d * 16 + h] = acc;
    }
  uint8_t *dV, *dP;
  float* dO;
  cudaMalloc(&dV, sizeof(hV));
  cudaMalloc(&dP, sizeof(hP));
  cudaMalloc(&dO, sizeof(got));
  cudaMemcpy(dV, hV, sizeof(hV), cudaMemcpyHostToDevice);
  cudaMemcpy(dP, hP, sizeof(hP), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  struct { int mn; uint32_t lbo, sbo; const char* name; } cases[] = {
      {0, 0, 0, "K-major A (control)"},        {1, 1024, 1024, "MN-major A, LBO=SBO=1024"},
      {1, 16, 1024, "MN-major A, LBO=16 SBO=1024"}, {1, 1024, 16, "MN-major A, LBO=1024 SBO=16"},
  };
  for (auto& c : cases) {
    cudaMemset(dO, 0, sizeof(got));
    probe<<<1, 128, 32768>>>(dV, dP, dO, c.mn, c.lbo, c.sbo);
    const cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(got, dO, sizeof(got), cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int i = 0; i < 128 * 16; ++i) maxerr = fmax(maxerr, fabs(got[i] - ref[i]));
    printf("%-34s max abs err %g  (%s)\n", c.name, maxerr, cudaGetErrorString(e));
    if (e != cudaSuccess) break;
  }
  return 0;
}
