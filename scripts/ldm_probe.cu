// Probe (sm_100a): fragment layout of ldmatrix.m16n16.trans.b8 and a known-answer check of
// mma.sync.m16n8k32 e4m3 (which ptxas lowers to F2FP unpacks + HMMA on sm_100a - see the SASS).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o scripts/_bin/ldm_probe scripts/ldm_probe.cu && scripts/_bin/ldm_probe
// Result on B200: lane (g = lane/4, t = lane%4) receives source rows 4t..4t+3 at column g (r0) and at column
// g+8 (r1); .x2 adds the same of a second matrix whose row addresses come from lanes 16-31.
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

__global__ void probe(uint32_t* out) {
  __shared__ __align__(128) uint8_t sm[32 * 16];
  for (int i = threadIdx.x; i < 32 * 16; i += 32) sm[i] = (uint8_t)i;  // byte = row*16 + col (mod 256)
  __syncwarp();
  const int lane = threadIdx.x;
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(sm + (lane & 15) * 16);
  uint32_t r0, r1;
  asm volatile("ldmatrix.sync.aligned.m16n16.x1.trans.shared.b8 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(a));
  out[lane * 2] = r0;
  out[lane * 2 + 1] = r1;
}

// C[16x8] = A[16x32] * B[32x8]; assumed fragments: a0:(g, 4t..4t+3) a1:(g+8, ..) a2:(g, 16+4t..) a3:(g+8, 16+4t..)
// b0:(k 4t.., n g) b1:(k 16+4t.., n g)
__global__ void mma_probe(const uint8_t* A, const uint8_t* B, float* C) {
  const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
  auto ld4 = [](const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); };
  const uint32_t a0 = ld4(A + g * 32 + 4 * t), a1 = ld4(A + (g + 8) * 32 + 4 * t), a2 = ld4(A + g * 32 + 16 + 4 * t),
                 a3 = ld4(A + (g + 8) * 32 + 16 + 4 * t);
  const uint32_t b0 = ld4(B + g * 32 + 4 * t), b1 = ld4(B + g * 32 + 16 + 4 * t);  // B stored [n][k]
  float c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c0), "+f"(c1), "+f"(c2), "+f"(c3)
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  C[g * 8 + 2 * t] = c0;
  C[g * 8 + 2 * t + 1] = c1;
  C[(g + 8) * 8 + 2 * t] = c2;
  C[(g + 8) * 8 + 2 * t + 1] = c3;
}

int main() {
  uint32_t* d;
  cudaMalloc(&d, 4 * 64);
  probe<<<1, 32>>>(d);
  uint32_t h[64];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("ldmatrix.m16n16.x1.trans.b8: per lane the (source row, source col) of each byte of r0 | r1\n");
  for (int l = 0; l < 32; ++l) {
    printf("lane %2d:", l);
    for (int r = 0; r < 2; ++r) {
      for (int b = 0; b < 4; ++b) {
        const int v = (h[l * 2 + r] >> (8 * b)) & 255;
        printf(" (%2d,%2d)", v / 16, v % 16);
      }
      printf(" |");
    }
    printf("\n");
  }
  uint8_t hA[16 * 32], hB[8 * 32];
  float fA[16 * 32], fB[8 * 32];
  for (int i = 0; i < 16 * 32; ++i) {
    __nv_fp8_e4m3 q((float)((i * 7 + 3) % 9 - 4) * 0.5f);
    hA[i] = *reinterpret_cast<uint8_t*>(&q);
    fA[i] = (float)q;
  }
  for (int i = 0; i < 8 * 32; ++i) {
    __nv_fp8_e4m3 q((float)((i * 5 + 1) % 7 - 3) * 0.25f);
    hB[i] = *reinterpret_cast<uint8_t*>(&q);
    fB[i] = (float)q;
  }
  uint8_t *dA, *dB;
  float* dC;
  cudaMalloc(&dA, sizeof(hA));
  cudaMalloc(&dB, sizeof(hB));
  cudaMalloc(&dC, 16 * 8 * 4);
  cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB, sizeof(hB), cudaMemcpyHostToDevice);
  mma_probe<<<1, 32>>>(dA, dB, dC);
  float hC[128];
  cudaMemcpy(hC, dC, sizeof(hC), cudaMemcpyDeviceToHost);
  double maxerr = 0;
  for (int i = 0; i < 16; ++i)
    for (int n = 0; n < 8; ++n) {
      float ref = 0;
      for (int k = 0; k < 32; ++k) ref += fA[i * 32 + k] * fB[n * 32 + k];
      maxerr = fmax(maxerr, fabs(ref - hC[i * 8 + n]));
    }
  printf("mma m16n8k32 e4m3 known-answer max abs err = %g (%s)\n", maxerr, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
