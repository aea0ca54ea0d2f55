// NVFP4 quantisation arithmetic shared by nvfp4_quant.cu and the fused all-reduce epilogue (lamport_allreduce.cu).
// Reference: csrc/libtorch_stable/quantization/fp4/nvfp4_utils.cuh:221-302 (math), :168-199 (scale-factor offset);
// kept operation for operation so codes and scales are bit-identical to the reference kernel on the same GPU.
#pragma once
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ float rcp_approx_ftz(float a) {
  float b;
  asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(b) : "f"(a));
  return b;
}
// 8 floats -> 8 e2m1 codes packed in a u32 (element 0 in the lowest nibble)
__device__ __forceinline__ uint32_t pack_e2m1x8(const float (&x)[8]) {
  uint32_t val;
  asm volatile(
      "{\n"
      ".reg .b8 b0, b1, b2, b3;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b0, %2, %1;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b1, %4, %3;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b2, %6, %5;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b3, %8, %7;\n"
      "mov.b32 %0, {b0, b1, b2, b3};\n"
      "}"
      : "=r"(val)
      : "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]), "f"(x[4]), "f"(x[5]), "f"(x[6]), "f"(x[7]));
  return val;
}

// offset (bytes) of SF (row, kblock) in the [mTile][kTile][32][4][4] layout
__device__ __forceinline__ int64_t sf_swizzled_offset(int row, int kb, int num_k_tiles) {
  return ((int64_t)(row >> 7) * num_k_tiles + (kb >> 2)) * 512 + ((row & 31) << 4) + (((row >> 5) & 3) << 2) +
         (kb & 3);
}

// One thread's 8 elements (fp32 copies of T values) -> 8 e2m1 codes; lane pairs (xor 1) share one e4m3 scale.
// Every lane of the warp must call this (shuffle).  *sf8 = the scale byte (same in both lanes of the pair).
__device__ __forceinline__ uint32_t nvfp4_quant8(float (&f)[8], float gs, float rcp6, float rcp_gs, uint8_t* sf8_out) {
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(f[i]));
  amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
  float sf = gs * (amax * rcp6);
  const __nv_fp8_storage_t sf8 = __nv_cvt_float_to_fp8(sf, __NV_SATFINITE, __NV_E4M3);
  sf = __half2float(__half(__nv_cvt_fp8_to_halfraw(sf8, __NV_E4M3)));
  const float oscale = sf != 0.f ? rcp_approx_ftz(sf * rcp_gs) : 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] *= oscale;
  *sf8_out = (uint8_t)sf8;
  return pack_e2m1x8(f);
}

}  // namespace b200
