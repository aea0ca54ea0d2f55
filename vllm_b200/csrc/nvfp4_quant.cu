// NVFP4 activation quantisation: e2m1 codes + one e4m3 scale per 16 elements.
// Reference behaviour: csrc/libtorch_stable/quantization/fp4/nvfp4_quant_kernels.cu:39-113
// (swizzled SF), :115-186 (row-major SF), launcher :188-281; arithmetic
// csrc/libtorch_stable/quantization/fp4/nvfp4_utils.cuh:221-302; SF offset :168-199.
//
// The arithmetic sequence (rcp.approx.ftz(6), e4m3 RN-satfinite scale,
// rcp.approx.ftz of the dequantised scale, cvt.rn.satfinite.e2m1x2) is kept
// operation-for-operation so codes and scales are bit-identical to the
// reference kernel on the same GPU.
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

template <typename T, bool kSwizzled>
__global__ void __launch_bounds__(256) nvfp4_quant_kernel(uint32_t* __restrict__ out, uint8_t* __restrict__ sf_out,
                                                          const T* __restrict__ in,
                                                          const float* __restrict__ global_scale, int m, int n,
                                                          int out_n, int rows_total, int vecs_total) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  // rows_total / vecs_total include the padding the swizzled layout must zero.
  const float gs = global_scale ? global_scale[0] : 1.0f;
  const float rcp6 = rcp_approx_ftz(6.0f);
  const float rcp_gs = rcp_approx_ftz(gs);
  const int num_k_tiles = (out_n + 63) / 64;
  const int sf_cols = out_n / 16;
  const int vecs_rounded = (vecs_total + 31) & ~31;  // keep warps converged for the shuffle
  for (int row = blockIdx.x; row < rows_total; row += gridDim.x) {
    for (int cv = blockIdx.y * blockDim.x + threadIdx.x; cv < vecs_rounded; cv += gridDim.y * blockDim.x) {
      const int e0 = cv * 8;
      const bool valid_in = (row < m) && (e0 < n);
      uint4 raw = make_uint4(0, 0, 0, 0);
      if (valid_in) raw = __ldcg(reinterpret_cast<const uint4*>(in + (int64_t)row * n + e0));
      const T* x = reinterpret_cast<const T*>(&raw);
      float f[8];
      float amax = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f[i] = to_float(x[i]);
        amax = fmaxf(amax, fabsf(f[i]));
      }
      amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
      float sf = gs * (amax * rcp6);
      const __nv_fp8_storage_t sf8 = __nv_cvt_float_to_fp8(sf, __NV_SATFINITE, __NV_E4M3);
      sf = __half2float(__half(__nv_cvt_fp8_to_halfraw(sf8, __NV_E4M3)));
      const float oscale = sf != 0.f ? rcp_approx_ftz(sf * rcp_gs) : 0.f;
      if (cv < vecs_total) {
        if ((cv & 1) == 0) {
          const int kb = cv >> 1;
          if constexpr (kSwizzled) sf_out[sf_swizzled_offset(row, kb, num_k_tiles)] = (uint8_t)sf8;
          else if (row < m && kb < sf_cols) sf_out[(int64_t)row * sf_cols + kb] = (uint8_t)sf8;
        }
        if (row < m && e0 < out_n) {
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] *= oscale;
          out[((int64_t)row * out_n + e0) >> 3] = pack_e2m1x8(f);
        }
      }
    }
  }
}

}  // namespace b200

extern "C" int b200_scaled_fp4_quant(void* output, void* output_sf, const void* input,
                                     const float* input_global_scale, int m, int n, int out_n, int swizzled,
                                     int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(m >= 0 && n > 0, B200_ERR_INVALID_ARG, "scaled_fp4_quant: bad sizes");
  B200_CHECK_ARG(n % 16 == 0, B200_ERR_INVALID_ARG, "The N dimension must be multiple of 16.");
  B200_CHECK_ARG(out_n % 16 == 0 && out_n >= n, B200_ERR_INVALID_ARG,
                 "The output tensor width must be a multiple of 16.");
  B200_CHECK_ARG(dtype == B200_DT_F16 || dtype == B200_DT_BF16, B200_ERR_INVALID_ARG,
                 "Unsupported input data type for quantize_to_fp4.");
  B200_CHECK_ARG(output && output_sf && (input || m == 0), B200_ERR_INVALID_ARG, "scaled_fp4_quant: null pointer");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(input) % 16 == 0 && reinterpret_cast<uintptr_t>(output) % 4 == 0,
                 B200_ERR_ALIGNMENT, "scaled_fp4_quant: input must be 16-byte aligned");
  const int rows_total = swizzled ? ((m + 127) / 128) * 128 : m;
  if (rows_total == 0) return 0;
  const int sf_cols_padded = swizzled ? ((out_n / 16 + 3) / 4) * 4 : out_n / 16;
  const int vecs_total = sf_cols_padded * 2;
  const int threads = vecs_total >= 256 ? 256 : ((vecs_total + 31) / 32) * 32;
  int gy = (vecs_total + threads - 1) / threads;
  if (gy > 8) gy = 8;
  int gx = (sm_count() * 8) / gy;
  if (gx > rows_total) gx = rows_total;
  if (gx < 1) gx = 1;
  dim3 grid(gx, gy), block(threads);
  cudaStream_t s = (cudaStream_t)stream;
#define B200_Q4(T, SW)                                                                                             \
  launch_kernel(nvfp4_quant_kernel<T, SW>, dim3(grid), dim3(block), 0, s, 1, (uint32_t*)output, (uint8_t*)output_sf, (const T*)input,        \
                                                   input_global_scale, m, n, out_n, rows_total, vecs_total)
  if (dtype == B200_DT_BF16) {
    if (swizzled) B200_Q4(__nv_bfloat16, true);
    else B200_Q4(__nv_bfloat16, false);
  } else {
    if (swizzled) B200_Q4(__half, true);
    else B200_Q4(__half, false);
  }
#undef B200_Q4
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}
