// NVFP4 activation quantisation: e2m1 codes + one e4m3 scale per 16 elements.
// Reference behaviour: csrc/libtorch_stable/quantization/fp4/nvfp4_quant_kernels.cu:39-113
// (swizzled SF), :115-186 (row-major SF), launcher :188-281; arithmetic
// csrc/libtorch_stable/quantization/fp4/nvfp4_utils.cuh:221-302; SF offset :168-199.
//
// The arithmetic sequence (rcp.approx.ftz(6), e4m3 RN-satfinite scale,
// rcp.approx.ftz of the dequantised scale, cvt.rn.satfinite.e2m1x2) is kept
// operation-for-operation so codes and scales are bit-identical to the
// reference kernel on the same GPU.
#include "nvfp4_quant.cuh"

namespace b200 {

// kFuse = 0: in is [m, n].  kFuse = 1 (silu_and_mul_nvfp4_quant): in is [m, 2n] = [gate | up] and the quantised
// value is T(silu(gate) * up) with the product formed in fp32 and rounded to T ONCE (the reference's fused kernel,
// activation_nvfp4_quant_fusion_kernels.cu:39-113 + nvfp4_utils.cuh:304-329; note this is one rounding fewer than
// silu_and_mul followed by scaled_fp4_quant).
template <typename T, bool kSwizzled, int kFuse>
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
  const int64_t in_stride = kFuse == 1 ? 2 * (int64_t)n : n;
  for (int row = blockIdx.x; row < rows_total; row += gridDim.x) {
    for (int cv = blockIdx.y * blockDim.x + threadIdx.x; cv < vecs_rounded; cv += gridDim.y * blockDim.x) {
      const int e0 = cv * 8;
      const bool valid_in = (row < m) && (e0 < n);
      uint4 raw = make_uint4(0, 0, 0, 0);
      if (valid_in) raw = __ldcg(reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + e0));
      const T* x = reinterpret_cast<const T*>(&raw);
      float f[8];
      if constexpr (kFuse == 1) {
        uint4 raw2 = make_uint4(0, 0, 0, 0);
        if (valid_in) raw2 = __ldcg(reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + n + e0));
        const T* u = reinterpret_cast<const T*>(&raw2);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float g = to_float(x[i]);
          const float s = __fdividef(g, 1.f + __expf(-g));
          f[i] = to_float(from_float<T>(s * to_float(u[i])));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = to_float(x[i]);
      }
      uint8_t sf8;
      const uint32_t codes = nvfp4_quant8(f, gs, rcp6, rcp_gs, &sf8);
      if (cv < vecs_total) {
        if ((cv & 1) == 0) {
          const int kb = cv >> 1;
          if constexpr (kSwizzled) sf_out[sf_swizzled_offset(row, kb, num_k_tiles)] = sf8;
          else if (row < m && kb < sf_cols) sf_out[(int64_t)row * sf_cols + kb] = sf8;
        }
        if (row < m && e0 < out_n) out[((int64_t)row * out_n + e0) >> 3] = codes;
      }
    }
  }
}

// (fused add +) RMSNorm + NVFP4 quant, one CTA of 256 threads per row.  No reference kernel fuses these two; the
// result is bit-identical to b200_rms_norm (merge_elementwise.cu rms_norm_kernel: same element-to-thread map and
// summation order for the variance, same roundings) followed by b200_scaled_fp4_quant with swizzled scales.
// Rows m .. rows_total-1 (the 128-row padding of the scale layout) only zero their scales.
template <typename T>
__global__ void __launch_bounds__(256) rms_norm_nvfp4_quant_kernel(uint32_t* __restrict__ out,
                                                                   uint8_t* __restrict__ sf_out,
                                                                   const T* __restrict__ in, const T* __restrict__ w,
                                                                   T* __restrict__ residual,
                                                                   const float* __restrict__ global_scale, int m,
                                                                   int hidden, int64_t in_stride, float eps,
                                                                   int vecs_total) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sh[8];
  extern __shared__ __align__(16) uint8_t row_buf[];
  float* rowf = reinterpret_cast<float*>(row_buf);
  const int row = blockIdx.x;
  const int num_k_tiles = (hidden + 63) / 64;
  const int vecs_rounded = (vecs_total + 31) & ~31;
  if (row >= m) {
    for (int cv = threadIdx.x; cv < vecs_total; cv += blockDim.x)
      if ((cv & 1) == 0) sf_out[sf_swizzled_offset(row, cv >> 1, num_k_tiles)] = 0;
    return;
  }
  const float gs = global_scale ? global_scale[0] : 1.0f;
  const float rcp6 = rcp_approx_ftz(6.0f);
  const float rcp_gs = rcp_approx_ftz(gs);
  const T* src = in + (int64_t)row * in_stride;
  T* res = residual ? residual + (int64_t)row * hidden : nullptr;
  const int nv = hidden / 8;
  for (int v = threadIdx.x; v < nv; v += blockDim.x) {  // 16-byte loads; the sum below runs in rms_norm's order
    const uint4 a = *reinterpret_cast<const uint4*>(src + v * 8);
    const T* av = reinterpret_cast<const T*>(&a);
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = to_float(av[i]);
    if (res) {
      const uint4 r = *reinterpret_cast<const uint4*>(res + v * 8);
      const T* rv = reinterpret_cast<const T*>(&r);
      uint4 o;
      T* ov = reinterpret_cast<T*>(&o);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        ov[i] = from_float<T>(x[i] + to_float(rv[i]));  // the sum is rounded to T before the norm
        x[i] = to_float(ov[i]);
      }
      *reinterpret_cast<uint4*>(res + v * 8) = o;
    }
    *reinterpret_cast<float4*>(rowf + v * 8) = make_float4(x[0], x[1], x[2], x[3]);
    *reinterpret_cast<float4*>(rowf + v * 8 + 4) = make_float4(x[4], x[5], x[6], x[7]);
  }
  __syncthreads();
  float ss = 0.f;
  for (int e = threadIdx.x; e < hidden; e += blockDim.x) ss += rowf[e] * rowf[e];
  {  // block_sum of merge_elementwise.cu
    ss = warp_sum(ss);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) sh[warp] = ss;
    __syncthreads();
    float r = lane < 8 ? sh[lane] : 0.f;
    ss = warp_sum(r);
  }
  const float inv = rsqrtf(ss / hidden + eps);
  for (int cv = threadIdx.x; cv < vecs_rounded; cv += blockDim.x) {
    float f[8];
    if (cv < nv) {
      const uint4 wr = *reinterpret_cast<const uint4*>(w + cv * 8);
      const T* wv = reinterpret_cast<const T*>(&wr);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        f[i] = to_float(from_float<T>(to_float(from_float<T>(rowf[cv * 8 + i] * inv)) * to_float(wv[i])));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = 0.f;
    }
    uint8_t sf8;
    const uint32_t codes = nvfp4_quant8(f, gs, rcp6, rcp_gs, &sf8);
    if (cv < vecs_total) {
      if ((cv & 1) == 0) sf_out[sf_swizzled_offset(row, cv >> 1, num_k_tiles)] = sf8;
      if (cv < nv) out[((int64_t)row * hidden >> 3) + cv] = codes;
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
  launch_kernel(nvfp4_quant_kernel<T, SW, 0>, dim3(grid), dim3(block), 0, s, 1, (uint32_t*)output, (uint8_t*)output_sf, (const T*)input,        \
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

extern "C" int b200_silu_and_mul_nvfp4_quant(void* result, void* result_block_scale, const void* input,
                                             const float* input_global_scale, int m, int d, int dtype,
                                             b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(m >= 0 && d > 0, B200_ERR_INVALID_ARG, "silu_and_mul_nvfp4_quant: bad sizes");
  B200_CHECK_ARG(d % 16 == 0, B200_ERR_INVALID_ARG, "The N dimension must be multiple of 16.");
  B200_CHECK_ARG(dtype == B200_DT_F16 || dtype == B200_DT_BF16, B200_ERR_INVALID_ARG,
                 "Unsupported input data type for quantize_to_fp4.");
  B200_CHECK_ARG(result && result_block_scale && (input || m == 0), B200_ERR_INVALID_ARG,
                 "silu_and_mul_nvfp4_quant: null pointer");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(input) % 16 == 0 && reinterpret_cast<uintptr_t>(result) % 4 == 0,
                 B200_ERR_ALIGNMENT, "silu_and_mul_nvfp4_quant: input must be 16-byte aligned");
  const int rows_total = ((m + 127) / 128) * 128;
  if (rows_total == 0) return 0;
  const int vecs_total = ((d / 16 + 3) / 4) * 4 * 2;
  const int threads = vecs_total >= 256 ? 256 : ((vecs_total + 31) / 32) * 32;
  int gy = (vecs_total + threads - 1) / threads;
  if (gy > 8) gy = 8;
  int gx = (sm_count() * 8) / gy;
  if (gx > rows_total) gx = rows_total;
  if (gx < 1) gx = 1;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    launch_kernel(nvfp4_quant_kernel<__nv_bfloat16, true, 1>, dim3(gx, gy), dim3(threads), 0, s, 1, (uint32_t*)result,
                  (uint8_t*)result_block_scale, (const __nv_bfloat16*)input, input_global_scale, m, d, d, rows_total,
                  vecs_total);
  else
    launch_kernel(nvfp4_quant_kernel<__half, true, 1>, dim3(gx, gy), dim3(threads), 0, s, 1, (uint32_t*)result,
                  (uint8_t*)result_block_scale, (const __half*)input, input_global_scale, m, d, d, rows_total,
                  vecs_total);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

extern "C" int b200_rms_norm_nvfp4_quant(void* result, void* result_block_scale, const void* input,
                                         const void* weight, void* residual, const float* input_global_scale, int m,
                                         int hidden, int64_t in_row_stride, float eps, int dtype,
                                         b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(m >= 0 && hidden > 0 && hidden <= 32768, B200_ERR_INVALID_ARG, "rms_norm_nvfp4_quant: bad sizes");
  B200_CHECK_ARG(hidden % 16 == 0, B200_ERR_INVALID_ARG, "The N dimension must be multiple of 16.");
  B200_CHECK_ARG(dtype == B200_DT_F16 || dtype == B200_DT_BF16, B200_ERR_INVALID_ARG,
                 "Unsupported input data type for quantize_to_fp4.");
  B200_CHECK_ARG(result && result_block_scale && weight && (input || m == 0), B200_ERR_INVALID_ARG,
                 "rms_norm_nvfp4_quant: null pointer");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(input) % 16 == 0 && reinterpret_cast<uintptr_t>(weight) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(residual) % 16 == 0 && in_row_stride % 8 == 0 &&
                     reinterpret_cast<uintptr_t>(result) % 4 == 0,
                 B200_ERR_ALIGNMENT, "rms_norm_nvfp4_quant: tensors must be 16-byte aligned");
  const int rows_total = ((m + 127) / 128) * 128;
  if (rows_total == 0) return 0;
  const int vecs_total = ((hidden / 16 + 3) / 4) * 4 * 2;
  cudaStream_t s = (cudaStream_t)stream;
  auto go = [&](auto tag) -> int {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(rms_norm_nvfp4_quant_kernel<T>), 32768 * 4))
      return rc_attr;
    launch_kernel(rms_norm_nvfp4_quant_kernel<T>, dim3((unsigned)rows_total), dim3(256), hidden * 4, s, 1,
                  (uint32_t*)result, (uint8_t*)result_block_scale, (const T*)input, (const T*)weight, (T*)residual,
                  input_global_scale, m, hidden, in_row_stride, eps, vecs_total);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  };
  return dtype == B200_DT_BF16 ? go(__nv_bfloat16()) : go(__half());
}
