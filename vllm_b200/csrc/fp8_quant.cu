// FP8 (e4m3) activation quantisation: static, dynamic per-tensor, dynamic per-token.
// Reference behaviour: csrc/libtorch_stable/quantization/w8a8/fp8/common.cu:13-179
// (kernels) and :183-420 (launchers); conversion csrc/quantization/w8a8/fp8/common.cuh:57-76
// (x*inv_scale or x/scale in fp32, clamp to +-448, hardware RN cvt);
// min scale 1/(448*512) (csrc/quantization/utils.cuh).
//
// B200 design: bandwidth-trivial, launch-latency-bound rows (K <= 28672); one
// CTA per token row, 16-byte loads, the row is kept in registers between the
// amax pass and the convert pass when it fits (K <= 8192 for 16-bit inputs) so
// HBM/L2 sees each activation once.
#include "common.cuh"

namespace b200 {

constexpr float kFp8Max = 448.f;
constexpr float kMinScale = 1.0f / (448.f * 512.f);

__device__ __forceinline__ uint8_t cvt_e4m3(float x) {
  const float r = fmaxf(-kFp8Max, fminf(x, kFp8Max));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}

__device__ __forceinline__ float block_max_256(float v, float* sh) {
  v = warp_max(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = (lane < nw) ? sh[lane] : 0.f;
  r = warp_max(r);
  __syncthreads();
  return r;  // every thread has the block max
}

template <typename T> struct Vec16 {  // 16 bytes of T
  static constexpr int N = 16 / sizeof(T);
  T v[N];
};

// ---- dynamic per token ----------------------------------------------------
template <typename T, int kCache>
__global__ void __launch_bounds__(256) per_token_quant_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                                              const T* __restrict__ in,
                                                              const float* __restrict__ scale_ub, int hidden,
                                                              int64_t in_stride, int64_t out_stride) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  __shared__ float sh[8];
  constexpr int N = Vec16<T>::N;
  const int64_t row = blockIdx.x;
  const T* src = in + row * in_stride;
  uint8_t* dst = out + row * out_stride;
  const bool vec_ok = (hidden % N == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dst) & (N - 1)) == 0);
  float amax = 0.f;
  Vec16<T> cache[kCache > 0 ? kCache : 1];
  const int nvec = vec_ok ? hidden / N : 0;
  if (vec_ok) {
#pragma unroll
    for (int c = 0; c < kCache; ++c) {
      const int v = threadIdx.x + c * blockDim.x;
      if (v < nvec) {
        cache[c] = *reinterpret_cast<const Vec16<T>*>(src + (int64_t)v * N);
#pragma unroll
        for (int i = 0; i < N; ++i) amax = fmaxf(amax, fabsf(to_float(cache[c].v[i])));
      }
    }
    for (int v = threadIdx.x + kCache * blockDim.x; v < nvec; v += blockDim.x) {
      Vec16<T> x = *reinterpret_cast<const Vec16<T>*>(src + (int64_t)v * N);
#pragma unroll
      for (int i = 0; i < N; ++i) amax = fmaxf(amax, fabsf(to_float(x.v[i])));
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) amax = fmaxf(amax, fabsf(to_float(src[e])));
  }
  amax = block_max_256(amax, sh);
  float s = scale_ub ? fminf(amax, *scale_ub) : amax;
  s = fmaxf(__fdiv_rn(s, kFp8Max), kMinScale);
  if (threadIdx.x == 0) scales[row] = s;
  if (vec_ok) {
    auto emit = [&](const Vec16<T>& x, int v) {
      uint8_t q[N];
#pragma unroll
      for (int i = 0; i < N; ++i) q[i] = cvt_e4m3(__fdiv_rn(to_float(x.v[i]), s));
      if constexpr (N == 8) *reinterpret_cast<uint2*>(dst + (int64_t)v * N) = *reinterpret_cast<uint2*>(q);
      else *reinterpret_cast<uint32_t*>(dst + (int64_t)v * N) = *reinterpret_cast<uint32_t*>(q);
    };
#pragma unroll
    for (int c = 0; c < kCache; ++c) {
      const int v = threadIdx.x + c * blockDim.x;
      if (v < nvec) emit(cache[c], v);
    }
    for (int v = threadIdx.x + kCache * blockDim.x; v < nvec; v += blockDim.x)
      emit(*reinterpret_cast<const Vec16<T>*>(src + (int64_t)v * N), v);
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) dst[e] = cvt_e4m3(__fdiv_rn(to_float(src[e]), s));
  }
}

// ---- dynamic per tensor: amax pass (atomic max of row_amax/448), then scale pass
template <typename T>
__global__ void __launch_bounds__(256) tensor_amax_kernel(float* __restrict__ scale, const T* __restrict__ in,
                                                          int hidden, int64_t in_stride) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  __shared__ float sh[8];
  constexpr int N = Vec16<T>::N;
  const T* src = in + (int64_t)blockIdx.x * in_stride;
  float amax = 0.f;
  if ((hidden % N == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
      Vec16<T> x = *reinterpret_cast<const Vec16<T>*>(src + (int64_t)v * N);
#pragma unroll
      for (int i = 0; i < N; ++i) amax = fmaxf(amax, fabsf(to_float(x.v[i])));
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) amax = fmaxf(amax, fabsf(to_float(src[e])));
  }
  amax = block_max_256(amax, sh);
  // non-negative floats order like their bit patterns
  if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(scale), __float_as_int(__fdiv_rn(amax, kFp8Max)));
}

// out = e4m3(clamp(x * (1/scale[group]))): covers per-tensor (static or after the
// amax pass), per-token, per-channel and 2-D group scales.
template <typename T>
__global__ void __launch_bounds__(256) scaled_quant_kernel(uint8_t* __restrict__ out, const T* __restrict__ in,
                                                           const float* __restrict__ scale, int hidden,
                                                           int64_t in_stride, int64_t out_stride, int group_m,
                                                           int group_n, int64_t ss_i, int64_t ss_j) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  constexpr int N = Vec16<T>::N;
  const int64_t row = blockIdx.x;
  const T* src = in + row * in_stride;
  uint8_t* dst = out + row * out_stride;
  const float* srow = scale + (ss_i ? (row / group_m) * ss_i : 0);
  const bool vec_ok = (hidden % N == 0) && (group_n % N == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dst) & (N - 1)) == 0);
  if (vec_ok) {
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
      const float inv = __fdiv_rn(1.0f, srow[ss_j ? ((v * N) / group_n) * ss_j : 0]);
      Vec16<T> x = *reinterpret_cast<const Vec16<T>*>(src + (int64_t)v * N);
      uint8_t q[N];
#pragma unroll
      for (int i = 0; i < N; ++i) q[i] = cvt_e4m3(to_float(x.v[i]) * inv);
      if constexpr (N == 8) *reinterpret_cast<uint2*>(dst + (int64_t)v * N) = *reinterpret_cast<uint2*>(q);
      else *reinterpret_cast<uint32_t*>(dst + (int64_t)v * N) = *reinterpret_cast<uint32_t*>(q);
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) {
      const float inv = __fdiv_rn(1.0f, srow[ss_j ? (e / group_n) * ss_j : 0]);
      dst[e] = cvt_e4m3(to_float(src[e]) * inv);
    }
  }
}

template <typename F> static int dispatch_dtype(int dtype, F&& f) {
  switch (dtype) {
    case B200_DT_BF16: return f(__nv_bfloat16());
    case B200_DT_F16: return f(__half());
    case B200_DT_F32: return f(float());
    default: set_error("fp8 quant: unsupported input dtype %d", dtype); return B200_ERR_INVALID_ARG;
  }
}

}  // namespace b200

extern "C" int b200_static_scaled_fp8_quant(void* out, const void* input, const float* scale, int64_t num_tokens,
                                            int hidden, int64_t in_row_stride, int64_t out_row_stride, int group_m,
                                            int group_n, int64_t scale_stride_i, int64_t scale_stride_j, int dtype,
                                            b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID_ARG, "static_scaled_fp8_quant: bad sizes");
  B200_CHECK_ARG(group_m > 0 && group_n > 0, B200_ERR_INVALID_ARG, "static_scaled_fp8_quant: bad group shape");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(out && input && scale, B200_ERR_INVALID_ARG, "static_scaled_fp8_quant: null pointer");
  return dispatch_dtype(dtype, [&](auto tag) {
    using T = decltype(tag);
    launch_kernel(scaled_quant_kernel<T>, dim3((unsigned)num_tokens), dim3(256), 0, (cudaStream_t)stream, 1, 
        (uint8_t*)out, (const T*)input, scale, hidden, in_row_stride, out_row_stride, group_m, group_n,
        scale_stride_i, scale_stride_j);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale, int64_t num_tokens,
                                             int hidden, int64_t in_row_stride, int64_t out_row_stride, int dtype,
                                             b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID_ARG, "dynamic_scaled_fp8_quant: bad sizes");
  B200_CHECK_ARG(scale != nullptr, B200_ERR_INVALID_ARG, "dynamic_scaled_fp8_quant: null scale");
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(scale, 0, sizeof(float), s) != cudaSuccess) {
    set_error("dynamic_scaled_fp8_quant: memset failed");
    return B200_ERR_CUDA;
  }
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(out && input, B200_ERR_INVALID_ARG, "dynamic_scaled_fp8_quant: null pointer");
  return dispatch_dtype(dtype, [&](auto tag) {
    using T = decltype(tag);
    launch_kernel(tensor_amax_kernel<T>, dim3((unsigned)num_tokens), dim3(256), 0, s, 1, scale, (const T*)input, hidden, in_row_stride);
    B200_CHECK_LAUNCH();
    launch_kernel(scaled_quant_kernel<T>, dim3((unsigned)num_tokens), dim3(256), 0, s, 1, (uint8_t*)out, (const T*)input, scale, hidden,
                                                               in_row_stride, out_row_stride, (int)num_tokens,
                                                               hidden, 0, 0);
    B200_CHECK_LAUNCH();
    count_launch(2);
    return 0;
  });
}

extern "C" int b200_dynamic_per_token_scaled_fp8_quant(void* out, const void* input, float* scales,
                                                       const float* scale_ub, int64_t num_tokens, int hidden,
                                                       int64_t in_row_stride, int64_t out_row_stride, int dtype,
                                                       b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID_ARG, "dynamic_per_token_scaled_fp8_quant: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(out && input && scales, B200_ERR_INVALID_ARG, "dynamic_per_token_scaled_fp8_quant: null pointer");
  return dispatch_dtype(dtype, [&](auto tag) {
    using T = decltype(tag);
    constexpr int N = 16 / sizeof(T);
    cudaStream_t s = (cudaStream_t)stream;
    const int threads = hidden / N >= 256 ? 256 : (hidden / N > 32 ? ((hidden / N + 31) / 32) * 32 : 32);
    if (hidden <= 4 * 256 * N)
      launch_kernel(per_token_quant_kernel<T, 4>, dim3((unsigned)num_tokens), dim3(threads), 0, s, 1, 
          (uint8_t*)out, scales, (const T*)input, scale_ub, hidden, in_row_stride, out_row_stride);
    else
      launch_kernel(per_token_quant_kernel<T, 0>, dim3((unsigned)num_tokens), dim3(threads), 0, s, 1, 
          (uint8_t*)out, scales, (const T*)input, scale_ub, hidden, in_row_stride, out_row_stride);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}
