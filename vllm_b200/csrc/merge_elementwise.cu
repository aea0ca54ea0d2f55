// (1) two-way (out, lse) merge — reference csrc/libtorch_stable/attention/merge_attn_states.cu:19-257
// (2) element-wise companions of the quantised linears on the decode step:
//     rms_norm / fused_add_rms_norm (csrc/libtorch_stable/layernorm_kernels.cu),
//     silu_and_mul (csrc/libtorch_stable/activation_kernels.cu),
//     neox rotary embedding (csrc/libtorch_stable/pos_encoding_kernels.cu).
// All are HBM/launch-bound: 16-byte vectors, one pass, fp32 math.
#include "common.cuh"

namespace b200 {

// one thread per 16-byte pack of one (token, head); LSEs are [heads][tokens]
template <typename T>
__global__ void __launch_bounds__(128) merge_states_kernel(
    T* __restrict__ out, float* __restrict__ out_lse, const T* __restrict__ p_out, const float* __restrict__ p_lse,
    const T* __restrict__ s_out, const float* __restrict__ s_lse, int64_t num_tokens, int num_heads, int head_size,
    int64_t o_st, int64_t o_sh, int64_t p_st, int64_t p_sh, int64_t s_st, int64_t s_sh, int64_t lse_sh) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  constexpr int N = 16 / sizeof(T);
  const int packs = head_size / N;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= num_tokens * num_heads * packs) return;
  const int pack = gid % packs;
  const int64_t th = gid / packs;
  const int head = th % num_heads;
  const int64_t tok = th / num_heads;
  float pl = p_lse[head * lse_sh + tok], sl = s_lse[head * lse_sh + tok];
  pl = isinf(pl) ? -INFINITY : pl;
  sl = isinf(sl) ? -INFINITY : sl;
  const float mx = fmaxf(pl, sl);
  const T* pp = p_out + tok * p_st + head * p_sh + pack * N;
  const T* sp = s_out + tok * s_st + head * s_sh + pack * N;
  T* op = out + tok * o_st + head * o_sh + pack * N;
  if (isinf(mx)) {  // both empty: emit the prefix (zeros by contract) and -inf
    *reinterpret_cast<uint4*>(op) = *reinterpret_cast<const uint4*>(pp);
    if (out_lse && pack == 0) out_lse[head * lse_sh + tok] = mx;
    return;
  }
  const float pe = expf(pl - mx), se = expf(sl - mx);
  const float sum = pe + se;
  const float pw = pe / sum, sw = se / sum;
  uint4 pr = *reinterpret_cast<const uint4*>(pp), sr = *reinterpret_cast<const uint4*>(sp), orr;
  const T* pv = reinterpret_cast<const T*>(&pr);
  const T* sv = reinterpret_cast<const T*>(&sr);
  T* ov = reinterpret_cast<T*>(&orr);
#pragma unroll
  for (int i = 0; i < N; ++i) ov[i] = from_float<T>(to_float(pv[i]) * pw + to_float(sv[i]) * sw);
  *reinterpret_cast<uint4*>(op) = orr;
  if (out_lse && pack == 0) out_lse[head * lse_sh + tok] = logf(sum) + mx;
}

__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = lane < nw ? sh[lane] : 0.f;
  r = warp_sum(r);
  __syncthreads();
  return r;
}

// out = x * rsqrt(mean(x^2)+eps) * w ; with residual: x += residual, residual = x first.
template <typename T>
__global__ void __launch_bounds__(256) rms_norm_kernel(T* __restrict__ out, const T* __restrict__ in,
                                                       const T* __restrict__ w, T* __restrict__ residual,
                                                       int hidden, int64_t in_stride, float eps) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  __shared__ float sh[8];
  extern __shared__ __align__(16) uint8_t row_buf[];  // fp32 copy of the row
  float* rowf = reinterpret_cast<float*>(row_buf);
  const int64_t row = blockIdx.x;
  const T* src = in + row * in_stride;
  T* res = residual ? residual + row * (int64_t)hidden : nullptr;
  float ss = 0.f;
  for (int e = threadIdx.x; e < hidden; e += blockDim.x) {
    float x = to_float(src[e]);
    if (res) {
      x += to_float(res[e]);
      const T xr = from_float<T>(x);  // reference rounds the sum to T before the norm
      res[e] = xr;
      x = to_float(xr);
    }
    rowf[e] = x;
    ss += x * x;
  }
  ss = block_sum(ss, sh);
  const float inv = rsqrtf(ss / hidden + eps);
  T* dst = out + row * (int64_t)hidden;
  for (int e = threadIdx.x; e < hidden; e += blockDim.x)
    dst[e] = from_float<T>(to_float(from_float<T>(rowf[e] * inv)) * to_float(w[e]));
}

template <typename T>
__global__ void __launch_bounds__(256) silu_and_mul_kernel(T* __restrict__ out, const T* __restrict__ in, int d) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  constexpr int N = 16 / sizeof(T);
  const int64_t row = blockIdx.x;
  const T* g = in + row * 2 * (int64_t)d;
  const T* u = g + d;
  T* o = out + row * (int64_t)d;
  if (d % N == 0) {
    for (int v = threadIdx.x; v < d / N; v += blockDim.x) {
      uint4 gr = *reinterpret_cast<const uint4*>(g + v * N), ur = *reinterpret_cast<const uint4*>(u + v * N), orr;
      const T* gv = reinterpret_cast<const T*>(&gr);
      const T* uv = reinterpret_cast<const T*>(&ur);
      T* ov = reinterpret_cast<T*>(&orr);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const float x = to_float(gv[i]);
        const T s = from_float<T>(x / (1.0f + expf(-x)));  // silu rounded to T, then * up (reference order)
        ov[i] = from_float<T>(to_float(s) * to_float(uv[i]));
      }
      *reinterpret_cast<uint4*>(o + v * N) = orr;
    }
  } else {
    for (int e = threadIdx.x; e < d; e += blockDim.x) {
      const float x = to_float(g[e]);
      const T s = from_float<T>(x / (1.0f + expf(-x)));
      o[e] = from_float<T>(to_float(s) * to_float(u[e]));
    }
  }
}

// neox style: x1 = x[:rot/2], x2 = x[rot/2:rot]; cache row = [cos(rot/2) | sin(rot/2)]
template <typename T>
__global__ void __launch_bounds__(256) rope_neox_kernel(T* __restrict__ q, T* __restrict__ k,
                                                        const int64_t* __restrict__ positions,
                                                        const T* __restrict__ cos_sin, int num_q_heads,
                                                        int num_kv_heads, int head_size, int rot_dim,
                                                        int64_t q_st, int64_t k_st) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  const int64_t tok = blockIdx.x;
  const int half = rot_dim / 2;
  const T* cs = cos_sin + positions[tok] * rot_dim;
  const int total = (num_q_heads + num_kv_heads) * half;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int head = i / half, j = i - head * half;
    T* base = head < num_q_heads ? q + tok * q_st + (int64_t)head * head_size
                                 : k + tok * k_st + (int64_t)(head - num_q_heads) * head_size;
    const float c = to_float(cs[j]), s = to_float(cs[half + j]);
    const float x = to_float(base[j]), y = to_float(base[half + j]);
    base[j] = from_float<T>(x * c - y * s);
    base[half + j] = from_float<T>(y * c + x * s);
  }
}

template <typename F> static int dispatch16(int dtype, const char* what, F&& f) {
  switch (dtype) {
    case B200_DT_BF16: return f(__nv_bfloat16());
    case B200_DT_F16: return f(__half());
    default: set_error("%s: dtype must be fp16 or bf16", what); return B200_ERR_INVALID_ARG;
  }
}

}  // namespace b200

extern "C" int b200_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                                      const float* prefix_lse, const void* suffix_output, const float* suffix_lse,
                                      int64_t num_tokens, int num_heads, int head_size, int64_t out_stride_tok,
                                      int64_t out_stride_head, int64_t p_stride_tok, int64_t p_stride_head,
                                      int64_t s_stride_tok, int64_t s_stride_head, int64_t lse_stride_head,
                                      int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && num_heads > 0 && head_size > 0, B200_ERR_INVALID_ARG,
                 "merge_attn_states: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(output && prefix_output && prefix_lse && suffix_output && suffix_lse, B200_ERR_INVALID_ARG,
                 "merge_attn_states: null pointer");
  return dispatch16(dtype, "merge_attn_states", [&](auto tag) {
    using T = decltype(tag);
    constexpr int N = 16 / sizeof(T);
    B200_CHECK_ARG(head_size % N == 0, B200_ERR_INVALID_ARG,
                   "merge_attn_states: head_size must be a multiple of pack_size %d", N);
    auto al = [&](const void* p, int64_t a, int64_t b) {
      return reinterpret_cast<uintptr_t>(p) % 16 == 0 && a % N == 0 && b % N == 0;
    };
    B200_CHECK_ARG(al(output, out_stride_tok, out_stride_head) && al(prefix_output, p_stride_tok, p_stride_head) &&
                       al(suffix_output, s_stride_tok, s_stride_head),
                   B200_ERR_ALIGNMENT, "merge_attn_states: tensors must be 16-byte aligned");
    const int64_t total = num_tokens * num_heads * (head_size / N);
    launch_kernel(merge_states_kernel<T>, dim3((unsigned)((total + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, 1, 
        (T*)output, output_lse, (const T*)prefix_output, prefix_lse, (const T*)suffix_output, suffix_lse, num_tokens,
        num_heads, head_size, out_stride_tok, out_stride_head, p_stride_tok, p_stride_head, s_stride_tok,
        s_stride_head, lse_stride_head);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_rms_norm(void* out, const void* input, const void* weight, void* residual, int64_t num_tokens,
                             int hidden, int64_t in_row_stride, float eps, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden <= 32768, B200_ERR_INVALID_ARG, "rms_norm: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(out && input && weight, B200_ERR_INVALID_ARG, "rms_norm: null pointer");
  return dispatch16(dtype, "rms_norm", [&](auto tag) {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(rms_norm_kernel<T>), 32768 * 4)) return rc_attr;
    launch_kernel(rms_norm_kernel<T>, dim3((unsigned)num_tokens), dim3(256), hidden * 4, (cudaStream_t)stream, 1, 
        (T*)out, (const T*)input, (const T*)weight, (T*)residual, hidden, in_row_stride, eps);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d, int dtype,
                                 b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && d > 0, B200_ERR_INVALID_ARG, "silu_and_mul: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(out && input, B200_ERR_INVALID_ARG, "silu_and_mul: null pointer");
  return dispatch16(dtype, "silu_and_mul", [&](auto tag) {
    using T = decltype(tag);
    launch_kernel(silu_and_mul_kernel<T>, dim3((unsigned)num_tokens), dim3(256), 0, (cudaStream_t)stream, 1, (T*)out, (const T*)input, d);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_rotary_embedding_neox(void* query, void* key, const int64_t* positions,
                                          const void* cos_sin_cache, int64_t num_tokens, int num_q_heads,
                                          int num_kv_heads, int head_size, int rot_dim, int64_t q_stride_tok,
                                          int64_t k_stride_tok, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size, B200_ERR_INVALID_ARG,
                 "rotary_embedding: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(query && key && positions && cos_sin_cache, B200_ERR_INVALID_ARG, "rotary_embedding: null pointer");
  return dispatch16(dtype, "rotary_embedding", [&](auto tag) {
    using T = decltype(tag);
    launch_kernel(rope_neox_kernel<T>, dim3((unsigned)num_tokens), dim3(256), 0, (cudaStream_t)stream, 1, 
        (T*)query, (T*)key, positions, (const T*)cos_sin_cache, num_q_heads, num_kv_heads, head_size, rot_dim,
        q_stride_tok, k_stride_tok);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}
