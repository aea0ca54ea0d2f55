// Fused element-wise neighbours of the quantised linears and of the cache write.
// Each removes one or two launch-bound kernels and an activation round trip per layer:
//   (1) rms_norm_dynamic_per_token_quant   (fused add +) RMSNorm + FP8 per-token quant
//         reference: csrc/libtorch_stable/quantization/fused_kernels/fused_layernorm_dynamic_per_token_quant.cu:50-91,
//         math layernorm_utils.cuh (compute_rms / compute_dynamic_per_token_scales / norm_and_quant),
//         schema csrc/libtorch_stable/torch_bindings.cpp:394-396
//   (2) silu_and_mul_dynamic_per_token_quant  SiLU(gate)*up + FP8 per-token quant
//         = silu_and_mul (csrc/libtorch_stable/activation_kernels.cu) followed by
//           dynamic_per_token_scaled_fp8_quant (fp8/common.cu:137-179), bit-identical to that composition
//   (3) rotary_embedding_and_cache_write    neox RoPE on q,k (in place) + K/V cache write over slot_mapping
//         = the fused path behind AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:932-980),
//           bit-identical to rotary_embedding followed by reshape_and_cache_flash
// One CTA per token row; the row lives in shared memory between the statistics pass and the convert pass,
// so HBM/L2 sees every activation once; 16-byte vectors throughout.
#include "peer_sync.cuh"

namespace b200 {

constexpr float kFp8MaxF = 448.f;
constexpr float kMinScaleF = 1.0f / (448.f * 512.f);

__device__ __forceinline__ uint8_t cvt_e4m3_sat(float x) {
  const float r = fmaxf(-kFp8MaxF, fminf(x, kFp8MaxF));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}

template <bool kMax> __device__ __forceinline__ float block_reduce(float v, float* sh) {
  v = kMax ? warp_max(v) : warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // sh may still be read from a previous reduction
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = lane < nw ? sh[lane] : 0.f;
  return kMax ? warp_max(r) : warp_sum(r);
}

// T x T product rounded to T, as `static_cast<scalar_t>(a) * b` does in the reference
template <typename T> __device__ __forceinline__ float mul_round(float a, float b) {
  return to_float(from_float<T>(to_float(from_float<T>(a)) * b));
}

// ---- (1) ---------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) rms_norm_quant_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                                              const T* __restrict__ in, const T* __restrict__ w,
                                                              T* __restrict__ residual,
                                                              const float* __restrict__ scale_ub, int hidden,
                                                              int64_t in_stride, float eps) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  constexpr int N = 16 / sizeof(T);
  __shared__ float sh[32];
  extern __shared__ __align__(16) uint8_t row_buf[];
  float* rowf = reinterpret_cast<float*>(row_buf);  // x (then y) in fp32
  const int64_t row = blockIdx.x;
  const T* src = in + row * in_stride;
  T* res = residual ? residual + row * (int64_t)hidden : nullptr;
  uint8_t* dst = out + row * (int64_t)hidden;
  const bool vec = (hidden % N == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                   (!res || (reinterpret_cast<uintptr_t>(res) & 15) == 0) && ((reinterpret_cast<uintptr_t>(w) & 15) == 0);
  float ss = 0.f;
  if (vec) {
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
      uint4 a = *reinterpret_cast<const uint4*>(src + v * N);
      const T* av = reinterpret_cast<const T*>(&a);
      float x[N];
#pragma unroll
      for (int i = 0; i < N; ++i) x[i] = to_float(av[i]);
      if (res) {
        uint4 r = *reinterpret_cast<const uint4*>(res + v * N);
        const T* rv = reinterpret_cast<const T*>(&r);
        uint4 o;
        T* ov = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int i = 0; i < N; ++i) {
          x[i] += to_float(rv[i]);  // the norm sees the unrounded fp32 sum (layernorm_utils.cuh compute_rms)
          ov[i] = from_float<T>(x[i]);
        }
        *reinterpret_cast<uint4*>(res + v * N) = o;
      }
#pragma unroll
      for (int i = 0; i < N; ++i) {
        rowf[v * N + i] = x[i];
        ss += x[i] * x[i];
      }
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) {
      float x = to_float(src[e]);
      if (res) {
        x += to_float(res[e]);
        res[e] = from_float<T>(x);
      }
      rowf[e] = x;
      ss += x * x;
    }
  }
  ss = block_reduce<false>(ss, sh);
  const float rms = rsqrtf(ss / hidden + eps);
  float amax = 0.f;
  if (vec) {  // 8 elements per thread and access (the scalar loop was 8 dependent rounds of 2-byte weight loads)
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
      const uint4 wr = *reinterpret_cast<const uint4*>(w + v * N);
      const T* wv = reinterpret_cast<const T*>(&wr);
      float* rp = rowf + v * N;
#pragma unroll
      for (int i = 0; i < N; i += 4) {
        float4 x = *reinterpret_cast<const float4*>(rp + i);
        x.x = mul_round<T>(x.x * rms, to_float(wv[i]));
        x.y = mul_round<T>(x.y * rms, to_float(wv[i + 1]));
        x.z = mul_round<T>(x.z * rms, to_float(wv[i + 2]));
        x.w = mul_round<T>(x.w * rms, to_float(wv[i + 3]));
        *reinterpret_cast<float4*>(rp + i) = x;
        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(x.x), fabsf(x.y)), fmaxf(fabsf(x.z), fabsf(x.w))));
      }
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) {
      const float y = mul_round<T>(rowf[e] * rms, to_float(w[e]));
      rowf[e] = y;
      amax = fmaxf(amax, fabsf(y));
    }
  }
  amax = block_reduce<true>(amax, sh);
  float s = scale_ub ? fminf(amax, *scale_ub) : amax;
  s = fmaxf(s / kFp8MaxF, kMinScaleF);
  if (threadIdx.x == 0) scales[row] = s;
  if ((hidden & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
    for (int v = threadIdx.x; v < hidden / 4; v += blockDim.x) {
      const float4 y = *reinterpret_cast<const float4*>(rowf + v * 4);
      uint32_t q = cvt_e4m3_sat(__fdiv_rn(y.x, s)) | (cvt_e4m3_sat(__fdiv_rn(y.y, s)) << 8) |
                   (cvt_e4m3_sat(__fdiv_rn(y.z, s)) << 16) | (cvt_e4m3_sat(__fdiv_rn(y.w, s)) << 24);
      *reinterpret_cast<uint32_t*>(dst + v * 4) = q;
    }
  } else {
    for (int e = threadIdx.x; e < hidden; e += blockDim.x) dst[e] = cvt_e4m3_sat(__fdiv_rn(rowf[e], s));
  }
}

// ---- (1b) the same fused with the tensor-parallel all-reduce that precedes it ------------------------------
// hidden = sum over ranks of the row-parallel GEMM output (one-shot: every rank reads all peer buffers, rank order,
// fp32 accumulation, rounded to T like the stand-alone all-reduce), then exactly kernel (1) with a residual.
// One launch and one activation round trip less per row-parallel linear.  Start barrier only: the caller
// alternates between two buffers (see b200_peer_all_reduce_out).
template <typename T>
__global__ void __launch_bounds__(1024) peer_allreduce_rms_norm_quant_kernel(
    uint8_t* out, float* scales, const ArArgs a, const T* __restrict__ w, T* residual, int num_tokens, int hidden,
    float eps) {
  pdl_trigger();
  pdl_wait();
  constexpr int N = 8;
  __shared__ float sh[32];
  __shared__ uint32_t* s_pads[kArMaxWorld];
  extern __shared__ __align__(16) uint8_t row_buf[];
  float* rowf = reinterpret_cast<float*>(row_buf);
  if ((int)threadIdx.x < a.world) s_pads[threadIdx.x] = a.pads[threadIdx.x];
  const int rank = ar_rank(a);
  if (a.emu) {  // single-GPU emulation: every emulated rank has its own outputs and residual
    out += (int64_t)rank * a.emu_stride;
    residual += (int64_t)rank * a.emu_stride;
    scales += (int64_t)rank * num_tokens;
  }
  uint32_t* my_epoch = a.pads[rank] + blockIdx.x * kArSlotWords + 2 * kArMaxWorld;
  const uint32_t epoch = *my_epoch + 1;
  ar_barrier(s_pads, rank, a.world, 0, epoch, a.spin_limit);
  const int nv = hidden / N;
  if (a.mc) {
    // NVLS, two-shot per CTA: of THIS CTA's rows, rank r asks the switch for the sum of column slice r
    // (multimem.ld_reduce, fp32 accumulation, rounded to T once) and multicasts it into every rank's buffer; after
    // the second barrier (same CTA index on every rank = same rows) the whole rows are in the local buffer.
    const int cpr = nv / a.world;  // 16-byte vectors of one row per rank (host: hidden % (8 * world) == 0)
    const int my_rows = (num_tokens - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    uint4* mc = reinterpret_cast<uint4*>(const_cast<void*>(a.mc));
    for (int i = threadIdx.x; i < my_rows * cpr; i += blockDim.x) {
      const int k = i / cpr, c = i - k * cpr;
      const int64_t idx = (int64_t)(blockIdx.x + k * gridDim.x) * nv + rank * cpr + c;
      const uint4 sum = multimem_ld_reduce_add<T>(mc + idx);
      multimem_st<T>(mc + idx, sum);
    }
    ar_barrier(s_pads, rank, a.world, 1, epoch, a.spin_limit);
  }
  for (int row = blockIdx.x; row < num_tokens; row += gridDim.x) {
    T* res = residual + (int64_t)row * hidden;
    uint8_t* dst = out + (int64_t)row * hidden;
    float ss = 0.f;
    for (int v = threadIdx.x; v < nv; v += blockDim.x) {
      float acc[N];
      if (a.mc) {  // the summed row, multicast into the local buffer above (L2: other SMs' L1s never saw it)
        const uint4 sm = __ldcg(reinterpret_cast<const uint4*>(a.bufs[rank]) + (int64_t)row * nv + v);
        const T* e = reinterpret_cast<const T*>(&sm);
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] = to_float(e[i]);
      } else {
        uint4 in[kArMaxWorld];
#pragma unroll
        for (int q = 0; q < kArMaxWorld; ++q)
          if (q < a.world) in[q] = reinterpret_cast<const uint4*>(a.bufs[q])[(int64_t)row * nv + v];
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] = 0.f;
#pragma unroll
        for (int q = 0; q < kArMaxWorld; ++q)
          if (q < a.world) {
            const T* e = reinterpret_cast<const T*>(&in[q]);
#pragma unroll
            for (int i = 0; i < N; ++i) acc[i] += to_float(e[i]);
          }
      }
      const uint4 rr = *reinterpret_cast<const uint4*>(res + v * N);
      const T* rv = reinterpret_cast<const T*>(&rr);
      uint4 o;
      T* ov = reinterpret_cast<T*>(&o);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const float x = to_float(from_float<T>(acc[i])) + to_float(rv[i]);  // all-reduce output is T; then fused add
        ov[i] = from_float<T>(x);
        rowf[v * N + i] = x;
        ss += x * x;
      }
      *reinterpret_cast<uint4*>(res + v * N) = o;
    }
    ss = block_reduce<false>(ss, sh);
    const float rms = rsqrtf(ss / hidden + eps);
    float amax = 0.f;
    for (int v = threadIdx.x; v < nv; v += blockDim.x) {
      const uint4 wr = *reinterpret_cast<const uint4*>(w + v * N);
      const T* wv = reinterpret_cast<const T*>(&wr);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const float y = mul_round<T>(rowf[v * N + i] * rms, to_float(wv[i]));
        rowf[v * N + i] = y;
        amax = fmaxf(amax, fabsf(y));
      }
    }
    amax = block_reduce<true>(amax, sh);
    const float s = fmaxf(amax / kFp8MaxF, kMinScaleF);
    if (threadIdx.x == 0) scales[row] = s;
    for (int v = threadIdx.x; v < hidden / 4; v += blockDim.x) {
      const float4 y = *reinterpret_cast<const float4*>(rowf + v * 4);
      const uint32_t qv = cvt_e4m3_sat(__fdiv_rn(y.x, s)) | (cvt_e4m3_sat(__fdiv_rn(y.y, s)) << 8) |
                          (cvt_e4m3_sat(__fdiv_rn(y.z, s)) << 16) | (cvt_e4m3_sat(__fdiv_rn(y.w, s)) << 24);
      *reinterpret_cast<uint32_t*>(dst + v * 4) = qv;
    }
    __syncthreads();  // rowf is reused by the next row
  }
  if (threadIdx.x == 0) *my_epoch = epoch;
}

// ---- (2) ---------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) silu_mul_quant_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                                              const T* __restrict__ in, int d) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  constexpr int N = 16 / sizeof(T);
  __shared__ float sh[32];
  extern __shared__ __align__(16) uint8_t row_buf[];
  T* rowt = reinterpret_cast<T*>(row_buf);  // act row in T (what silu_and_mul would have stored)
  const int64_t row = blockIdx.x;
  const T* g = in + row * 2 * (int64_t)d;
  const T* u = g + d;
  uint8_t* dst = out + row * (int64_t)d;
  float amax = 0.f;
  auto act = [](T gv, T uv) {
    const float x = to_float(gv);
    const T s = from_float<T>(x / (1.0f + expf(-x)));
    return from_float<T>(to_float(s) * to_float(uv));
  };
  if (d % N == 0) {
    for (int v = threadIdx.x; v < d / N; v += blockDim.x) {
      uint4 gr = *reinterpret_cast<const uint4*>(g + v * N), ur = *reinterpret_cast<const uint4*>(u + v * N), orr;
      const T* gv = reinterpret_cast<const T*>(&gr);
      const T* uv = reinterpret_cast<const T*>(&ur);
      T* ov = reinterpret_cast<T*>(&orr);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        ov[i] = act(gv[i], uv[i]);
        amax = fmaxf(amax, fabsf(to_float(ov[i])));
      }
      *reinterpret_cast<uint4*>(rowt + v * N) = orr;
    }
  } else {
    for (int e = threadIdx.x; e < d; e += blockDim.x) {
      const T y = act(g[e], u[e]);
      rowt[e] = y;
      amax = fmaxf(amax, fabsf(to_float(y)));
    }
  }
  amax = block_reduce<true>(amax, sh);
  const float s = fmaxf(__fdiv_rn(amax, kFp8MaxF), kMinScaleF);
  if (threadIdx.x == 0) scales[row] = s;
  if (d % N == 0 && (reinterpret_cast<uintptr_t>(dst) & (N - 1)) == 0) {
    for (int v = threadIdx.x; v < d / N; v += blockDim.x) {
      uint4 yr = *reinterpret_cast<const uint4*>(rowt + v * N);
      const T* yv = reinterpret_cast<const T*>(&yr);
      uint8_t q[N];
#pragma unroll
      for (int i = 0; i < N; ++i) q[i] = cvt_e4m3_sat(__fdiv_rn(to_float(yv[i]), s));
      *reinterpret_cast<uint2*>(dst + v * N) = *reinterpret_cast<uint2*>(q);
    }
  } else {
    for (int e = threadIdx.x; e < d; e += blockDim.x) dst[e] = cvt_e4m3_sat(__fdiv_rn(to_float(rowt[e]), s));
  }
}

// ---- (3) ---------------------------------------------------------------------------------
template <int kKv> __device__ __forceinline__ uint8_t kv_to_fp8(float x, float scale) {
  return (uint8_t)__nv_cvt_float_to_fp8(__fdiv_rn(x, scale), __NV_SATFINITE,
                                        kKv == B200_KV_FP8_E5M2 ? __NV_E5M2 : __NV_E4M3);
}

template <typename T, int kKv>
__global__ void __launch_bounds__(256) rope_cache_kernel(
    T* __restrict__ q, T* __restrict__ k, const T* __restrict__ v, const int64_t* __restrict__ positions,
    const T* __restrict__ cos_sin, void* __restrict__ key_cache, void* __restrict__ value_cache,
    const int64_t* __restrict__ slot_mapping, int64_t num_slots, int num_q_heads, int num_kv_heads, int head_size,
    int rot_dim, int64_t q_st, int64_t k_st, int64_t v_st, int block_size, int64_t block_stride,
    int64_t page_stride, int64_t head_stride, const float* __restrict__ k_scale, const float* __restrict__ v_scale,
    int kv_scale_stride) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  using cache_t = typename std::conditional<kKv == B200_KV_AUTO, T, uint8_t>::type;
  extern __shared__ __align__(16) uint8_t kv_buf[];
  T* ks = reinterpret_cast<T*>(kv_buf);  // rotated key row [Hkv][D]
  const int64_t tok = blockIdx.x;
  const int half = rot_dim / 2;
  const T* cs = cos_sin + positions[tok] * rot_dim;
  const int kd = num_kv_heads * head_size;
  const int64_t slot = tok < num_slots ? slot_mapping[tok] : -1;  // -1: PAD_SLOT_ID / token without a slot
  const int64_t dst_base = slot >= 0 ? (slot / block_size) * block_stride + (slot % block_size) * page_stride : 0;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  // 16-byte path: 8 elements per thread and access.  (The scalar path below did 10 dependent rounds of 2-byte
  // accesses per thread at Llama shapes and took 15 us for 128 tokens; this one needs 2.)
  const bool vec = (half % 8 == 0) && (head_size % 8 == 0) && (q_st % 8 == 0) && (k_st % 8 == 0) && (v_st % 8 == 0) &&
                   (rot_dim % 8 == 0) && al16(q) && al16(k) && al16(v) && al16(cos_sin) &&
                   (page_stride % 8 == 0) && (block_stride % 8 == 0) && (head_stride % 8 == 0) &&
                   (reinterpret_cast<uintptr_t>(key_cache) % 16 == 0) && (reinterpret_cast<uintptr_t>(value_cache) % 16 == 0);
  if (vec) {
    // stage the key row (rotated in place below; the cache write reads it back)
    for (int i = threadIdx.x; i < kd / 8; i += blockDim.x)
      reinterpret_cast<uint4*>(ks)[i] = *reinterpret_cast<const uint4*>(k + tok * k_st + i * 8);
    __syncthreads();
    const int hv = half / 8;
    for (int i = threadIdx.x; i < (num_q_heads + num_kv_heads) * hv; i += blockDim.x) {
      const int head = i / hv, j = (i - head * hv) * 8;
      const uint4 c8 = *reinterpret_cast<const uint4*>(cs + j), s8 = *reinterpret_cast<const uint4*>(cs + half + j);
      const bool is_q = head < num_q_heads;
      T* base = is_q ? q + tok * q_st + (int64_t)head * head_size : ks + (head - num_q_heads) * head_size;
      const uint4 x8 = *reinterpret_cast<const uint4*>(base + j), y8 = *reinterpret_cast<const uint4*>(base + half + j);
      const T *cv = reinterpret_cast<const T*>(&c8), *sv = reinterpret_cast<const T*>(&s8);
      const T *xv = reinterpret_cast<const T*>(&x8), *yv = reinterpret_cast<const T*>(&y8);
      uint4 r0, r1;
      T *r0v = reinterpret_cast<T*>(&r0), *r1v = reinterpret_cast<T*>(&r1);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float c = to_float(cv[e]), sn = to_float(sv[e]), x = to_float(xv[e]), y = to_float(yv[e]);
        r0v[e] = from_float<T>(x * c - y * sn);
        r1v[e] = from_float<T>(y * c + x * sn);
      }
      *reinterpret_cast<uint4*>(base + j) = r0;
      *reinterpret_cast<uint4*>(base + half + j) = r1;
      if (!is_q) {
        T* gb = k + tok * k_st + (int64_t)(head - num_q_heads) * head_size;
        *reinterpret_cast<uint4*>(gb + j) = r0;
        *reinterpret_cast<uint4*>(gb + half + j) = r1;
      }
    }
    __syncthreads();
    if (slot < 0) return;
    for (int i = threadIdx.x; i < 2 * kd / 8; i += blockDim.x) {
      const int which = i / (kd / 8), rem = (i - which * (kd / 8)) * 8;
      const int head = rem / head_size, d = rem - head * head_size;
      const uint4 x8 = which ? *reinterpret_cast<const uint4*>(v + tok * v_st + rem) : *reinterpret_cast<const uint4*>(ks + rem);
      cache_t* dst = reinterpret_cast<cache_t*>(which ? value_cache : key_cache) + dst_base + (int64_t)head * head_stride + d;
      if constexpr (kKv == B200_KV_AUTO) {
        *reinterpret_cast<uint4*>(dst) = x8;
      } else {
        const float sc = (which ? v_scale : k_scale)[head * kv_scale_stride];
        const T* xv = reinterpret_cast<const T*>(&x8);
        uint32_t w[2] = {0u, 0u};
#pragma unroll
        for (int e = 0; e < 8; ++e) w[e >> 2] |= (uint32_t)kv_to_fp8<kKv>(to_float(xv[e]), sc) << (8 * (e & 3));
        *reinterpret_cast<uint2*>(dst) = make_uint2(w[0], w[1]);
      }
    }
    return;
  }
  // ---- scalar path (odd sizes / unaligned views) ----
  for (int i = threadIdx.x; i < kd; i += blockDim.x) ks[i] = k[tok * k_st + i];
  __syncthreads();
  const int total = (num_q_heads + num_kv_heads) * half;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int head = i / half, j = i - head * half;
    const float c = to_float(cs[j]), s = to_float(cs[half + j]);
    if (head < num_q_heads) {
      T* base = q + tok * q_st + (int64_t)head * head_size;
      const float x = to_float(base[j]), y = to_float(base[half + j]);
      base[j] = from_float<T>(x * c - y * s);
      base[half + j] = from_float<T>(y * c + x * s);
    } else {
      const int h = head - num_q_heads;
      T* sb = ks + h * head_size;
      const float x = to_float(sb[j]), y = to_float(sb[half + j]);
      const T r0 = from_float<T>(x * c - y * s), r1 = from_float<T>(y * c + x * s);
      sb[j] = r0;
      sb[half + j] = r1;
      T* gb = k + tok * k_st + (int64_t)h * head_size;
      gb[j] = r0;
      gb[half + j] = r1;
    }
  }
  __syncthreads();
  if (slot < 0) return;
  for (int i = threadIdx.x; i < 2 * kd; i += blockDim.x) {
    const int which = i / kd, rem = i - which * kd;
    const int head = rem / head_size, d = rem - head * head_size;
    const T x = which ? v[tok * v_st + rem] : ks[rem];
    cache_t* dst = reinterpret_cast<cache_t*>(which ? value_cache : key_cache) + dst_base + (int64_t)head * head_stride;
    if constexpr (kKv == B200_KV_AUTO) {
      dst[d] = x;
    } else {
      dst[d] = kv_to_fp8<kKv>(to_float(x), (which ? v_scale : k_scale)[head * kv_scale_stride]);
    }
  }
}

template <typename F> static int dispatch16f(int dtype, const char* what, F&& f) {
  switch (dtype) {
    case B200_DT_BF16: return f(__nv_bfloat16());
    case B200_DT_F16: return f(__half());
    default: set_error("%s: dtype must be fp16 or bf16", what); return B200_ERR_INVALID_ARG;
  }
}
static int row_threads(int vecs) {
  int t = (vecs + 31) / 32 * 32;
  return t < 64 ? 64 : (t > 1024 ? 1024 : t);
}

}  // namespace b200

extern "C" int b200_rms_norm_dynamic_per_token_quant(void* result, const void* input, const void* weight,
                                                     float* scales, float epsilon, const float* scale_ub,
                                                     void* residual, int64_t num_tokens, int hidden,
                                                     int64_t in_row_stride, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && hidden > 0 && hidden <= 49152, B200_ERR_INVALID_ARG,
                 "rms_norm_dynamic_per_token_quant: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(result && input && weight && scales, B200_ERR_INVALID_ARG,
                 "rms_norm_dynamic_per_token_quant: null pointer");
  return dispatch16f(dtype, "rms_norm_dynamic_per_token_quant", [&](auto tag) {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(rms_norm_quant_kernel<T>), 49152 * 4)) return rc_attr;
    const int threads = row_threads(hidden / (int)(16 / sizeof(T)));
    launch_kernel(rms_norm_quant_kernel<T>, dim3((unsigned)num_tokens), dim3(threads), hidden * 4, (cudaStream_t)stream, 1, 
        (uint8_t*)result, scales, (const T*)input, (const T*)weight, (T*)residual, scale_ub, hidden, in_row_stride,
        epsilon);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_silu_and_mul_dynamic_per_token_quant(void* result, float* scales, const void* input,
                                                         int64_t num_tokens, int d, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && d > 0 && d <= 98304, B200_ERR_INVALID_ARG,
                 "silu_and_mul_dynamic_per_token_quant: bad sizes");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(result && scales && input, B200_ERR_INVALID_ARG, "silu_and_mul_dynamic_per_token_quant: null pointer");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(input) % 16 == 0, B200_ERR_ALIGNMENT,
                 "silu_and_mul_dynamic_per_token_quant: input must be 16-byte aligned");
  return dispatch16f(dtype, "silu_and_mul_dynamic_per_token_quant", [&](auto tag) {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(silu_mul_quant_kernel<T>), 98304 * 2)) return rc_attr;
    const int threads = row_threads(d / (int)(16 / sizeof(T)));
    launch_kernel(silu_mul_quant_kernel<T>, dim3((unsigned)num_tokens), dim3(threads), d * sizeof(T), (cudaStream_t)stream, 1, 
        (uint8_t*)result, scales, (const T*)input, d);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_rotary_embedding_and_cache_write(
    void* query, void* key, const void* value, const int64_t* positions, const void* cos_sin_cache,
    void* key_cache, void* value_cache, const int64_t* slot_mapping, int64_t num_tokens, int64_t num_slots,
    int num_q_heads, int num_kv_heads, int head_size, int rot_dim, int64_t q_stride_tok, int64_t k_stride_tok,
    int64_t v_stride_tok, int block_size, int64_t block_stride, int64_t page_stride, int64_t head_stride, int dtype,
    int kv_dtype, const float* k_scale, const float* v_scale, int kv_scale_stride, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size && block_size > 0 &&
                     num_q_heads > 0 && num_kv_heads > 0,
                 B200_ERR_INVALID_ARG, "rotary_embedding_and_cache_write: bad sizes");
  B200_CHECK_ARG(kv_dtype == B200_KV_AUTO || kv_dtype == B200_KV_FP8_E4M3 || kv_dtype == B200_KV_FP8_E5M2,
                 B200_ERR_INVALID_ARG, "rotary_embedding_and_cache_write: Unsupported data type of kv cache: %d", kv_dtype);
  B200_CHECK_ARG(kv_dtype == B200_KV_AUTO || (k_scale && v_scale), B200_ERR_INVALID_ARG,
                 "rotary_embedding_and_cache_write: fp8 cache needs k_scale and v_scale");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(query && key && value && positions && cos_sin_cache && key_cache && value_cache && slot_mapping,
                 B200_ERR_INVALID_ARG, "rotary_embedding_and_cache_write: null pointer");
  const size_t smem = (size_t)num_kv_heads * head_size * 2;
  B200_CHECK_ARG(smem <= 48 * 1024, B200_ERR_UNSUPPORTED, "rotary_embedding_and_cache_write: key row too large");
  return dispatch16f(dtype, "rotary_embedding_and_cache_write", [&](auto tag) {
    using T = decltype(tag);
    cudaStream_t s = (cudaStream_t)stream;
#define B200_RC(KV)                                                                                             \
  launch_kernel(rope_cache_kernel<T, KV>, dim3((unsigned)num_tokens), dim3(256), smem, s, 1,                                              \
      (T*)query, (T*)key, (const T*)value, positions, (const T*)cos_sin_cache, key_cache, value_cache,          \
      slot_mapping, num_slots, num_q_heads, num_kv_heads, head_size, rot_dim, q_stride_tok, k_stride_tok,        \
      v_stride_tok, block_size, block_stride, page_stride, head_stride, k_scale, v_scale, kv_scale_stride)
    if (kv_dtype == B200_KV_AUTO) B200_RC(B200_KV_AUTO);
    else if (kv_dtype == B200_KV_FP8_E4M3) B200_RC(B200_KV_FP8_E4M3);
    else B200_RC(B200_KV_FP8_E5M2);
#undef B200_RC
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

static int ar_rms_norm_quant_impl(void* result, float* scales, const uint64_t* peer_buffers,
                                  const uint64_t* peer_signal_pads, int rank, int world_size, int64_t byte_offset,
                                  int64_t signal_pad_bytes, const void* weight, void* residual, float epsilon,
                                  int64_t num_tokens, int hidden, int dtype, b200_stream_t stream, uint64_t multicast_ptr) {
  using namespace b200;
  B200_CHECK_ARG(result && scales && peer_buffers && peer_signal_pads && weight && residual, B200_ERR_INVALID_ARG,
                 "peer_all_reduce_rms_norm_quant: null pointer");
  const bool emu = rank == -1;  // single-GPU emulation of all ranks (see include/b200_vllm.h)
  B200_CHECK_ARG(world_size >= 2 && world_size <= kArMaxWorld && (emu || (rank >= 0 && rank < world_size)),
                 B200_ERR_INVALID_ARG,
                 "peer_all_reduce_rms_norm_quant: world size %d / rank %d not supported (2..8 ranks)", world_size, rank);
  B200_CHECK_ARG(num_tokens > 0 && hidden > 0 && hidden % 8 == 0 && hidden <= 49152 && byte_offset % 16 == 0,
                 B200_ERR_INVALID_ARG, "peer_all_reduce_rms_norm_quant: hidden must be a multiple of 8 (<= 49152)");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(weight) % 16 == 0 && reinterpret_cast<uintptr_t>(residual) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(result) % 4 == 0,
                 B200_ERR_ALIGNMENT, "peer_all_reduce_rms_norm_quant: weight/residual must be 16-byte aligned");
  // one row per CTA up to 128 CTAs (every row pays a peer-load / switch round trip: rows in flight, not bytes, set
  // the duration), bounded by the signal pad and by one co-resident wave
  int64_t grid = num_tokens < 128 ? num_tokens : 128;
  const int64_t max_grid = signal_pad_bytes / (kArSlotWords * 4);
  if (emu && grid > sm_count() / world_size) grid = sm_count() / world_size;  // one cooperative wave
  if (grid > max_grid) grid = max_grid;
  B200_CHECK_ARG(grid >= 1, B200_ERR_INVALID_ARG, "peer_all_reduce_rms_norm_quant: signal pad too small");
  ArArgs a;
  for (int q = 0; q < kArMaxWorld; ++q) {
    a.bufs[q] = q < world_size ? reinterpret_cast<void*>(peer_buffers[q] + byte_offset) : nullptr;
    a.pads[q] = q < world_size ? reinterpret_cast<uint32_t*>(peer_signal_pads[q]) : nullptr;
  }
  a.rank = emu ? 0 : rank;
  a.world = world_size;
  a.numel = num_tokens * hidden;
  a.spin_limit = ar_spin_limit();
  a.emu = emu ? 1 : 0;
  a.emu_stride = num_tokens * hidden;
  a.mc = (multicast_ptr && !emu) ? reinterpret_cast<const void*>(multicast_ptr + byte_offset) : nullptr;
  B200_CHECK_ARG(!a.mc || hidden % (8 * world_size) == 0, B200_ERR_INVALID_ARG,
                 "nvls_all_reduce_rms_norm_quant: hidden must be a multiple of 8 * world_size");
  return dispatch16f(dtype, "peer_all_reduce_rms_norm_quant", [&](auto tag) {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(peer_allreduce_rms_norm_quant_kernel<T>), 49152 * 4)) return rc_attr;
    const int threads = row_threads(hidden / 8);
    launch_kernel_ex(peer_allreduce_rms_norm_quant_kernel<T>, dim3((unsigned)grid, emu ? world_size : 1), dim3(threads),
                     hidden * 4, (cudaStream_t)stream, 1, emu, (uint8_t*)result, scales, a, (const T*)weight,
                     (T*)residual, (int)num_tokens, hidden, epsilon);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  });
}

extern "C" int b200_peer_all_reduce_rms_norm_quant(void* result, float* scales, const uint64_t* peer_buffers,
                                                   const uint64_t* peer_signal_pads, int rank, int world_size,
                                                   int64_t byte_offset, int64_t signal_pad_bytes, const void* weight,
                                                   void* residual, float epsilon, int64_t num_tokens, int hidden,
                                                   int dtype, b200_stream_t stream) {
  return ar_rms_norm_quant_impl(result, scales, peer_buffers, peer_signal_pads, rank, world_size, byte_offset,
                                signal_pad_bytes, weight, residual, epsilon, num_tokens, hidden, dtype, stream, 0);
}

extern "C" int b200_nvls_all_reduce_rms_norm_quant(void* result, float* scales, uint64_t multicast_ptr,
                                                   const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                                                   int rank, int world_size, int64_t byte_offset,
                                                   int64_t signal_pad_bytes, const void* weight, void* residual,
                                                   float epsilon, int64_t num_tokens, int hidden, int dtype,
                                                   b200_stream_t stream) {
  B200_CHECK_ARG(multicast_ptr != 0 && (multicast_ptr + byte_offset) % 16 == 0, B200_ERR_INVALID_ARG,
                 "nvls_all_reduce_rms_norm_quant: no (aligned) multicast mapping - NVLS unavailable");
  return ar_rms_norm_quant_impl(result, scales, peer_buffers, peer_signal_pads, rank, world_size, byte_offset,
                                signal_pad_bytes, weight, residual, epsilon, num_tokens, hidden, dtype, stream,
                                multicast_ptr);
}
