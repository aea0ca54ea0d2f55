// KV-cache write (reshape_and_cache_flash semantics) and slot-mapping kernels.
// Reference behaviour: csrc/libtorch_stable/cache_kernels.cu:320-406,801-871,
// fp8 conversion csrc/quantization/w8a8/fp8/nvidia/quant_utils.cuh:480-532,
// slot mapping vllm/v1/worker/block_table.py:412-475.
//
// B200 design: the op is launch-latency bound in decode (T = batch tokens), so
// the kernel does the minimum: one CTA per token, one 16-byte vector per
// thread-iteration covering K and V of all heads in a single pass (the
// reference makes two passes), loads via the read-only path with no L1
// allocation, scalar fallback chosen per (token, head) segment when either side
// is not 16-byte aligned (H=13, D=46 style shapes).
#include "common.cuh"

namespace b200 {

template <int kKv> __device__ __forceinline__ uint8_t to_fp8(float x, float scale) {
  // IEEE fp32 divide then RN-even satfinite: identical to the reference's
  // __nv_cvt_float_to_fp8(float(x) / scale, __NV_SATFINITE, ...).
  float q = __fdiv_rn(x, scale);
  return (uint8_t)__nv_cvt_float_to_fp8(q, __NV_SATFINITE, kKv == B200_KV_FP8_E5M2 ? __NV_E5M2 : __NV_E4M3);
}

__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// scalar_t: source element; kKv: B200_KV_*.  Cache element = scalar_t (auto)
// or uint8_t (fp8).
template <typename scalar_t, int kKv>
__global__ void __launch_bounds__(256) cache_write_kernel(
    const scalar_t* __restrict__ key, const scalar_t* __restrict__ value, void* __restrict__ key_cache,
    void* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int num_heads, int head_size,
    int block_size, int64_t key_stride, int64_t value_stride, int64_t block_stride, int64_t page_stride,
    int64_t head_stride, const float* __restrict__ k_scale, const float* __restrict__ v_scale,
    int kv_scale_stride) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  using cache_t = typename std::conditional<kKv == B200_KV_AUTO, scalar_t, uint8_t>::type;
  constexpr int kVec = 16 / sizeof(scalar_t);  // source elements per 16-byte vector
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // PAD_SLOT_ID
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int64_t dst_base = block_idx * block_stride + block_off * page_stride;

  const bool d_vec = (head_size % kVec) == 0;
  const int vec_per_head = d_vec ? head_size / kVec : 0;
  // ---- vector path: segments (which, head) whose src and dst are aligned ----
  const int total_vec = 2 * num_heads * vec_per_head;
  for (int v = threadIdx.x; v < total_vec; v += blockDim.x) {
    const int which = v / (num_heads * vec_per_head);  // 0 = K, 1 = V
    const int rem = v - which * num_heads * vec_per_head;
    const int head = rem / vec_per_head;
    const int chunk = rem - head * vec_per_head;
    const scalar_t* src = (which ? value + token * value_stride : key + token * key_stride) + head * head_size;
    cache_t* dst =
        reinterpret_cast<cache_t*>(which ? value_cache : key_cache) + dst_base + (int64_t)head * head_stride;
    const bool aligned = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(dst) & (kKv == B200_KV_AUTO ? 15 : (kVec - 1))) == 0);
    if (!aligned) continue;  // handled by the scalar pass below
    uint4 raw = ld_nc_v4(src + chunk * kVec);
    if constexpr (kKv == B200_KV_AUTO) {
      *reinterpret_cast<uint4*>(dst + chunk * kVec) = raw;
    } else {
      const float s = (which ? v_scale : k_scale)[head * kv_scale_stride];
      const scalar_t* e = reinterpret_cast<const scalar_t*>(&raw);
      uint8_t o[kVec];
#pragma unroll
      for (int i = 0; i < kVec; ++i) o[i] = to_fp8<kKv>(to_float(e[i]), s);
      if constexpr (kVec == 8) {
        *reinterpret_cast<uint2*>(dst + chunk * kVec) = *reinterpret_cast<uint2*>(o);
      } else {
        *reinterpret_cast<uint32_t*>(dst + chunk * kVec) = *reinterpret_cast<uint32_t*>(o);
      }
    }
  }
  // ---- scalar path: unaligned segments ------------------------------------
  const int total = 2 * num_heads * head_size;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int which = i / (num_heads * head_size);
    const int rem = i - which * num_heads * head_size;
    const int head = rem / head_size;
    const int d = rem - head * head_size;
    const scalar_t* src = (which ? value + token * value_stride : key + token * key_stride) + head * head_size;
    cache_t* dst =
        reinterpret_cast<cache_t*>(which ? value_cache : key_cache) + dst_base + (int64_t)head * head_stride;
    const bool aligned = d_vec && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(dst) & (kKv == B200_KV_AUTO ? 15 : (kVec - 1))) == 0);
    if (aligned) continue;
    if constexpr (kKv == B200_KV_AUTO) {
      dst[d] = src[d];
    } else {
      const float s = (which ? v_scale : k_scale)[head * kv_scale_stride];
      dst[d] = to_fp8<kKv>(to_float(src[d]), s);
    }
  }
}

// slot[t] = block_table[req][pos/bs]*bs + pos%bs, padding tail = -1.
__global__ void __launch_bounds__(256) slot_mapping_kernel(int64_t* __restrict__ slot_mapping, int64_t num_slots,
                                                           int64_t num_tokens,
                                                           const int32_t* __restrict__ query_start_loc,
                                                           int num_reqs, const int64_t* __restrict__ positions,
                                                           const int32_t* __restrict__ block_table,
                                                           int64_t block_table_stride, int block_size) {
  pdl_trigger();  // let the next kernel's prologue start
  pdl_wait();     // predecessors' writes are visible from here on
  // blockIdx.x < num_reqs: one request each; blockIdx.x == num_reqs: padding tail.
  if ((int)blockIdx.x == num_reqs) {
    for (int64_t t = num_tokens + threadIdx.x; t < num_slots; t += blockDim.x) slot_mapping[t] = -1;
    return;
  }
  const int req = blockIdx.x;
  const int start = query_start_loc[req], end = query_start_loc[req + 1];
  for (int t = start + threadIdx.x; t < end; t += blockDim.x) {
    const int64_t pos = positions[t];
    const int64_t blk = block_table[(int64_t)req * block_table_stride + pos / block_size];
    slot_mapping[t] = blk * block_size + pos % block_size;
  }
}

template <typename scalar_t>
static int launch_cache_write(const void* key, const void* value, void* key_cache, void* value_cache,
                              const int64_t* slot_mapping, int64_t num_tokens, int num_heads, int head_size,
                              int block_size, int64_t key_stride, int64_t value_stride, int64_t block_stride,
                              int64_t page_stride, int64_t head_stride, int kv_dtype, const float* k_scale,
                              const float* v_scale, int kv_scale_stride, cudaStream_t stream) {
  constexpr int kVec = 16 / sizeof(scalar_t);
  int work = 2 * num_heads * ((head_size % kVec) == 0 ? head_size / kVec : head_size);
  int threads = work < 32 ? 32 : (work > 256 ? 256 : ((work + 31) / 32) * 32);
  dim3 grid((unsigned)num_tokens), block(threads);
#define B200_CW(KV)                                                                                          \
  launch_kernel(cache_write_kernel<scalar_t, KV>, dim3(grid), dim3(block), 0, stream, 1,                                               \
      (const scalar_t*)key, (const scalar_t*)value, key_cache, value_cache, slot_mapping, num_heads,        \
      head_size, block_size, key_stride, value_stride, block_stride, page_stride, head_stride, k_scale,     \
      v_scale, kv_scale_stride)
  if (kv_dtype == B200_KV_AUTO) B200_CW(B200_KV_AUTO);
  else if (kv_dtype == B200_KV_FP8_E4M3) B200_CW(B200_KV_FP8_E4M3);
  else B200_CW(B200_KV_FP8_E5M2);
#undef B200_CW
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

}  // namespace b200

extern "C" int b200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                            void* value_cache, const int64_t* slot_mapping, int64_t num_tokens,
                                            int num_heads, int head_size, int block_size, int64_t key_stride,
                                            int64_t value_stride, int64_t block_stride, int64_t page_stride,
                                            int64_t head_stride, int src_dtype, int kv_dtype,
                                            const float* k_scale, const float* v_scale, int kv_scale_stride,
                                            b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0, B200_ERR_INVALID_ARG,
                 "reshape_and_cache_flash: bad sizes tokens=%lld heads=%d head_size=%d block_size=%d",
                 (long long)num_tokens, num_heads, head_size, block_size);
  B200_CHECK_ARG(kv_dtype == B200_KV_AUTO || kv_dtype == B200_KV_FP8_E4M3 || kv_dtype == B200_KV_FP8_E5M2,
                 B200_ERR_INVALID_ARG, "reshape_and_cache_flash: Unsupported data type of kv cache: %d", kv_dtype);
  B200_CHECK_ARG(kv_scale_stride == 0 || kv_scale_stride == 1, B200_ERR_INVALID_ARG,
                 "k_scale and v_scale must be of shape [1] or [num_heads]");
  B200_CHECK_ARG(kv_dtype == B200_KV_AUTO || (k_scale && v_scale), B200_ERR_INVALID_ARG,
                 "reshape_and_cache_flash: fp8 cache needs k_scale and v_scale");
  if (num_tokens == 0) return 0;
  B200_CHECK_ARG(key && value && key_cache && value_cache && slot_mapping, B200_ERR_INVALID_ARG,
                 "reshape_and_cache_flash: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  switch (src_dtype) {
    case B200_DT_BF16:
      return launch_cache_write<__nv_bfloat16>(key, value, key_cache, value_cache, slot_mapping, num_tokens,
                                               num_heads, head_size, block_size, key_stride, value_stride,
                                               block_stride, page_stride, head_stride, kv_dtype, k_scale, v_scale,
                                               kv_scale_stride, s);
    case B200_DT_F16:
      return launch_cache_write<__half>(key, value, key_cache, value_cache, slot_mapping, num_tokens, num_heads,
                                        head_size, block_size, key_stride, value_stride, block_stride,
                                        page_stride, head_stride, kv_dtype, k_scale, v_scale, kv_scale_stride, s);
    case B200_DT_F32:
      return launch_cache_write<float>(key, value, key_cache, value_cache, slot_mapping, num_tokens, num_heads,
                                       head_size, block_size, key_stride, value_stride, block_stride, page_stride,
                                       head_stride, kv_dtype, k_scale, v_scale, kv_scale_stride, s);
    default:
      set_error("Unsupported input type of kv cache: %d", src_dtype);
      return B200_ERR_INVALID_ARG;
  }
}

extern "C" int b200_compute_slot_mapping(int64_t* slot_mapping, int64_t num_slots, int64_t num_tokens,
                                         const int32_t* query_start_loc, int num_reqs, const int64_t* positions,
                                         const int32_t* block_table, int64_t block_table_stride, int block_size,
                                         b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(num_reqs >= 0 && block_size > 0 && num_slots >= 0, B200_ERR_INVALID_ARG,
                 "compute_slot_mapping: bad sizes");
  if (num_slots == 0) return 0;
  B200_CHECK_ARG(slot_mapping && query_start_loc && positions && block_table, B200_ERR_INVALID_ARG,
                 "compute_slot_mapping: null pointer");
  launch_kernel(slot_mapping_kernel, dim3(num_reqs + 1), dim3(256), 0, (cudaStream_t)stream, 1, 
      slot_mapping, num_slots, num_tokens, query_start_loc, num_reqs, positions, block_table, block_table_stride, block_size);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}
