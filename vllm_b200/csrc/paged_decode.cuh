// Declarations shared by the two paged-attention decode kernels (paged_decode.cu: 16-bit caches on
// mma.sync; paged_decode_fp8.cu: fp8 caches on tcgen05).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace b200 {

constexpr int kChunk = 16;  // tokens per TMA box (block_size must be a multiple)
constexpr int kWarps = 4;   // consumer warps per CTA
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct DecodeArgs {
  void* out;
  const void* query;
  const int32_t* block_table;
  const int32_t* seq_lens;
  const float* k_scale;
  const float* v_scale;
  float* out_lse;
  const float* out_scale;  // non-null: `out` is e4m3 bytes, o * (1 / *out_scale) (fused static output quant)
  const int32_t* token_to_seq;  // optional: block_table row per query token
  float* part_o;   // [seq][hq][split][D]
  float2* part_ml; // [seq][hq][split]
  int64_t q_stride_seq, q_stride_head, out_stride_seq, out_stride_head, block_table_stride;
  int num_q_heads, num_kv_heads, group, qgroups;  // group = Hq/Hkv, qgroups = ceil(group/16)
  int block_size, num_splits, sliding_window;
  float scale, soft_cap;
  int max_seq_len;  // upper bound from the caller (<= 0: unknown)
  int eff_splits;   // fp8 kernel: splits that actually get work (<= num_splits; the others write empty partials)
};

// one fp32 value -> e4m3 byte, round-to-nearest-even, saturating (scaled_fp8_conversion, fp8/common.cuh:57-76)
__device__ __forceinline__ uint8_t to_e4m3_sat(float x) {
  return (uint8_t)__nv_cvt_float_to_fp8(fmaxf(-448.f, fminf(x, 448.f)), __NV_SATFINITE, __NV_E4M3);
}

__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// fp8-cache kernel (tcgen05): launches the main kernel only, the caller runs the split combine
template <typename QT, int D, int KV>
int launch_decode_fp8(const CUtensorMap& tk, const CUtensorMap& tv, const DecodeArgs& a, int num_seqs,
                      cudaStream_t stream);

}  // namespace b200
