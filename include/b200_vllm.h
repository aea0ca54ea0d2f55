/*
 * b200_vllm.h — C ABI of libb200vllm.so, the sm_100a implementation of vLLM's
 * decode-step hot path.  Every entry point takes plain device pointers, sizes,
 * strides (in ELEMENTS unless a name ends in _bytes) and a CUDA stream; no
 * torch types cross this boundary.  All outputs are caller-allocated, nothing
 * is allocated or synchronised inside (CUDA-graph capturable), and every
 * function returns 0 on success or a negative B200_ERR_* code *before* any
 * launch when a precondition fails (the Python shim turns it into
 * RuntimeError, like STD_TORCH_CHECK does in the reference).
 *
 * Each function cites the reference interface it replaces, as path:line under
 * /root/reference (vllm-project/vllm).
 */
#ifndef B200_VLLM_H_
#define B200_VLLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200_stream_t; /* cudaStream_t */

/* ---- error codes ------------------------------------------------------- */
#define B200_OK 0
#define B200_ERR_INVALID_ARG (-1)   /* shape / dtype / enum precondition     */
#define B200_ERR_ALIGNMENT (-2)     /* pointer or stride alignment           */
#define B200_ERR_UNSUPPORTED (-3)   /* valid in the reference, not built here */
#define B200_ERR_CUDA (-4)          /* a CUDA runtime / driver call failed   */
#define B200_ERR_WORKSPACE (-5)     /* workspace too small                   */

/* ---- dtype enums ------------------------------------------------------- */
/* activation / query / output scalar types */
#define B200_DT_F32 0
#define B200_DT_F16 1
#define B200_DT_BF16 2
/* kv_cache_dtype: "auto" (= same as activations), "fp8"/"fp8_e4m3", "fp8_e5m2"
 * (csrc/attention/dtype_fp8.cuh:21, quant_utils.cuh:575-613) */
#define B200_KV_AUTO 0
#define B200_KV_FP8_E4M3 1
#define B200_KV_FP8_E5M2 2

/* Library / build identification. */
const char* b200_version(void);          /* "b200vllm <n> sm_100a"            */
const char* b200_last_error(void);       /* thread-local detail for last <0  */
int b200_num_kernels_launched(void);     /* running launch counter (bench)   */

/* ------------------------------------------------------------------------
 * KV-cache write.
 * Replaces: reshape_and_cache_flash
 *   host   csrc/libtorch_stable/cache_kernels.cu:801-871
 *   kernel csrc/libtorch_stable/cache_kernels.cu:320-406
 *   schema csrc/libtorch_stable/torch_bindings.cpp:915-921
 *   python vllm/_custom_ops.py:2631-2650
 * For every t < num_tokens with slot = slot_mapping[t] >= 0:
 *   key_cache  [slot / block_size][slot % block_size][h][:] = cvt(key  [t][h][:])
 *   value_cache[slot / block_size][slot % block_size][h][:] = cvt(value[t][h][:])
 * cvt = identity (B200_KV_AUTO) or fp8(float(x) / scale) RN-even, satfinite
 * (quant_utils.cuh:480-503).  The cache is addressed with explicit strides so
 * NHD, HND and the K|V-interleaved page layouts all work:
 *   addr = cache + block*block_stride + offset*page_stride + h*head_stride.
 * k_scale/v_scale: device fp32, 1 element (kv_scale_stride==0) or one per
 * head (kv_scale_stride==1); ignored for B200_KV_AUTO.
 * ---------------------------------------------------------------------- */
int b200_reshape_and_cache_flash(
    const void* key, const void* value, void* key_cache, void* value_cache,
    const int64_t* slot_mapping, int64_t num_tokens, int num_heads,
    int head_size, int block_size, int64_t key_stride, int64_t value_stride,
    int64_t block_stride, int64_t page_stride, int64_t head_stride,
    int src_dtype, int kv_dtype, const float* k_scale, const float* v_scale,
    int kv_scale_stride, b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Slot mapping (producer of the cache-write indices; integer, bit-exact).
 * Replaces: BlockTable.compute_slot_mapping + its Triton kernel
 *   vllm/v1/worker/block_table.py:201-229,412-475
 * slot[t] = block_table[req(t)][pos[t] / block_size] * block_size
 *           + pos[t] % block_size          for t in [query_start_loc[r],
 *                                          query_start_loc[r+1]), r < num_reqs
 * slot[t] = -1 (PAD_SLOT_ID)               for num_tokens <= t < num_slots
 *                                          (CUDA-graph padding tail).
 * ---------------------------------------------------------------------- */
int b200_compute_slot_mapping(
    int64_t* slot_mapping, int64_t num_slots, int64_t num_tokens,
    const int32_t* query_start_loc, int num_reqs, const int64_t* positions,
    const int32_t* block_table,
    int64_t block_table_stride, int block_size, b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Paged-attention decode (one query token per sequence), split-KV + combine,
 * GQA-grouped: one CTA serves all q-heads of one kv-head so K/V pages are
 * read once.
 * Replaces (same math, same cache layout):
 *   legacy   paged_attention_v1/v2     vllm/_custom_ops.py:120-163 (ROCm port
 *            csrc/rocm/attention.cu:320-1520; design docs/design/paged_attention.md)
 *   current  kernel_unified_attention 3D + reduce_segments
 *            vllm/v1/attention/ops/triton_unified_attention.py:179-681,685-771
 *   caller   FlashAttentionImpl.forward vllm/v1/attention/backends/flash_attn.py:1085-1110
 *
 * out[s][h][:] = softmax(scale * q[s][h] . K[s][0:len_s]^T) V[s][0:len_s]
 * with K/V rows of sequence s found through block_table[s][0:ceil(len/bs)].
 * fp8 caches: K = k_scale * fp8, V = v_scale * fp8 (per-tensor device scalars).
 * Sequences with seq_lens[s] == 0 produce zeros.
 *
 * workspace: fp32 scratch for the split partials; size from
 * b200_paged_decode_workspace_bytes().  num_splits == 0 lets the library pick
 * (deterministic in (num_seqs, num_kv_heads, max_seq_len) only -> graph safe).
 * out_lse (optional, fp32 [num_seqs][num_q_heads]): natural-log LSE per head
 * (needed by decode-context-parallel callers, backend.py:791-816).
 * ---------------------------------------------------------------------- */
typedef struct b200_paged_decode_params {
  void* out;                 /* [num_seqs][num_q_heads][head_size], q dtype  */
  const void* query;         /* [num_seqs][num_q_heads][head_size]           */
  const void* key_cache;     /* paged, see strides                           */
  const void* value_cache;
  const int32_t* block_table; /* [num_seqs][block_table_stride]              */
  const int32_t* seq_lens;    /* [num_seqs]                                  */
  const float* k_scale;       /* device scalar or NULL (=1)                  */
  const float* v_scale;
  float* out_lse;             /* optional                                    */
  const int32_t* token_to_seq; /* optional [num_seqs]: block_table row of each
                                 query token (NULL = identity).  Lets a chunked-
                                 prefill / mixed batch run as one single-query
                                 attention per token with its own causal length
                                 in seq_lens[token].                          */
  void* workspace;
  size_t workspace_bytes;
  int64_t q_stride_seq, q_stride_head;     /* elements                      */
  int64_t out_stride_seq, out_stride_head; /* elements                      */
  int64_t kv_block_stride, kv_page_stride, kv_head_stride; /* elements, K  */
  int64_t v_block_stride, v_page_stride, v_head_stride;    /* elements, V  */
  int64_t block_table_stride;
  int64_t num_blocks;         /* extent of the cache (TMA bound)             */
  int num_seqs, num_q_heads, num_kv_heads, head_size, block_size;
  int max_seq_len;            /* upper bound on seq_lens (host knowledge)    */
  int q_dtype;                /* B200_DT_F16 / B200_DT_BF16                  */
  int kv_dtype;               /* B200_KV_*                                   */
  int num_splits;             /* 0 = auto                                    */
  float scale;                /* softmax scale                               */
  float logits_soft_cap;      /* 0 = off                                     */
  int sliding_window;         /* 0 = off; else attend to last W tokens       */
  /* Fused static FP8 output quant (AttentionImpl.forward(output_scale=...), vllm/v1/attention/backend.py:910-922,
   * quant key kFp8StaticTensorSym): when non-NULL, `out` is e4m3 BYTES (strides in bytes) and every output element
   * is stored as e4m3_rn_sat(o * (1 / *output_scale)) straight from the fp32 accumulator - no bf16 round trip and no
   * separate quant launch in front of o_proj.  NULL = out in the query dtype. */
  const float* output_scale;
} b200_paged_decode_params;

size_t b200_paged_decode_workspace_bytes(int num_seqs, int num_q_heads,
                                         int head_size, int num_splits);
int b200_paged_decode_pick_splits(int num_seqs, int num_kv_heads,
                                  int max_seq_len, int sm_count);
/* The same with the cache geometry (resident CTAs per SM depend on it); the 4-argument form assumes
 * head_size 128 with an fp8 cache. */
int b200_paged_decode_pick_splits_ex(int num_seqs, int num_kv_heads, int max_seq_len, int sm_count,
                                     int head_size, int kv_dtype);
int b200_paged_attention_decode(const b200_paged_decode_params* p,
                                b200_stream_t stream);

/* Allocation-free variant of b200_paged_attention_decode.  It has NO counterpart in the reference:
 * it was added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Same inputs, outputs and results as b200_paged_attention_decode; p->workspace / workspace_bytes are
 * IGNORED: the split-KV partials live in a file-scope `static __device__` array of
 * b200_paged_decode_static_scratch_bytes() = 32 MiB that every GPU which loads the library holds
 * until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree / allocator).
 * A problem whose partials (b200_paged_decode_workspace_bytes of its split count) exceed 32 MiB is
 * rejected with B200_ERR_WORKSPACE before any launch.
 * Every call on a device shares the one array, which keeps the previous call's contents (nothing
 * in it needs clearing: each slot read is written first by the same call).  Calls that OVERLAP on
 * the GPU (two streams, or two host threads launching concurrently) race on it; serialise them on
 * one stream or use the caller-workspace entry point. */
size_t b200_paged_decode_static_scratch_bytes(void);
int b200_paged_attention_decode_static_scratch(const b200_paged_decode_params* p,
                                               b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Two-way (out, lse) merge.
 * Replaces: merge_attn_states
 *   csrc/libtorch_stable/attention/merge_attn_states.cu:19-257,259-362
 *   schema csrc/libtorch_stable/torch_bindings.cpp:354-362
 * output[t][h][:] = (e^{lp-m} P + e^{ls-m} S) / (e^{lp-m} + e^{ls-m}),
 * output_lse = m + log(sum); +-inf LSEs are treated as -inf.
 * prefix/suffix outputs are [tokens][heads][head_size] (stride given in
 * elements per token / per head), LSEs are fp32 [heads][tokens].
 * ---------------------------------------------------------------------- */
int b200_merge_attn_states(void* output, float* output_lse,
                           const void* prefix_output, const float* prefix_lse,
                           const void* suffix_output, const float* suffix_lse,
                           int64_t num_tokens, int num_heads, int head_size,
                           int64_t out_stride_tok, int64_t out_stride_head,
                           int64_t p_stride_tok, int64_t p_stride_head,
                           int64_t s_stride_tok, int64_t s_stride_head,
                           int64_t lse_stride_head, int dtype,
                           b200_stream_t stream);

/* ------------------------------------------------------------------------
 * FP8 (e4m3) activation quantisation.
 * Replaces: csrc/libtorch_stable/quantization/w8a8/fp8/common.cu
 *   static_scaled_fp8_quant            :183-306  (kernel :13-76)
 *   dynamic_scaled_fp8_quant           :308-354  (kernels :78-135)
 *   dynamic_per_token_scaled_fp8_quant :356-420  (kernel :137-179)
 *   conversion csrc/quantization/w8a8/fp8/common.cuh:57-76
 *   schemas csrc/libtorch_stable/torch_bindings.cpp:645-659
 * input [num_tokens][hidden] (row stride in elements), out e4m3 bytes.
 * ---------------------------------------------------------------------- */
/* out = e4m3(clamp(x * (1/scale[g]), +-448)); scale indexed
 * scale[(m / group_m) * scale_stride_i + (n / group_n) * scale_stride_j]. */
int b200_static_scaled_fp8_quant(void* out, const void* input,
                                 const float* scale, int64_t num_tokens,
                                 int hidden, int64_t in_row_stride,
                                 int64_t out_row_stride, int group_m,
                                 int group_n, int64_t scale_stride_i,
                                 int64_t scale_stride_j, int dtype,
                                 b200_stream_t stream);
/* scale[0] = max|x| / 448 over the whole tensor; out = e4m3(clamp(x*(1/scale))) */
int b200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale,
                                  int64_t num_tokens, int hidden,
                                  int64_t in_row_stride, int64_t out_row_stride,
                                  int dtype, b200_stream_t stream);
/* scale[m] = max(min(amax_m, ub) / 448, 1/(448*512)); out = e4m3(clamp(x/scale[m])) */
int b200_dynamic_per_token_scaled_fp8_quant(
    void* out, const void* input, float* scales, const float* scale_ub,
    int64_t num_tokens, int hidden, int64_t in_row_stride,
    int64_t out_row_stride, int dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------
 * NVFP4 activation quantisation (e2m1 codes + e4m3 scale per 16 elements).
 * Replaces: scaled_fp4_quant(.out)
 *   csrc/libtorch_stable/quantization/fp4/nvfp4_quant_kernels.cu:39-113,188-281
 *   math   csrc/libtorch_stable/quantization/fp4/nvfp4_utils.cuh:221-302
 *   SF offset nvfp4_utils.cuh:168-199
 *   schema csrc/libtorch_stable/torch_bindings.cpp:218-222
 * input [m][n] fp16/bf16 contiguous; output [m][out_n/2] bytes (two codes per
 * byte, even element in the low nibble), out_n >= n (columns past n encode 0).
 * swizzled != 0: output_sf is the [ceil(m/128)][ceil(out_n/64)][32][4][4]
 * tensor-core layout and every padded SF is written (0); else row-major
 * [m][out_n/16].
 * ---------------------------------------------------------------------- */
int b200_scaled_fp4_quant(void* output, void* output_sf, const void* input,
                          const float* input_global_scale, int m, int n,
                          int out_n, int swizzled, int dtype,
                          b200_stream_t stream);

/* Fused NVFP4 activation producers: the quantiser above with its producer folded in, always with the
 * swizzled tensor-core scale layout (every padded scale written as 0).
 *   b200_silu_and_mul_nvfp4_quant  input [m][2d] = [gate | up]; quantises T(silu(gate) * up), the product formed in
 *     fp32 and rounded to T once.  Replaces: silu_and_mul_nvfp4_quant
 *       csrc/libtorch_stable/quantization/fp4/activation_nvfp4_quant_fusion_kernels.cu:39-113 (+ launcher),
 *       math nvfp4_utils.cuh:304-329, schema csrc/libtorch_stable/torch_bindings.cpp:251-253.
 *   b200_rms_norm_nvfp4_quant      (residual += input;) quantises RMSNorm(.) * weight.  The reference has no such
 *     kernel (its NVFP4 path runs fused_add_rms_norm, csrc/layernorm_kernels.cu, then scaled_fp4_quant); this is
 *     bit-identical to b200_rms_norm followed by b200_scaled_fp4_quant.  residual may be NULL. */
int b200_silu_and_mul_nvfp4_quant(void* result, void* result_block_scale, const void* input,
                                  const float* input_global_scale, int m, int d, int dtype, b200_stream_t stream);
int b200_rms_norm_nvfp4_quant(void* result, void* result_block_scale, const void* input, const void* weight,
                              void* residual, const float* input_global_scale, int m, int hidden,
                              int64_t in_row_stride, float epsilon, int dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------
 * FP8 W8A8 GEMM (tcgen05 / TMEM / TMA).
 * Replaces: cutlass_scaled_mm
 *   host checks csrc/libtorch_stable/quantization/w8a8/cutlass/scaled_mm_entry.cu:197-270
 *   sm100 configs .../c3x/scaled_mm_sm100_fp8_dispatch.cuh:100-194,243-296
 *   epilogue csrc/libtorch_stable/cutlass_extensions/epilogue/scaled_mm_epilogues_c3x.hpp:150-183
 *   schema csrc/libtorch_stable/torch_bindings.cpp:116-119
 * out[M][N] = cast( a_scales[m|0] * b_scales[n|0] * sum_k A[m][k] B[k][n] ) + bias[n]
 * A [M][K] e4m3 row-major (lda elements), B "column-major [K][N]" i.e. stored
 * [N][K] K-contiguous (ldb elements), out row-major (ldc), fp32 accumulate.
 * a_scales has 1 or M entries, b_scales 1 or N.  K % 16 == 0, N % 16 == 0.
 * Differs from the reference in one precondition: for M <= 256 the WEIGHT operand `b` is prefetched before the
 * kernel waits for its stream predecessors (programmatic dependent launch, see b200_set_pdl), so `b` must not be
 * written by a kernel of this library that is still in flight on the same stream when the GEMM is launched
 * (weights quantised at load time - the only case in vLLM - satisfy this; b200_set_pdl(0) removes the condition).
 * ---------------------------------------------------------------------- */
int b200_scaled_mm_fp8(void* out, const void* a, const void* b,
                       const float* a_scales, const float* b_scales,
                       const void* bias, int M, int N, int K, int64_t lda,
                       int64_t ldb, int64_t ldc, int a_scales_numel,
                       int b_scales_numel, int out_dtype, void* workspace,
                       size_t workspace_bytes, b200_stream_t stream);
/* Scratch size for b200_scaled_mm_fp8: always 0 since round 2 - the decode regime's split-K fix-up is a
 * reduce-scatter over distributed shared memory inside one thread-block cluster, and the large-M kernels need
 * none.  `workspace` / `workspace_bytes` of b200_scaled_mm_fp8 are accepted and ignored (kept so that callers
 * written against the round-1 ABI keep working). */
size_t b200_scaled_mm_workspace_bytes(int M, int N);

/* ------------------------------------------------------------------------
 * NVFP4 block-scaled GEMM (tcgen05.mma kind::mxf4nvf4, scale_vec::4X).
 * Replaces: cutlass_scaled_fp4_mm
 *   entry  csrc/libtorch_stable/quantization/fp4/nvfp4_scaled_mm_entry.cu:41-69
 *   kernel csrc/libtorch_stable/quantization/fp4/nvfp4_scaled_mm_kernels.cu:46-137,190-248
 *   checks nvfp4_scaled_mm_kernels.cu:279-347
 *   schema csrc/libtorch_stable/torch_bindings.cpp:187-190
 * out[M][N] = cast( alpha * sum_k (A4[m][k] SFA[m][k/16]) (B4[n][k] SFB[n][k/16]) )
 * A [M][K/2] bytes, B [N][K/2] bytes (K-contiguous, two e2m1 per byte),
 * SFA / SFB e4m3 in the swizzled 128x4 layout with padded shapes
 * [ru(M,128)][ru(K/16,4)] / [ru(N,128)][ru(K/16,4)].  K % 32 == 0, N % 32 == 0.
 * alpha: device fp32 scalar.
 * ---------------------------------------------------------------------- */
int b200_scaled_mm_nvfp4(void* out, const void* a, const void* b,
                         const void* sfa, const void* sfb, const float* alpha,
                         int M, int N, int K, int64_t ldc, int out_dtype,
                         b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Element-wise companions of the quantised linears on the decode step
 * (SURVEY 8f-1).  Not on the first-bar path; provided so a whole decode layer
 * can run on this library.
 *   rms_norm / fused_add_rms_norm  csrc/libtorch_stable/layernorm_kernels.cu
 *   silu_and_mul                   csrc/libtorch_stable/activation_kernels.cu
 *   rotary_embedding (neox)        csrc/libtorch_stable/pos_encoding_kernels.cu
 * ---------------------------------------------------------------------- */
int b200_rms_norm(void* out, const void* input, const void* weight,
                  void* residual /* nullable: fused add, updated in place */,
                  int64_t num_tokens, int hidden, int64_t in_row_stride,
                  float eps, int dtype, b200_stream_t stream);
int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d,
                      int dtype, b200_stream_t stream);
int b200_rotary_embedding_neox(void* query, void* key, const int64_t* positions,
                               const void* cos_sin_cache, int64_t num_tokens,
                               int num_q_heads, int num_kv_heads, int head_size,
                               int rot_dim, int64_t q_stride_tok,
                               int64_t k_stride_tok, int dtype,
                               b200_stream_t stream);

/* ------------------------------------------------------------------------
 * Fused element-wise neighbours: one launch instead of two or
 * three, activations touched once.
 *
 * b200_rms_norm_dynamic_per_token_quant replaces
 *   rms_norm_dynamic_per_token_quant
 *   csrc/libtorch_stable/quantization/fused_kernels/fused_layernorm_dynamic_per_token_quant.cu:50-91,174-230
 *   math   .../fused_kernels/layernorm_utils.cuh (compute_rms, compute_dynamic_per_token_scales, norm_and_quant)
 *   schema csrc/libtorch_stable/torch_bindings.cpp:394-396
 *   x = input (+ residual, fp32; residual <- T(x)); rms = rsqrt(mean(x^2) + eps);
 *   y = T(T(x * rms) * weight); scale[m] = max(min(amax|y|, ub) / 448, 1/(448*512));
 *   result = e4m3(y / scale[m]).  result/residual rows are contiguous [hidden].
 *
 * b200_silu_and_mul_dynamic_per_token_quant: bit-identical to silu_and_mul
 *   (csrc/libtorch_stable/activation_kernels.cu) followed by
 *   dynamic_per_token_scaled_fp8_quant (fp8/common.cu:137-179); input [tokens][2d].
 *
 * b200_rotary_embedding_and_cache_write: bit-identical to rotary_embedding (neox,
 *   csrc/libtorch_stable/pos_encoding_kernels.cu) on query/key in place followed by
 *   reshape_and_cache_flash of (key, value); the fused entry behind
 *   AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:932-980).
 *   Tokens t >= num_slots or with slot -1 are rotated but not cached.
 * ---------------------------------------------------------------------- */
int b200_rms_norm_dynamic_per_token_quant(void* result, const void* input, const void* weight,
                                          float* scales, float epsilon, const float* scale_ub,
                                          void* residual /* nullable, updated in place */,
                                          int64_t num_tokens, int hidden, int64_t in_row_stride,
                                          int dtype, b200_stream_t stream);
int b200_silu_and_mul_dynamic_per_token_quant(void* result, float* scales, const void* input,
                                              int64_t num_tokens, int d, int dtype,
                                              b200_stream_t stream);
int b200_rotary_embedding_and_cache_write(
    void* query, void* key, const void* value, const int64_t* positions, const void* cos_sin_cache,
    void* key_cache, void* value_cache, const int64_t* slot_mapping, int64_t num_tokens,
    int64_t num_slots, int num_q_heads, int num_kv_heads, int head_size, int rot_dim,
    int64_t q_stride_tok, int64_t k_stride_tok, int64_t v_stride_tok, int block_size,
    int64_t block_stride, int64_t page_stride, int64_t head_stride, int dtype, int kv_dtype,
    const float* k_scale, const float* v_scale, int kv_scale_stride, b200_stream_t stream);

/* -------------------------------------------------------------------------
 * In-place two-shot all-reduce (sum) over NVLink peer memory, for the row-parallel linears of a
 * tensor-parallel step.  Replaces tensor_model_parallel_all_reduce (vllm/distributed/communication_op.py;
 * the reference's small-message fast path is csrc/custom_all_reduce.cuh).
 *   peer_buffers[world_size]      device addresses of every rank's buffer (same size, peer-mapped by the host,
 *                                 e.g. torch symmetric memory); the tensor starts byte_offset into each
 *   peer_signal_pads[world_size]  device addresses of every rank's zero-initialised signal pad
 * numel 16-bit elements, multiple of 8 * world_size.  Every rank must call it in the same order.  The sum is
 * taken in rank order with fp32 accumulation: identical on all ranks and run to run.
 *
 * rank == -1 (all three entries): single-GPU emulation for tests on a one-GPU box.  All `world_size` buffers and
 * signal pads live on the current device and ONE cooperative launch runs every rank (blockIdx.y = rank, so the
 * cross-rank barriers cannot deadlock on residency).  Per-rank outputs are stacked: `out` is [world][numel];
 * `result` / `residual` are [world][tokens][hidden] and `scales` [world][tokens].
 * ---------------------------------------------------------------------- */
int b200_peer_all_reduce(const uint64_t* peer_buffers, const uint64_t* peer_signal_pads, int rank,
                         int world_size, int64_t numel, int dtype, int64_t byte_offset,
                         int64_t signal_pad_bytes, b200_stream_t stream);
/* One-shot variant: every rank reads all buffers and writes the sum to its local `out` (the buffers are left
 * untouched).  Needs only a start barrier, PROVIDED consecutive calls alternate between at least two buffers
 * (byte_offset); meant for 2-4 ranks where reading (world-1) x the message is cheaper than a second barrier. */
int b200_peer_all_reduce_out(void* out, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                             int rank, int world_size, int64_t numel, int dtype, int64_t byte_offset,
                             int64_t signal_pad_bytes, b200_stream_t stream);
/* One-shot all-reduce of a row-parallel GEMM output fused with what follows it in the decoder layer:
 * residual += sum; result, scales = dynamic per-token FP8 quant of RMSNorm(residual) * weight
 * (= b200_peer_all_reduce_out followed by b200_rms_norm_dynamic_per_token_quant with a residual, bit-identical to
 * that composition).  Same buffer-alternation rule as b200_peer_all_reduce_out. */
int b200_peer_all_reduce_rms_norm_quant(void* result, float* scales, const uint64_t* peer_buffers,
                                        const uint64_t* peer_signal_pads, int rank, int world_size,
                                        int64_t byte_offset, int64_t signal_pad_bytes, const void* weight,
                                        void* residual, float epsilon, int64_t num_tokens, int hidden, int dtype,
                                        b200_stream_t stream);

/* NVLS variants (in-switch reduction): ONE `multimem.ld_reduce` per 16 bytes instead of world_size peer loads -
 * the NVSwitch adds the ranks' copies with fp32 accumulation and returns the sum rounded once to the element type.
 *   b200_nvls_all_reduce                 in place, TWO-shot through the switch: rank r reduces slice r and writes it
 *                                        back with one multicast store (`multimem.st`) that the switch replicates to
 *                                        every rank; per-GPU link traffic ~ the message size whatever the world size.
 *   b200_nvls_all_reduce_out             one-shot into a local `out` (every rank reduces the whole message).
 *   b200_nvls_all_reduce_rms_norm_quant  two-shot per CTA (each CTA owns rows: reduce + multicast its column slice,
 *                                        second barrier, then residual add + RMSNorm + FP8 quant on the local rows);
 *                                        hidden % (8 * world_size) == 0.  The summed GEMM output is left in the buffer.  multicast_ptr = the multicast mapping of the same symmetric allocation
 * (torch.distributed._symmetric_memory handle.multicast_ptr; 0 when the fabric has no multicast support -> error).
 * Reference analogue: vllm/distributed/device_communicators/symm_mem.py:137-150 (torch.ops.symm_mem.
 * multimem_all_reduce_) and the fused AR+RMSNorm of flashinfer_all_reduce.py.  Same call-order and
 * buffer-alternation rules as the peer-load entries; world sizes 2..8, incl. the fused form at 8 ranks. */
int b200_nvls_all_reduce(uint64_t multicast_ptr, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                         int rank, int world_size, int64_t numel, int dtype, int64_t byte_offset,
                         int64_t signal_pad_bytes, b200_stream_t stream);
int b200_nvls_all_reduce_out(void* out, uint64_t multicast_ptr, const uint64_t* peer_buffers,
                             const uint64_t* peer_signal_pads, int rank, int world_size, int64_t numel,
                             int dtype, int64_t byte_offset, int64_t signal_pad_bytes, b200_stream_t stream);
int b200_nvls_all_reduce_rms_norm_quant(void* result, float* scales, uint64_t multicast_ptr,
                                        const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                                        int rank, int world_size, int64_t byte_offset, int64_t signal_pad_bytes,
                                        const void* weight, void* residual, float epsilon, int64_t num_tokens,
                                        int hidden, int dtype, b200_stream_t stream);

/* Barrier-free ("Lamport") all-reduce over peer memory: a pushed datum is its own arrival flag (receive buffers are
 * pre-filled with a sentinel no pushed word can have), so a shot costs ONE one-way NVLink latency instead of a
 * release + flag handshake (profiles/r02_allreduce.md).  One-shot up to 4 ranks, two-shot (reduce-scatter push, then
 * all-gather push) above.  Reference analogue: FlashInfer trtllm_allreduce_fusion behind
 * vllm/distributed/device_communicators/flashinfer_all_reduce.py and compilation/passes/fusion/allreduce_rms_fusion.py.
 *   workspace: one region per rank of b200_lamport_workspace_bytes(world, max_numel) bytes, mapped into every rank
 *     (torch symmetric memory), initialised ONCE with b200_lamport_workspace_init followed by a cross-rank barrier;
 *     peer_workspaces[q] = rank q's region as addressable from this rank.  The call counter lives in the workspace
 *     (device memory), so captured CUDA graphs replay correctly.  All ranks must issue the same sequence of calls.
 *   input: this rank's [rows][cols] partial sums in ordinary local memory (cols % (8 * world_size) == 0);
 *   out:   the sum (fp32 accumulation in rank order, rounded once - same values as b200_peer_all_reduce_out), local.
 *   _rms_norm_quant: residual += sum; result, scales = per-token FP8 quant of RMSNorm(residual) * weight, bit-identical
 *     to the all-reduce followed by b200_rms_norm_dynamic_per_token_quant.
 * rank == -1 emulates all ranks on one GPU (tests), as for the entries above; input / out / residual / result then
 * hold world_size copies rows * cols elements apart (scales: rows apart). */
/*   _rms_norm_nvfp4_quant: the same with the NVFP4 quantiser (swizzled scales, padding written), bit-identical to the
 *     all-reduce followed by b200_rms_norm_nvfp4_quant. */
int64_t b200_lamport_workspace_bytes(int world_size, int64_t max_numel);
int b200_lamport_workspace_init(void* workspace, int64_t workspace_bytes, b200_stream_t stream);
int b200_lamport_all_reduce(void* out, const void* input, const uint64_t* peer_workspaces, int64_t workspace_bytes,
                            int rank, int world_size, int64_t rows, int cols, int dtype, b200_stream_t stream);
int b200_lamport_all_reduce_rms_norm_quant(void* result, float* scales, const void* input,
                                           const uint64_t* peer_workspaces, int64_t workspace_bytes, int rank,
                                           int world_size, const void* weight, void* residual, float epsilon,
                                           int64_t num_tokens, int hidden, int dtype, b200_stream_t stream);

int b200_lamport_all_reduce_rms_norm_nvfp4_quant(void* result, void* result_block_scale, const void* input,
                                                 const uint64_t* peer_workspaces, int64_t workspace_bytes, int rank,
                                                 int world_size, const void* weight, void* residual,
                                                 const float* input_global_scale, float epsilon, int64_t num_tokens,
                                                 int hidden, int dtype, b200_stream_t stream);

/* Spin budget of one cross-rank barrier wait, in polls (default 2^28, minutes); 0 = wait for ever.  A rank
 * that exceeds it prints a message and traps instead of hanging the GPU. */
void b200_peer_set_barrier_timeout(unsigned long long max_polls);

/* -------------------------------------------------------------------------
 * Programmatic dependent launch (no reference counterpart).  On by default (B200_PDL=0 in the
 * environment or b200_set_pdl(0) turns it off): every kernel is launched so that the next kernel's
 * prologue overlaps this kernel's tail, and every kernel waits for its predecessors' memory before
 * touching global data.  Results are identical either way.
 * ---------------------------------------------------------------------- */
void b200_set_pdl(int enabled);
int b200_get_pdl(void);

/* -------------------------------------------------------------------------
 * Debug aid (no reference counterpart): when set to a device buffer of
 * [grid][16] uint64, the decode-regime scaled_mm kernel stamps %globaltimer at its
 * phase boundaries (scripts/prof_gemm.py --trace).  Pass NULL to switch it off.
 * ---------------------------------------------------------------------- */
void b200_debug_gemm_trace(unsigned long long* device_buffer);
/* Same for the fp8 paged-attention kernel: [cta][32] stamps of one steady-state tile and per-barrier wait
 * cycles (scripts/prof_attn.py). */
void b200_debug_attn_trace(unsigned long long* device_buffer);
/* fp8-cache decode kernel selection (tests / tuning): 1 = tcgen05 kernel, 0 = register-level kernel,
 * -1 = default (register-level unless B200_ATTN_FP8_TCGEN05=1). */
void b200_debug_attn_fp8_tcgen05(int mode);
/* Force the decode-regime (M <= 256) kernel / work partition (tests / tuning): 0 whole tiles per CTA pair, 3 cluster
 * split-K over distributed shared memory, 4 the persistent 2-CTA kernel (UMMA M = 128 / 256 across the pair, no K
 * split), -1 automatic. */
void b200_debug_gemm_mode(int mode);
/* Debug aid: when set, the 2-CTA FP8 kernel also writes the raw TMEM image of its LAST tile, fp32
 * [2 CTAs][128 lanes][accumulator columns], there (how the UMMA M = 128 data-path layout was verified). */
void b200_debug_gemm_dump(float* device_buffer);

#ifdef __cplusplus
}
#endif
#endif /* B200_VLLM_H_ */
