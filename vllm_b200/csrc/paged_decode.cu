// Paged-attention decode for sm_100a: split-KV + combine, GQA-grouped.
//
// Reference semantics: legacy paged_attention_v1/v2 (vllm/_custom_ops.py:120-163,
// docs/design/paged_attention.md, ROCm port csrc/rocm/attention.cu:320-1520) and the
// in-tree successor kernel_unified_attention 3D + reduce_segments
// (vllm/v1/attention/ops/triton_unified_attention.py:179-681,685-771).
//
// B200 design (HBM-bound):
//  * one CTA per (split, kv-head[, q-head group of 16], sequence); the G = Hq/Hkv
//    query heads of a kv-head share every K/V byte, so KV is read once (the
//    legacy kernel re-read it G times);
//  * K/V pages are gathered with TMA (cp.async.bulk.tensor.4d over the
//    [D, H, block_size, num_blocks] view of the cache with its true strides, so
//    NHD / HND / K|V-interleaved layouts all work) into 128B-swizzled shared
//    memory; each of the 4 warps owns a private ring of slots and mbarriers and
//    re-arms a slot as soon as it has consumed it -> no block-wide barrier in the
//    main loop, 12-16 page-chunks in flight per CTA, 2-3 CTAs per SM;
//  * QK^T and PV for the (<=16) grouped heads run on mma.sync m16n8k16 with fp32
//    accumulation: the kernel is bandwidth bound, the tensor pipe is mostly idle;
//  * online softmax in the exp2 domain, quad shuffles for row max / row sum,
//    warps merged through shared memory at the end, split partials merged by a
//    second tiny kernel.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "paged_decode.cuh"

namespace b200 {

template <int D, int KV> struct DecodeCfg {
  static constexpr int ES = (KV == B200_KV_AUTO) ? 2 : 1;  // bytes per cache element
  static constexpr int ROWB = D * ES;                      // bytes per (token, head) row
  static_assert(ROWB % 128 == 0, "row must be a whole number of 128-byte swizzle spans");
  static constexpr int NSUB = ROWB / 128;           // TMA boxes per 16-token K (or V) piece
  static constexpr int SUB_BYTES = kChunk * 128;    // 2 KB, 1024-aligned
  // fp8 caches: a ring slot holds 32 tokens (two 16-token pieces).  The kernel is issue-bound there (64 converts
  // per 16 tokens) and ~150 of its ~450 instructions per 16 tokens were per-slot overhead (barrier wait, TMA
  // issue, masks, max/rescale bookkeeping) - twice the tokens per slot halves that share.
  static constexpr int TOK = (KV == B200_KV_AUTO) ? 16 : 32;
  static constexpr int NBOX = TOK / kChunk;
  static constexpr int BOX_BYTES = NSUB * SUB_BYTES;  // one 16-token piece of K (or V)
  static constexpr int KCH_BYTES = NBOX * BOX_BYTES;
  static constexpr int SLOT_BYTES = 2 * KCH_BYTES;  // K part + V part
#ifndef B200_FP8_NSLOT
#define B200_FP8_NSLOT 2
#endif
  static constexpr int NSLOT = (KV != B200_KV_AUTO) ? (ROWB >= 256 ? 2 : B200_FP8_NSLOT) : (ROWB >= 512 ? 2 : (ROWB == 256 ? 3 : 4));
  static constexpr int RING_BYTES = kWarps * NSLOT * SLOT_BYTES;
  static constexpr int OPAD = D + 8;  // fp32 row pitch of the merge buffer
  static constexpr int EPI_BYTES = kWarps * 16 * OPAD * 4 + 2 * kWarps * 16 * 4;
  static constexpr int BAR_BYTES = kWarps * NSLOT * 8;
  static constexpr int SMEM_BYTES = (RING_BYTES > EPI_BYTES ? RING_BYTES : EPI_BYTES) + BAR_BYTES + 1024;
  static constexpr int BOX_D = 128 / ES;  // elements per 128-byte span
};

__device__ __forceinline__ uint32_t pack2(float a, float b, __nv_bfloat16) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pack2(float a, float b, __half) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// two fp8 bytes (low 16 bits of x) -> packed f16x2, exact
template <int KV> __device__ __forceinline__ uint32_t fp8x2_to_h2(uint32_t x) {
  uint32_t r;
  if constexpr (KV == B200_KV_FP8_E5M2) {
    asm("{\n\t.reg .b16 lo;\n\tcvt.u16.u32 lo, %1;\n\tcvt.rn.f16x2.e5m2x2 %0, lo;\n\t}" : "=r"(r) : "r"(x));
  } else {
    asm("{\n\t.reg .b16 lo;\n\tcvt.u16.u32 lo, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, lo;\n\t}" : "=r"(r) : "r"(x));
  }
  return r;
}
// four fp8 bytes -> f16x2 of bytes (0,1) and f16x2 of bytes (2,3); the halves are sub-registers, no shift
template <int KV> __device__ __forceinline__ void fp8x4_to_h2x2(uint32_t x, uint32_t& lo, uint32_t& hi) {
  if constexpr (KV == B200_KV_FP8_E5M2) {
    asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %2;\n\tcvt.rn.f16x2.e5m2x2 %0, l;\n\tcvt.rn.f16x2.e5m2x2 %1, h;\n\t}"
        : "=r"(lo), "=r"(hi)
        : "r"(x));
  } else {
    asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %2;\n\tcvt.rn.f16x2.e4m3x2 %0, l;\n\tcvt.rn.f16x2.e4m3x2 %1, h;\n\t}"
        : "=r"(lo), "=r"(hi)
        : "r"(x));
  }
}
template <typename QT> __device__ __forceinline__ uint32_t load_q_pair_as_half(const QT* p) {
  // fp8-KV path runs the MMAs in f16 (e4m3/e5m2 are exactly representable);
  // bf16 queries are converted with saturation to the f16 range.
  if constexpr (std::is_same<QT, __half>::value) {
    return *reinterpret_cast<const uint32_t*>(p);
  } else {
    __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(p);
    float2 f = __bfloat1622float2(v);
    f.x = fminf(fmaxf(f.x, -65504.f), 65504.f);
    f.y = fminf(fmaxf(f.y, -65504.f), 65504.f);
    __half2 h = __floats2half2_rn(f.x, f.y);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}

template <int SMEM> struct OccFor {
  static constexpr int value = SMEM <= 74 * 1024 ? 3 : (SMEM <= 112 * 1024 ? 2 : 1);
};

// R8: every CTA has at most 8 live query heads (GQA group <= 8): rows g+8 of the MMA tiles are padding and
// their share of the softmax is skipped.
template <typename QT, int D, int KV, bool R8>
__global__ void __launch_bounds__(kWarps * 32, OccFor<DecodeCfg<D, KV>::SMEM_BYTES>::value)
    paged_decode_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v,
                        const DecodeArgs a) {
  using Cfg = DecodeCfg<D, KV>;
  constexpr bool kFp8 = (KV != B200_KV_AUTO);
  using MT = typename std::conditional<kFp8, __half, QT>::type;  // MMA operand type
  constexpr bool kBf16 = std::is_same<MT, __nv_bfloat16>::value;
  constexpr int NSLOT = Cfg::NSLOT;
  constexpr int SUB = Cfg::SUB_BYTES;
  constexpr int KS = D / 16;  // k-steps of QK^T == 16-wide d groups of PV
  constexpr int TOK = Cfg::TOK, NBOX = Cfg::NBOX, NT = TOK / 8;

  pdl_trigger();
  pdl_wait();  // query, caches and seq_lens come from earlier kernels

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  constexpr int kDataBytes = Cfg::RING_BYTES > Cfg::EPI_BYTES ? Cfg::RING_BYTES : Cfg::EPI_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kDataBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int split = blockIdx.x, seq = blockIdx.z;
  const int kvh = blockIdx.y / a.qgroups, qg = blockIdx.y - kvh * a.qgroups;
  const int rows = min(16, a.group - qg * 16);  // live query heads of this CTA
  const int qh0 = kvh * a.group + qg * 16;
  const int len = a.seq_lens[seq];
  const int lo = (a.sliding_window > 0) ? max(0, len - a.sliding_window) : 0;
  const int c_begin = lo / TOK, c_end = (len + TOK - 1) / TOK;
  const int nc = max(0, c_end - c_begin);
  int cps = (nc + a.num_splits - 1) / a.num_splits;
  cps = (cps + kWarps - 1) / kWarps * kWarps;
  const int c0 = c_begin + split * cps;
  const int c1 = min(c_end, c0 + cps);

  if (c0 >= c1) {  // nothing to attend to in this split
    if (a.num_splits > 1) {
      if ((int)threadIdx.x < rows)
        a.part_ml[((int64_t)seq * a.num_q_heads + qh0 + threadIdx.x) * a.num_splits + split] =
            make_float2(-INFINITY, 0.f);
    } else {
      if (a.out_scale) {  // e4m3 zero is the zero byte
        uint8_t* o = reinterpret_cast<uint8_t*>(a.out) + (int64_t)seq * a.out_stride_seq + (int64_t)qh0 * a.out_stride_head;
        for (int i = threadIdx.x; i < rows * D; i += blockDim.x) o[(int64_t)(i / D) * a.out_stride_head + (i % D)] = 0;
      } else {
        QT* o = reinterpret_cast<QT*>(a.out) + (int64_t)seq * a.out_stride_seq + (int64_t)qh0 * a.out_stride_head;
        for (int i = threadIdx.x; i < rows * D; i += blockDim.x)
          o[(int64_t)(i / D) * a.out_stride_head + (i % D)] = from_float<QT>(0.f);
      }
      if (a.out_lse && (int)threadIdx.x < rows)
        a.out_lse[(int64_t)seq * a.num_q_heads + qh0 + threadIdx.x] = -INFINITY;
    }
    return;
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWarps * NSLOT; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
    prefetch_tensormap(&tmap_k);
    prefetch_tensormap(&tmap_v);
  }
  __syncthreads();

  // ---- Q fragments (A operand, rows = query heads of the group) -------------
  uint32_t qa[KS][4];
  {
    const QT* qb = reinterpret_cast<const QT*>(a.query) + (int64_t)seq * a.q_stride_seq +
                   (int64_t)qh0 * a.q_stride_head;
    const QT* q_lo = qb + (int64_t)g * a.q_stride_head;
    const QT* q_hi = qb + (int64_t)(g + 8) * a.q_stride_head;
    const bool v_lo = g < rows, v_hi = (g + 8) < rows;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      // fp8 path: MMA k-slots (2t,2t+1 | 2t+8,2t+9) map to d = 16ks+4t+(0,1 | 2,3)
      const int ca = kFp8 ? 16 * ks + 4 * t : 16 * ks + 2 * t;
      const int cb = kFp8 ? ca + 2 : ca + 8;
      if constexpr (kFp8) {
        qa[ks][0] = v_lo ? load_q_pair_as_half(q_lo + ca) : 0u;
        qa[ks][1] = v_hi ? load_q_pair_as_half(q_hi + ca) : 0u;
        qa[ks][2] = v_lo ? load_q_pair_as_half(q_lo + cb) : 0u;
        qa[ks][3] = v_hi ? load_q_pair_as_half(q_hi + cb) : 0u;
      } else {
        qa[ks][0] = v_lo ? *reinterpret_cast<const uint32_t*>(q_lo + ca) : 0u;
        qa[ks][1] = v_hi ? *reinterpret_cast<const uint32_t*>(q_hi + ca) : 0u;
        qa[ks][2] = v_lo ? *reinterpret_cast<const uint32_t*>(q_lo + cb) : 0u;
        qa[ks][3] = v_hi ? *reinterpret_cast<const uint32_t*>(q_hi + cb) : 0u;
      }
    }
  }

  const float kscale = (kFp8 && a.k_scale) ? *a.k_scale : 1.f;
  const float sc_nat = a.scale * kscale;  // natural-domain logit scale
  const float sc = sc_nat * kLog2e;       // exp2-domain logit scale
  const bool use_cap = a.soft_cap > 0.f;
  const float cap_inv = use_cap ? 1.f / a.soft_cap : 0.f;
  const float cap_l2 = a.soft_cap * kLog2e;

  // ---- per-warp chunk list, block ids prefetched 32 at a time -----------------
  const int span = c1 - c0;
  const int n_my = span > warp ? (span - warp + kWarps - 1) / kWarps : 0;
  uint8_t* ring = smem + warp * NSLOT * Cfg::SLOT_BYTES;
  uint64_t* wbar = bars + warp * NSLOT;
  const int32_t* bt_row =
      a.block_table + (int64_t)(a.token_to_seq ? a.token_to_seq[seq] : seq) * a.block_table_stride;
  // block id of 16-token piece h of this lane's slot in `batch` (0 where the piece starts past the sequence:
  // such a piece is not loaded, and the table may not extend that far)
  // block_size is a power of two in practice (16..128): shift/mask instead of the ~20-instruction integer
  // division, which sat in the per-slot TMA issue path of every warp
  const int bs_shift = (a.block_size & (a.block_size - 1)) == 0 ? 31 - __clz(a.block_size) : -1;
  auto div_bs = [&](int tok) { return bs_shift >= 0 ? (tok >> bs_shift) : tok / a.block_size; };
  auto mod_bs = [&](int tok) { return bs_shift >= 0 ? (tok & (a.block_size - 1)) : tok % a.block_size; };
  auto load_id = [&](int batch, int h) -> int {
    const int i = batch * 32 + lane;
    if (i < n_my) {
      const int tok = (c0 + warp + kWarps * i) * TOK + h * kChunk;
      if (tok < len) return bt_row[div_bs(tok)];
    }
    return 0;
  };
  int ids_cur[NBOX], ids_next[NBOX], issue_batch = 0;
#pragma unroll
  for (int h = 0; h < NBOX; ++h) {
    ids_cur[h] = load_id(0, h);
    ids_next[h] = load_id(1, h);
  }
  auto issue = [&](int ii, int slot) {
    const int b = ii >> 5;
    if (b != issue_batch) {
#pragma unroll
      for (int h = 0; h < NBOX; ++h) {
        ids_cur[h] = ids_next[h];
        ids_next[h] = load_id(b + 1, h);
      }
      issue_batch = b;
    }
    int blk[NBOX];
#pragma unroll
    for (int h = 0; h < NBOX; ++h) blk[h] = __shfl_sync(0xffffffffu, ids_cur[h], ii & 31);
    if (lane == 0) {
      const int tok0 = (c0 + warp + kWarps * ii) * TOK;
      uint8_t* dst = ring + slot * Cfg::SLOT_BYTES;
      int nload = 0;
#pragma unroll
      for (int h = 0; h < NBOX; ++h) nload += (tok0 + h * kChunk < len);
      mbar_arrive_expect_tx(&wbar[slot], (uint32_t)nload * 2u * Cfg::BOX_BYTES);
#pragma unroll
      for (int h = 0; h < NBOX; ++h) {
        if (tok0 + h * kChunk >= len) continue;
        const int tok_in_blk = mod_bs(tok0 + h * kChunk);
#pragma unroll
        for (int sb = 0; sb < Cfg::NSUB; ++sb)
          tma_load_4d(dst + h * Cfg::BOX_BYTES + sb * SUB, &tmap_k, &wbar[slot], sb * Cfg::BOX_D, kvh, tok_in_blk,
                      blk[h], kEvictFirst);
#pragma unroll
        for (int sb = 0; sb < Cfg::NSUB; ++sb)
          tma_load_4d(dst + Cfg::KCH_BYTES + h * Cfg::BOX_BYTES + sb * SUB, &tmap_v, &wbar[slot], sb * Cfg::BOX_D,
                      kvh, tok_in_blk, blk[h], kEvictFirst);
      }
    }
  };
  {
    const int pre = min(NSLOT, n_my);
    for (int ii = 0; ii < pre; ++ii) issue(ii, ii);
  }

  // ---- online-softmax state -------------------------------------------------
  float o[2 * KS][4];
#pragma unroll
  for (int n = 0; n < 2 * KS; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;  // rows g and g+8

  // per-lane ldmatrix address pieces
  const int li = lane >> 3, lr = lane & 7;
  // 16-bit KV:  K matrices (tok-lo,d-lo)(tok-lo,d-hi)(tok-hi,d-lo)(tok-hi,d-hi); V.trans (tok-lo,d)(tok-hi,d)(tok-lo,d+1)(tok-hi,d+1)
  // fp8 KV:     K and V both (tok-lo,c)(tok-hi,c)(tok-lo,c+1)(tok-hi,c+1) with c = 16-byte column
  // fp8: S column c of n-tile nt holds token 4*(c/2) + 2*nt + (c&1), so that a thread's four S values of a
  // 16-token piece are the tokens 4t..4t+3 - the order in which the transposed byte ldmatrix of V delivers them
  const int k_tok = kFp8 ? 4 * (lr >> 1) + 2 * (li & 1) + (lr & 1) : ((li >> 1) << 3) + lr;
  const int k_cof = kFp8 ? (li >> 1) : (li & 1);
  const int v_tok = ((li & 1) << 3) + lr;
  const int v_cof = li >> 1;

  int slot = 0;
  uint32_t phase = 0;
  for (int i = 0; i < n_my; ++i) {
    const int c = c0 + warp + kWarps * i;
    const uint32_t kbase = smem_u32(ring + slot * Cfg::SLOT_BYTES);
    const uint32_t vbase = kbase + Cfg::KCH_BYTES;
    mbar_wait(&wbar[slot], phase);

    const int tok0 = c * TOK;
    if (tok0 + TOK > len) {
      // boundary slot: rows past seq_len hold stale / never-loaded bytes; p is 0 there but 0 * NaN would
      // poison the accumulator, so clear those V rows (K garbage only reaches masked logits).
      const int valid = len - tok0;
      uint8_t* vs = ring + slot * Cfg::SLOT_BYTES + Cfg::KCH_BYTES;
      for (int idx = lane; idx < (TOK - valid) * Cfg::NSUB * 8; idx += 32) {
        const int row = valid + idx / (Cfg::NSUB * 8);
        const int rem = idx % (Cfg::NSUB * 8);
        *reinterpret_cast<uint4*>(vs + (row >> 4) * Cfg::BOX_BYTES + (rem >> 3) * SUB + (row & 15) * 128 +
                                  (rem & 7) * 16) = make_uint4(0, 0, 0, 0);
      }
      __syncwarp();
    }

    // ---- S = Q K^T (16 x 16 tokens) ---------------------------------------
    float s[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
    if constexpr (!kFp8) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const int dch = 2 * ks + k_cof;
        uint32_t b[4];
        ldmatrix_x4(b, kbase + (dch >> 3) * SUB + k_tok * 128 + (((dch & 7) ^ lr) << 4));
        mma_16816<kBf16>(s[0], qa[ks], b[0], b[1]);
        mma_16816<kBf16>(s[1], qa[ks], b[2], b[3]);
      }
    } else {
#pragma unroll
      for (int hb = 0; hb < NBOX; ++hb) {  // 16-token pieces of the slot
#pragma unroll
        for (int jj = 0; jj < KS / 2; ++jj) {
          const int dch = 2 * jj + k_cof;  // 16-byte column = one k-step of 16 fp8
          uint32_t b[4];
          ldmatrix_x4(b, kbase + hb * Cfg::BOX_BYTES + (dch >> 3) * SUB + k_tok * 128 + (((dch & 7) ^ (k_tok & 7)) << 4));
          uint32_t c[4][2];
#pragma unroll
          for (int m = 0; m < 4; ++m) fp8x4_to_h2x2<KV>(b[m], c[m][0], c[m][1]);
          mma_16816<false>(s[2 * hb], qa[2 * jj], c[0][0], c[0][1]);
          mma_16816<false>(s[2 * hb + 1], qa[2 * jj], c[1][0], c[1][1]);
          mma_16816<false>(s[2 * hb], qa[2 * jj + 1], c[2][0], c[2][1]);
          mma_16816<false>(s[2 * hb + 1], qa[2 * jj + 1], c[3][0], c[3][1]);
        }
      }
    }

    // ---- scale, soft-cap, mask ------------------------------------------------
    const bool need_mask = (tok0 + TOK > len) || (tok0 < lo);
    constexpr int NE = R8 ? 2 : 4;  // accumulator elements per n-tile that belong to live rows
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < NE; ++e) {
        float v = use_cap ? cap_l2 * fast_tanh(s[nt][e] * sc_nat * cap_inv) : s[nt][e] * sc;
        if (need_mask) {
          const int tok = kFp8 ? tok0 + 16 * (nt >> 1) + 4 * t + 2 * (nt & 1) + (e & 1) : tok0 + nt * 8 + 2 * t + (e & 1);
          if (tok >= len || tok < lo) v = -INFINITY;
        }
        s[nt][e] = v;
      }
    float mx0 = fmaxf(s[0][0], s[0][1]), mx1 = R8 ? 0.f : fmaxf(s[0][2], s[0][3]);
#pragma unroll
    for (int nt = 1; nt < NT; ++nt) {
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      if constexpr (!R8) mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    if constexpr (!R8) {
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    }
    const float mn0 = fmaxf(m0, mx0), mn1 = R8 ? m1 : fmaxf(m1, mx1);
    const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
    const float al0 = fast_exp2(m0 - ms0), al1 = R8 ? 1.f : fast_exp2(m1 - ms1);
    m0 = mn0;
    m1 = mn1;
    float p[NT][4];
    float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      p[nt][0] = fast_exp2(s[nt][0] - ms0);
      p[nt][1] = fast_exp2(s[nt][1] - ms0);
      ps0 += p[nt][0] + p[nt][1];
      if constexpr (R8) {
        p[nt][2] = p[nt][3] = 0.f;
      } else {
        p[nt][2] = fast_exp2(s[nt][2] - ms1);
        p[nt][3] = fast_exp2(s[nt][3] - ms1);
        ps1 += p[nt][2] + p[nt][3];
      }
    }
    l0 = l0 * al0 + ps0;
    if constexpr (!R8) l1 = l1 * al1 + ps1;
    if (__any_sync(0xffffffffu, (al0 != 1.f) || (al1 != 1.f))) {
#pragma unroll
      for (int n = 0; n < 2 * KS; ++n) {
        o[n][0] *= al0;
        o[n][1] *= al0;
        if constexpr (!R8) {
          o[n][2] *= al1;
          o[n][3] *= al1;
        }
      }
    }

    // ---- O += P V -----------------------------------------------------------
    auto pack_p = [&](int hb, uint32_t (&pa)[4]) {  // A fragment of the 16 tokens of piece hb
      pa[0] = pack2(p[2 * hb][0], p[2 * hb][1], MT());
      pa[2] = pack2(p[2 * hb + 1][0], p[2 * hb + 1][1], MT());
      if constexpr (R8) {
        pa[1] = pa[3] = 0u;
      } else {
        pa[1] = pack2(p[2 * hb][2], p[2 * hb][3], MT());
        pa[3] = pack2(p[2 * hb + 1][2], p[2 * hb + 1][3], MT());
      }
    };
    if constexpr (!kFp8) {
      uint32_t pa[4];
      pack_p(0, pa);
#pragma unroll
      for (int j = 0; j < KS; ++j) {
        const int dch = 2 * j + v_cof;
        uint32_t b[4];
        ldmatrix_x4_trans(b, vbase + (dch >> 3) * SUB + v_tok * 128 + (((dch & 7) ^ lr) << 4));
        mma_16816<kBf16>(o[2 * j], pa, b[0], b[1]);
        mma_16816<kBf16>(o[2 * j + 1], pa, b[2], b[3]);
      }
    } else {
      // ldmatrix.m16n16.trans.b8 hands every thread four consecutive tokens of one d column per register: two
      // converts make the B fragment of one n-tile, no byte permutes (the k index of the MMA runs over the
      // tokens in the order 4t, 4t+1 | 4t+2, 4t+3, which is how the S accumulators hold them, see k_tok).
      const int vrow = lane & 15, vmat = lane >> 4;
#pragma unroll
      for (int hb = 0; hb < NBOX; ++hb) {
        uint32_t pa[4];
        pack_p(hb, pa);
#pragma unroll
        for (int jj = 0; jj < D / 32; ++jj) {  // 32 d columns = n-tiles 4jj .. 4jj+3
          const int dch = 2 * jj + vmat;       // 16-byte column
          uint32_t b[4];
          ldmatrix_b8_x2_trans(b, vbase + hb * Cfg::BOX_BYTES + (dch >> 3) * SUB + vrow * 128 + (((dch & 7) ^ (vrow & 7)) << 4));
#pragma unroll
          for (int qd = 0; qd < 4; ++qd) {
            uint32_t c0, c1;
            fp8x4_to_h2x2<KV>(b[qd], c0, c1);
            mma_16816<false>(o[4 * jj + qd], pa, c0, c1);
          }
        }
      }
    }

    __syncwarp();
    if (i + NSLOT < n_my) issue(i + NSLOT, slot);
    if (++slot == NSLOT) {
      slot = 0;
      phase ^= 1;
    }
  }

  // ---- merge the 4 warps through shared memory --------------------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  __syncthreads();  // every issued TMA has been consumed: the rings are free
  constexpr int OPAD = Cfg::OPAD;
  float* sm_o = reinterpret_cast<float*>(smem);  // [kWarps][16][OPAD]
  float* sm_m = sm_o + kWarps * 16 * OPAD;       // [kWarps][16]
  float* sm_l = sm_m + kWarps * 16;
  if (t == 0) {
    sm_m[warp * 16 + g] = m0;
    sm_m[warp * 16 + g + 8] = m1;
    sm_l[warp * 16 + g] = l0;
    sm_l[warp * 16 + g + 8] = l1;
  }
  {
    float* r_lo = sm_o + (warp * 16 + g) * OPAD;
    float* r_hi = sm_o + (warp * 16 + g + 8) * OPAD;
#pragma unroll
    for (int n = 0; n < 2 * KS; ++n) {
      const int d0 = n * 8 + 2 * t, d1 = d0 + 1;
      r_lo[d0] = o[n][0];
      r_lo[d1] = o[n][1];
      if (rows > 8) {
        r_hi[d0] = o[n][2];
        r_hi[d1] = o[n][3];
      }
    }
  }
  __syncthreads();
  const float vscale = (kFp8 && a.v_scale) ? *a.v_scale : 1.f;
  for (int idx = threadIdx.x; idx < rows * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;
    float mt = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) mt = fmaxf(mt, sm_m[w * 16 + row]);
    float acc = 0.f, lt = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float mw = sm_m[w * 16 + row];
      const float wgt = (mw == -INFINITY) ? 0.f : fast_exp2(mw - mt);
      acc += sm_o[(w * 16 + row) * OPAD + d] * wgt;
      lt += sm_l[w * 16 + row] * wgt;
    }
    const int qh = qh0 + row;
    if (a.num_splits > 1) {
      const int64_t pi = ((int64_t)seq * a.num_q_heads + qh) * a.num_splits + split;
      a.part_o[pi * D + d] = acc;
      if (d == 0) a.part_ml[pi] = make_float2(mt, lt);
    } else {
      const float ov = lt > 0.f ? acc / lt * vscale : 0.f;
      if (a.out_scale) {
        reinterpret_cast<uint8_t*>(a.out)[(int64_t)seq * a.out_stride_seq + (int64_t)qh * a.out_stride_head + d] =
            to_e4m3_sat(ov * (1.0f / *a.out_scale));
      } else {
        QT* op = reinterpret_cast<QT*>(a.out) + (int64_t)seq * a.out_stride_seq + (int64_t)qh * a.out_stride_head;
        op[d] = from_float<QT>(ov);
      }
      if (d == 0 && a.out_lse)
        a.out_lse[(int64_t)seq * a.num_q_heads + qh] = lt > 0.f ? (mt + log2f(lt)) * kLn2 : -INFINITY;
    }
  }
}

// ---- split combine:  m = max m_i;  o = sum o_i e^{m_i-m} / sum l_i e^{m_i-m} ------
// (reduce_segments math, triton_unified_attention.py:735-759; partials hold
// un-normalised accumulators like the Triton kernel, not the legacy v2 ones.)
template <typename QT>
__global__ void __launch_bounds__(128) paged_decode_combine_kernel(
    QT* __restrict__ out, float* __restrict__ out_lse, const float* __restrict__ part_o,
    const float2* __restrict__ part_ml, const float* __restrict__ v_scale, const float* __restrict__ out_scale,
    int num_q_heads, int num_splits, int D, int64_t out_stride_seq, int64_t out_stride_head) {
  pdl_trigger();
  pdl_wait();
  // one warp per (sequence, head), four heads per CTA; a lane owns 4 consecutive d (16-byte partial loads,
  // all splits' loads independent of each other)
  const int h = blockIdx.x * 4 + (threadIdx.x >> 5), seq = blockIdx.y, lane = threadIdx.x & 31;
  if (h >= num_q_heads) return;
  const int64_t base = ((int64_t)seq * num_q_heads + h) * num_splits;
  float mt = -INFINITY;
  for (int i = lane; i < num_splits; i += 32) mt = fmaxf(mt, part_ml[base + i].x);
  mt = warp_max(mt);
  float lt = 0.f;
  for (int i = lane; i < num_splits; i += 32) {
    const float2 ml = part_ml[base + i];
    lt += (ml.x == -INFINITY) ? 0.f : ml.y * fast_exp2(ml.x - mt);
  }
  lt = warp_sum(lt);
  const float vs = v_scale ? *v_scale : 1.f;
  const float inv = lt > 0.f ? vs / lt : 0.f;
  if (out_scale) {
    // fused static FP8 output quant: e4m3(o * (1 / output_scale)) from the fp32 accumulator, 4 bytes per lane
    uint8_t* ob = reinterpret_cast<uint8_t*>(out) + (int64_t)seq * out_stride_seq + (int64_t)h * out_stride_head;
    const float qs = inv * (1.0f / *out_scale);
    for (int d = lane * 4; d < D; d += 128) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int i = 0; i < num_splits; ++i) {
        const float mi = part_ml[base + i].x;
        if (mi == -INFINITY) continue;
        const float wgt = fast_exp2(mi - mt);
        for (int e = 0; e < 4 && d + e < D; ++e) acc[e] += part_o[(base + i) * D + d + e] * wgt;
      }
      for (int e = 0; e < 4 && d + e < D; ++e) ob[d + e] = to_e4m3_sat(acc[e] * qs);
    }
    if (lane == 0 && out_lse) out_lse[(int64_t)seq * num_q_heads + h] = lt > 0.f ? (mt + log2f(lt)) * kLn2 : -INFINITY;
    return;
  }
  QT* op = out + (int64_t)seq * out_stride_seq + (int64_t)h * out_stride_head;
  const bool vec = (D % 4 == 0) && ((reinterpret_cast<uintptr_t>(op) & 7) == 0);
  if (vec) {
    for (int d = lane * 4; d < D; d += 128) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int i = 0; i < num_splits; ++i) {
        const float mi = part_ml[base + i].x;
        if (mi == -INFINITY) continue;
        const float wgt = fast_exp2(mi - mt);
        const float4 p = *reinterpret_cast<const float4*>(part_o + (base + i) * D + d);
        acc.x += p.x * wgt, acc.y += p.y * wgt, acc.z += p.z * wgt, acc.w += p.w * wgt;
      }
      QT r[4] = {from_float<QT>(acc.x * inv), from_float<QT>(acc.y * inv), from_float<QT>(acc.z * inv),
                 from_float<QT>(acc.w * inv)};
      *reinterpret_cast<uint2*>(op + d) = *reinterpret_cast<const uint2*>(r);
    }
  } else {
    for (int d = lane; d < D; d += 32) {
      float acc = 0.f;
      for (int i = 0; i < num_splits; ++i) {
        const float mi = part_ml[base + i].x;
        if (mi != -INFINITY) acc += part_o[(base + i) * D + d] * fast_exp2(mi - mt);
      }
      op[d] = from_float<QT>(acc * inv);
    }
  }
  if (lane == 0 && out_lse)
    out_lse[(int64_t)seq * num_q_heads + h] = lt > 0.f ? (mt + log2f(lt)) * kLn2 : -INFINITY;
}

// ---- host side ------------------------------------------------------------------
struct TmapKey {
  const void* base;
  int64_t s0, s1, s2, nb;
  int d, h, bs, es;
  bool operator==(const TmapKey& o) const {
    return base == o.base && s0 == o.s0 && s1 == o.s1 && s2 == o.s2 && nb == o.nb && d == o.d && h == o.h &&
           bs == o.bs && es == o.es;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t x = reinterpret_cast<size_t>(k.base);
    auto mix = [&](uint64_t v) { x ^= v + 0x9e3779b97f4a7c15ull + (x << 6) + (x >> 2); };
    mix(k.s0); mix(k.s1); mix(k.s2); mix(k.nb); mix(k.d); mix(k.h); mix(k.bs); mix(k.es);
    return x;
  }
};

// [D, H, block_size, num_blocks] view of a paged cache, box = one 128-byte span x 16 tokens.
static int get_cache_tmap(CUtensorMap* out, const void* base, int D, int H, int block_size, int64_t num_blocks,
                          int64_t head_stride, int64_t page_stride, int64_t block_stride, int es) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  TmapKey key{base, head_stride, page_stride, block_stride, num_blocks, D, H, block_size, es};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return 0;
  }
  uint64_t dims[4] = {(uint64_t)D, (uint64_t)H, (uint64_t)block_size, (uint64_t)num_blocks};
  uint64_t strides[3] = {(uint64_t)head_stride * es, (uint64_t)page_stride * es, (uint64_t)block_stride * es};
  uint32_t box[4] = {(uint32_t)(128 / es), 1, (uint32_t)kChunk, 1};
  CUtensorMap tm;
  int rc = encode_tmap(&tm, es == 2 ? kTmaU16 : kTmaU8, 4, base, dims, strides, box, kSwizzle128);
  if (rc) return rc;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, tm);
  *out = tm;
  return 0;
}

static int g_fp8_tcgen05 = -1;  // b200_debug_attn_fp8_tcgen05: -1 environment, 0 register kernel, 1 tcgen05 kernel

template <typename QT, int D, int KV>
static int launch_decode(const CUtensorMap& tk, const CUtensorMap& tv, const DecodeArgs& a, int num_seqs,
                         cudaStream_t stream) {
  using Cfg = DecodeCfg<D, KV>;
  const bool r8 = a.group <= 8;
  auto kern = r8 ? paged_decode_kernel<QT, D, KV, true> : paged_decode_kernel<QT, D, KV, false>;
  if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc_attr;
  // fp8 caches: the register-level kernel below is the default.  The tcgen05 kernel (paged_decode_fp8.cu)
  // is parity-green but measured 5-15% slower on the BASELINE shapes (profiles/r01_attn_fp8_tcgen05.md:
  // bound by the TMA request rate for 128-byte cache rows and by the QK->softmax->PV latency chain);
  // B200_ATTN_FP8_TCGEN05=1 selects it.
  static const bool env_legacy = [] {
    const char* e = getenv("B200_ATTN_FP8_TCGEN05");
    return !(e && *e && atoi(e) != 0);
  }();
  const bool legacy_fp8 = a.out_scale ? true : (g_fp8_tcgen05 >= 0 ? (g_fp8_tcgen05 == 0) : env_legacy);
  if (KV != B200_KV_AUTO && !legacy_fp8) {
    // fp8 caches run on tcgen05 (paged_decode_fp8.cu)
    if constexpr (KV != B200_KV_AUTO && D != 64) {
      int rc = launch_decode_fp8<QT, D, KV>(tk, tv, a, num_seqs, stream);
      if (rc) return rc;
    }
  } else {
    dim3 grid(a.num_splits, a.num_kv_heads * a.qgroups, num_seqs);
    launch_kernel(kern, dim3(grid), dim3(kWarps * 32), Cfg::SMEM_BYTES, stream, 1, tk, tv, a);
    B200_CHECK_LAUNCH();
    count_launch();
  }
  if (a.num_splits > 1) {
    dim3 g2((a.num_q_heads + 3) / 4, num_seqs);
    launch_kernel(paged_decode_combine_kernel<QT>, dim3(g2), dim3(128), 0, stream, 1, 
        reinterpret_cast<QT*>(a.out), a.out_lse, a.part_o, a.part_ml, KV == B200_KV_AUTO ? nullptr : a.v_scale,
        a.out_scale, a.num_q_heads, a.num_splits, D, a.out_stride_seq, a.out_stride_head);
    B200_CHECK_LAUNCH();
    count_launch();
  }
  return 0;
}

}  // namespace b200

extern "C" void b200_debug_attn_fp8_tcgen05(int mode) { b200::g_fp8_tcgen05 = mode; }

extern "C" size_t b200_paged_decode_workspace_bytes(int num_seqs, int num_q_heads, int head_size, int num_splits) {
  if (num_splits <= 1) return 0;
  return (size_t)num_seqs * num_q_heads * num_splits * ((size_t)head_size * 4 + 8);
}

extern "C" int b200_paged_decode_pick_splits_ex(int num_seqs, int num_kv_heads, int max_seq_len, int sm_count,
                                                int head_size, int kv_dtype) {
  // Deterministic in its arguments only (CUDA-graph safe).  Model fitted to the split sweeps in
  // profiles/r01_attn_splits.md: a CTA costs ~128 tokens' worth of prologue + merge, and the grid runs in
  // rounds of (resident CTAs per SM) x SMs, so efficiency(s) = fill(base*s / slots) * work / (work + 128).
  if (sm_count <= 0) sm_count = 148;
  static const int force = [] {
    const char* e = getenv("B200_ATTN_SPLITS");  // tuning aid
    return (e && *e) ? atoi(e) : 0;
  }();
  if (force > 0) return force;
  const long base = (long)num_seqs * num_kv_heads;
  if (base <= 0) return 1;
  const int row_bytes = head_size * (kv_dtype == B200_KV_AUTO ? 2 : 1);
  const int per_sm = (kv_dtype == B200_KV_AUTO) ? (row_bytes <= 128 ? 3 : (row_bytes == 256 ? 2 : 1))
                                                : (row_bytes <= 128 ? 3 : 1);
  const double slots = (double)per_sm * sm_count;
  int cap = max_seq_len / 256;  // every split streams at least 256 tokens
  if (cap > 128) cap = 128;
  if (cap < 1) cap = 1;
  int best = 1;
  double best_eff = -1.0;
  for (int s = 1; s <= cap; ++s) {
    const double waves = (double)base * s / slots;
    const double fill = waves / (double)(long)(waves + 0.999999);
    const double work = (double)max_seq_len / s;
    const double eff = fill * work / (work + 128.0);
    if (eff > best_eff * 1.01) {  // prefer fewer splits on near-ties
      best_eff = eff;
      best = s;
    }
  }
  return best;
}

extern "C" int b200_paged_decode_pick_splits(int num_seqs, int num_kv_heads, int max_seq_len, int sm_count) {
  return b200_paged_decode_pick_splits_ex(num_seqs, num_kv_heads, max_seq_len, sm_count, 128, B200_KV_FP8_E4M3);
}

extern "C" int b200_paged_attention_decode(const b200_paged_decode_params* p, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(p != nullptr, B200_ERR_INVALID_ARG, "paged_attention_decode: null params");
  B200_CHECK_ARG(p->num_seqs >= 0 && p->num_q_heads > 0 && p->num_kv_heads > 0 && p->head_size > 0 &&
                     p->block_size > 0,
                 B200_ERR_INVALID_ARG, "paged_attention_decode: bad sizes");
  B200_CHECK_ARG(p->num_q_heads % p->num_kv_heads == 0, B200_ERR_INVALID_ARG,
                 "paged_attention_decode: num_q_heads (%d) must be a multiple of num_kv_heads (%d)", p->num_q_heads,
                 p->num_kv_heads);
  B200_CHECK_ARG(p->block_size % kChunk == 0, B200_ERR_UNSUPPORTED,
                 "paged_attention_decode: block_size %d must be a multiple of 16", p->block_size);
  B200_CHECK_ARG(p->q_dtype == B200_DT_BF16 || p->q_dtype == B200_DT_F16, B200_ERR_UNSUPPORTED,
                 "paged_attention_decode: query dtype must be fp16 or bf16");
  B200_CHECK_ARG(p->kv_dtype == B200_KV_AUTO || p->kv_dtype == B200_KV_FP8_E4M3 || p->kv_dtype == B200_KV_FP8_E5M2,
                 B200_ERR_INVALID_ARG, "paged_attention_decode: Unsupported data type of kv cache: %d", p->kv_dtype);
  if (p->num_seqs == 0) return 0;
  B200_CHECK_ARG(p->out && p->query && p->key_cache && p->value_cache && p->block_table && p->seq_lens,
                 B200_ERR_INVALID_ARG, "paged_attention_decode: null pointer");
  const int es = p->kv_dtype == B200_KV_AUTO ? 2 : 1;
  const int D = p->head_size;
  B200_CHECK_ARG((D == 64 && es == 2) || D == 128 || D == 256, B200_ERR_UNSUPPORTED,
                 "paged_attention_decode: head_size %d with %d-byte cache elements is not built "
                 "(supported: 64 (16-bit), 128, 256)",
                 D, es);
  B200_CHECK_ARG((p->q_stride_seq % 2 == 0) && (p->q_stride_head % 2 == 0) &&
                     (reinterpret_cast<uintptr_t>(p->query) % 4 == 0),
                 B200_ERR_ALIGNMENT, "paged_attention_decode: query must be 4-byte aligned with even strides");
  auto tma_ok = [&](const void* base, int64_t hs, int64_t ps, int64_t bs) {
    return reinterpret_cast<uintptr_t>(base) % 16 == 0 && (hs * es) % 16 == 0 && (ps * es) % 16 == 0 &&
           (bs * es) % 16 == 0;
  };
  B200_CHECK_ARG(tma_ok(p->key_cache, p->kv_head_stride, p->kv_page_stride, p->kv_block_stride) &&
                     tma_ok(p->value_cache, p->v_head_stride, p->v_page_stride, p->v_block_stride),
                 B200_ERR_ALIGNMENT, "paged_attention_decode: cache base and strides must be 16-byte multiples");
  B200_CHECK_ARG(p->num_blocks > 0, B200_ERR_INVALID_ARG, "paged_attention_decode: num_blocks must be > 0");

  DecodeArgs a;
  a.out = p->out;
  a.query = p->query;
  a.block_table = p->block_table;
  a.seq_lens = p->seq_lens;
  a.k_scale = p->k_scale;
  a.v_scale = p->v_scale;
  a.out_lse = p->out_lse;
  a.out_scale = p->output_scale;
  a.token_to_seq = p->token_to_seq;
  a.q_stride_seq = p->q_stride_seq;
  a.q_stride_head = p->q_stride_head;
  a.out_stride_seq = p->out_stride_seq;
  a.out_stride_head = p->out_stride_head;
  a.block_table_stride = p->block_table_stride;
  a.num_q_heads = p->num_q_heads;
  a.num_kv_heads = p->num_kv_heads;
  a.group = p->num_q_heads / p->num_kv_heads;
  a.qgroups = (a.group + 15) / 16;
  a.block_size = p->block_size;
  a.sliding_window = p->sliding_window > 0 ? p->sliding_window : 0;
  a.scale = p->scale;
  a.soft_cap = p->logits_soft_cap;
  int splits = p->num_splits;
  if (splits <= 0)
    splits = b200_paged_decode_pick_splits_ex(p->num_seqs, p->num_kv_heads * a.qgroups,
                                              p->max_seq_len > 0 ? p->max_seq_len : 1, sm_count(), p->head_size,
                                              p->kv_dtype);
  a.num_splits = splits;
  a.eff_splits = splits;
  a.max_seq_len = p->max_seq_len;
  a.part_o = nullptr;
  a.part_ml = nullptr;
  if (splits > 1) {
    const size_t need = b200_paged_decode_workspace_bytes(p->num_seqs, p->num_q_heads, D, splits);
    B200_CHECK_ARG(p->workspace && p->workspace_bytes >= need, B200_ERR_WORKSPACE,
                   "paged_attention_decode: workspace %zu bytes < required %zu (splits=%d)", p->workspace_bytes,
                   need, splits);
    B200_CHECK_ARG(reinterpret_cast<uintptr_t>(p->workspace) % 16 == 0, B200_ERR_ALIGNMENT,
                   "paged_attention_decode: workspace must be 16-byte aligned");
    a.part_o = reinterpret_cast<float*>(p->workspace);
    a.part_ml = reinterpret_cast<float2*>(a.part_o + (size_t)p->num_seqs * p->num_q_heads * splits * D);
  }
  CUtensorMap tk, tv;
  int rc = get_cache_tmap(&tk, p->key_cache, D, p->num_kv_heads, p->block_size, p->num_blocks, p->kv_head_stride,
                          p->kv_page_stride, p->kv_block_stride, es);
  if (rc) return rc;
  rc = get_cache_tmap(&tv, p->value_cache, D, p->num_kv_heads, p->block_size, p->num_blocks, p->v_head_stride,
                      p->v_page_stride, p->v_block_stride, es);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
#define B200_DISPATCH_D(QT, KV)                                                    \
  switch (D) {                                                                     \
    case 128: return launch_decode<QT, 128, KV>(tk, tv, a, p->num_seqs, s);        \
    case 256: return launch_decode<QT, 256, KV>(tk, tv, a, p->num_seqs, s);        \
    default: break;                                                                \
  }
  if (p->kv_dtype == B200_KV_AUTO) {
    if (p->q_dtype == B200_DT_BF16) {
      if (D == 64) return launch_decode<__nv_bfloat16, 64, B200_KV_AUTO>(tk, tv, a, p->num_seqs, s);
      B200_DISPATCH_D(__nv_bfloat16, B200_KV_AUTO)
    } else {
      if (D == 64) return launch_decode<__half, 64, B200_KV_AUTO>(tk, tv, a, p->num_seqs, s);
      B200_DISPATCH_D(__half, B200_KV_AUTO)
    }
  } else if (p->kv_dtype == B200_KV_FP8_E4M3) {
    if (p->q_dtype == B200_DT_BF16) {
      B200_DISPATCH_D(__nv_bfloat16, B200_KV_FP8_E4M3)
    } else {
      B200_DISPATCH_D(__half, B200_KV_FP8_E4M3)
    }
  } else {
    if (p->q_dtype == B200_DT_BF16) {
      B200_DISPATCH_D(__nv_bfloat16, B200_KV_FP8_E5M2)
    } else {
      B200_DISPATCH_D(__half, B200_KV_FP8_E5M2)
    }
  }
#undef B200_DISPATCH_D
  set_error("paged_attention_decode: no kernel for head_size %d", D);
  return B200_ERR_UNSUPPORTED;
}

// ---- allocation-free variant ---------------------------------------------------------------------
// b200_paged_attention_decode_static_scratch: the same kernels, with the split-KV partials in a
// file-scope device array instead of a caller workspace, for callers that cannot allocate (CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The array lives in every GPU context that loads the
// library until the process exits.
namespace b200 {

constexpr size_t kStaticScratchBytes = (size_t)32 << 20;  // 32 MiB: cfg3 TP1 needs 6.4, cfg4 5.3, cfg5 4.8
static __device__ __align__(256) unsigned char g_decode_scratch[kStaticScratchBytes];

// In host code `g_decode_scratch` is only a stand-in for the symbol, not a device address: resolve it
// once per device, with that device current.
static int static_scratch_ptr(void** out) {
  static std::mutex mu;
  static void* ptrs[64] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("paged_attention_decode_static_scratch: cudaGetDevice failed or device %d out of range: %s", dev,
              cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_decode_scratch);
    if (e != cudaSuccess || !p) {
      set_error("paged_attention_decode_static_scratch: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

}  // namespace b200

extern "C" size_t b200_paged_decode_static_scratch_bytes(void) { return b200::kStaticScratchBytes; }

extern "C" int b200_paged_attention_decode_static_scratch(const b200_paged_decode_params* p, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(p != nullptr, B200_ERR_INVALID_ARG, "paged_attention_decode_static_scratch: null params");
  B200_CHECK_ARG(p->num_seqs >= 0 && p->num_q_heads > 0 && p->num_kv_heads > 0 && p->head_size > 0,
                 B200_ERR_INVALID_ARG, "paged_attention_decode_static_scratch: bad sizes");
  b200_paged_decode_params q = *p;
  if (q.num_splits <= 0) {  // the split count the ordinary entry point would pick
    const int group = q.num_q_heads / q.num_kv_heads;
    q.num_splits = b200_paged_decode_pick_splits_ex(q.num_seqs, q.num_kv_heads * ((group + 15) / 16),
                                                    q.max_seq_len > 0 ? q.max_seq_len : 1, sm_count(), q.head_size,
                                                    q.kv_dtype);
  }
  const size_t need = b200_paged_decode_workspace_bytes(q.num_seqs, q.num_q_heads, q.head_size, q.num_splits);
  B200_CHECK_ARG(need <= kStaticScratchBytes, B200_ERR_WORKSPACE,
                 "paged_attention_decode_static_scratch: %d seqs x %d heads x %d splits x head_size %d needs %zu bytes "
                 "of split-KV scratch, more than the built-in %zu; use b200_paged_attention_decode with a caller "
                 "workspace (or fewer splits)",
                 q.num_seqs, q.num_q_heads, q.num_splits, q.head_size, need, kStaticScratchBytes);
  void* scratch = nullptr;
  if (int rc = static_scratch_ptr(&scratch)) return rc;
  // Nothing in the scratch has to start at zero: every (sequence, head, split) slot the combine kernel reads is
  // written by this call first (an empty split writes (m, l) = (-inf, 0) and its stale accumulator is skipped).
  q.workspace = scratch;
  q.workspace_bytes = kStaticScratchBytes;
  return b200_paged_attention_decode(&q, stream);
}
