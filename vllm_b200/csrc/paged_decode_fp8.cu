// Paged-attention decode over an FP8 (e4m3 / e5m2) KV cache on the 5th-generation tensor cores.
//
// Reference semantics: the same as paged_decode.cu (legacy paged_attention_v1/v2,
// vllm/_custom_ops.py:120-163; kernel_unified_attention, triton_unified_attention.py:179-681): K and V are
// dequantised with the per-tensor k_scale / v_scale, logits are q.k * scale, softmax over the context,
// out = P V.  Here the scales are folded into the logit scale and the final normalisation.
//
// Why a second kernel: legacy mma.sync has no fp8 operand path on sm_100a (ptxas lowers
// mma.m16n8k32.e4m3 to F2FP unpacks + HMMA, scripts/ldm_probe.cu), so a register-level kernel converts
// every cache byte on the CUDA cores: ~450 warp instructions per 16 tokens, issue-bound at 78% of HBM.
// tcgen05.mma kind::f8f6f4 consumes the cache bytes straight from the shared memory TMA wrote them to:
//   S^T[128 tokens x 16 heads] = K_tile[128 x D] . Q^T        (A = K, K-major; B = Q, K-major)
//   O^T[D x 16 heads]          = V_tile^T[D x 128] . P^T      (A = V, MN-major = the [token][d] rows as they
//                                                              lie in the cache; B = P^T, K-major)
// so no thread ever touches a K or V byte.  Tokens sit on the TMEM lanes, which gives the softmax one
// thread per token and only the live query heads of the GQA group to work on (4 for Llama-3.1-8B).
//
// Precision: Q and P enter the fp8 MMAs as two e4m3 terms each (hi + lo): Q after a per-head power-of-two
// scale that puts the head's largest |q| in [64,128), P as p*2^8 with hi = p truncated to 3 mantissa bits.
// Both carry >= 7 significant bits - what rounding Q and P to bf16 gives the reference kernels.  Products
// with the exact fp8 cache values are exact in the fp32 accumulator.
//
// Pipeline per CTA (one (split, kv head, sequence), 384 threads):
//   warp 0 / 2  TMA producers for K / V: per 128-token tile 8 boxes of 16 tokens x 128 B (x D/128 spans), one
//               lane per 16-token piece, into separate rings of 128-byte-swizzled tiles (a K tile is released
//               right after QK^T, a V tile only after the softmax and PV);
//   warp 1 / 3  one thread each issues the QK^T / PV UMMAs; accumulators in TMEM, S and O double-buffered;
//   warps 4-11  two softmax groups (even / odd tiles), thread = token: tcgen05.ld the head columns,
//               scale/mask, tile max (redux + 4-warp exchange), p = 2^(s-m+8), hi/lo bytes into the P^T operand
//               tile, running sum per thread; then, as thread = d, fold the group's PREVIOUS tile's O^T into
//               registers (o = o*alpha + O_tile), so TMEM accumulators are never rescaled in place.  The two
//               groups keep independent (m, l, O) states, merged once at the end like split-KV partials.
#include <cuda.h>

#include <cstdlib>

#include "paged_decode.cuh"
#include "umma.cuh"

namespace b200 {

constexpr int kTile = 128;       // tokens per tile
constexpr int kF8Threads = 384;  // K producer, QK issuer, V producer, PV issuer, 2 softmax groups of 4 warps
constexpr int kSvcWarps = 4;     // warps ahead of the softmax groups
constexpr float kPShift = 8.f;   // probabilities are carried as p * 2^8

// debug: [cta][32] %globaltimer stamps of one steady-state tile + wait cycles (scripts/prof_attn.py), or null
__device__ unsigned long long* g_attn_trace = nullptr;
constexpr int kTraceTile = 4;
enum { TA_ENTRY, TA_SETUP, TA_TMA_ISSUE, TA_FULL, TA_QK_ISSUED, TA_S_FULL, TA_S_LOADED, TA_MAX_DONE, TA_P_READY,
       TA_PV_START, TA_PV_ISSUED, TA_O_FULL, TA_FOLDED, TA_LOOP_END, TA_EXIT };
// wait accounting: SM cycles a role spent blocked at one of its barriers, summed over the CTA's tiles
enum { TW_K_EMPTY = 16, TW_V_EMPTY, TW_K_FULL, TW_S_EMPTY, TW_V_FULL, TW_P_FULL, TW_O_EMPTY, TW_S_FULL, TW_O_FULL,
       TW_GRP_BAR, TW_TOTAL };
__device__ __forceinline__ int trace_cta() { return blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z); }
__device__ __forceinline__ void attn_stamp(int slot) {
  unsigned long long* tr = g_attn_trace;
  if (tr && trace_cta() < 4096) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tr[(size_t)trace_cta() * 32 + slot] = t;
  }
}
struct WaitClock {
  long long acc = 0;
  __device__ __forceinline__ void wait(uint64_t* bar, uint32_t parity) {
    if (g_attn_trace) {
      const long long t0 = clock64();
      mbar_wait(bar, parity);
      acc += clock64() - t0;
    } else {
      mbar_wait(bar, parity);
    }
  }
  __device__ __forceinline__ void flush(int slot) {
    unsigned long long* tr = g_attn_trace;
    if (tr && trace_cta() < 4096) tr[(size_t)trace_cta() * 32 + slot] = (unsigned long long)acc;
  }
};

template <int D> struct Fp8Cfg {
  static_assert(D % 128 == 0, "fp8 cache rows must be whole 128-byte swizzle spans");
  static constexpr int NSUB = D / 128;              // 128-byte spans per cache row
  static constexpr int SUB_TILE = kTile * 128;      // 16 KB: [128 tokens][128 B]
  static constexpr int OP_BYTES = NSUB * SUB_TILE;  // one K (or V) tile
  // K and V have separate rings: a K tile is released as soon as QK^T of its tile has run, a V tile only
  // after the softmax and PV - coupling them would hold K for the whole chain (measured: the ring, not the
  // softmax, bounds the kernel: throughput = ring bytes / (load latency + hold time)).
#ifndef B200_ATTN_KST
#define B200_ATTN_KST 3
#define B200_ATTN_VST 3
#endif
  static constexpr int KST = B200_ATTN_KST, VST = B200_ATTN_VST;
  static constexpr int RING_BYTES = (KST + VST) * OP_BYTES;
  static constexpr int Q_BYTES = 2 * NSUB * 16 * 128;  // Q hi, Q lo: [16 heads][D bytes] per span
  static constexpr int P_BYTES = 2 * 2 * 16 * 128;     // two buffers x (hi, lo) x [16 heads][128 tokens]
  static constexpr int MISC_BYTES = 2048;              // barriers, scales, exchange buffers
  static constexpr int SMEM_BYTES = RING_BYTES + Q_BYTES + P_BYTES + MISC_BYTES + 1024 /*align*/;
  static constexpr int S_COLS = 16;         // one S^T accumulator: 16 head columns
  static constexpr int O_COLS = NSUB * 16;  // one O^T accumulator set
  static constexpr int TMEM_COLS = (2 * S_COLS + 2 * O_COLS) <= 64 ? 64 : 128;
};

__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t smem_addr) {
  // MN-major operand, 128-byte swizzle: a row (one K index = one token) is a 128-byte span of the M index,
  // groups of 8 rows every SBO = 1024 bytes; the tile is one span wide so LBO is never used
  // (scripts/umma_probe.cu checks this encoding against a CPU product).
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void grp_bar(int grp) { asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory"); }
__device__ __forceinline__ void all_softmax_bar() { asm volatile("bar.sync 3, 256;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// TMEM -> registers: this warp's 32 lanes x NH consecutive fp32 columns
template <int NH> __device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[NH]) {
  if constexpr (NH == 4) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
  } else if constexpr (NH == 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
  } else {
    umma::ld_32x32b_x16(taddr, r);
  }
}
template <int NH> __device__ __forceinline__ void tmem_ld_wait(uint32_t (&r)[NH]) {
  // wait::ld with a data dependency on the destination registers
  if constexpr (NH == 4) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3])::"memory");
  } else if constexpr (NH == 8) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])::"memory");
  } else {
    umma::ld_wait_dep(r);
  }
}
// max over the warp of a float (any sign, -inf allowed): order-preserving map to s32 + redux.sync
__device__ __forceinline__ float warp_max_redux(float x) {
  int i = __float_as_int(x);
  i ^= (i >> 31) & 0x7fffffff;
  i = __reduce_max_sync(0xffffffffu, i);
  i ^= (i >> 31) & 0x7fffffff;
  return __int_as_float(i);
}

template <typename QT, int D, int KV, int NH>
__global__ void __launch_bounds__(kF8Threads, (Fp8Cfg<D>::SMEM_BYTES <= 113 * 1024 && NH <= 8) ? 2 : 1)
    paged_decode_fp8_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v,
                            const DecodeArgs a) {
  using Cfg = Fp8Cfg<D>;
  constexpr int NSUB = Cfg::NSUB;
  constexpr int KST = Cfg::KST, VST = Cfg::VST;
  constexpr uint32_t kvfmt = (KV == B200_KV_FP8_E5M2) ? umma::kFmtE5M2 : umma::kFmtE4M3;

  pdl_trigger();
  pdl_wait();  // query, caches, block table and seq_lens come from earlier kernels
  if (threadIdx.x == 0) attn_stamp(TA_ENTRY);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x, seq = blockIdx.z;
  const int kvh = blockIdx.y / a.qgroups, qg = blockIdx.y - kvh * a.qgroups;
  const int rows = min(16, a.group - qg * 16);  // live query heads of this CTA (<= NH)
  const int qh0 = kvh * a.group + qg * 16;
  const int len = a.seq_lens[seq];
  const int lo = (a.sliding_window > 0) ? max(0, len - a.sliding_window) : 0;
  const int c_begin = lo / kTile, c_end = (len + kTile - 1) / kTile;
  const int nc = max(0, c_end - c_begin);
  const int cps = (nc + a.eff_splits - 1) / a.eff_splits;
  const int c0 = c_begin + split * cps;
  const int c1 = min(c_end, c0 + cps);
  const int n_tiles = c1 - c0;

  if (n_tiles <= 0) {  // nothing to attend to in this split
    if (a.num_splits > 1) {
      if ((int)threadIdx.x < rows)
        a.part_ml[((int64_t)seq * a.num_q_heads + qh0 + threadIdx.x) * a.num_splits + split] =
            make_float2(-INFINITY, 0.f);
    } else {
      QT* o = reinterpret_cast<QT*>(a.out) + (int64_t)seq * a.out_stride_seq + (int64_t)qh0 * a.out_stride_head;
      for (int i = threadIdx.x; i < rows * D; i += blockDim.x)
        o[(int64_t)(i / D) * a.out_stride_head + (i % D)] = from_float<QT>(0.f);
      if (a.out_lse && (int)threadIdx.x < rows)
        a.out_lse[(int64_t)seq * a.num_q_heads + qh0 + threadIdx.x] = -INFINITY;
    }
    return;
  }

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* k_ring = smem;
  uint8_t* v_ring = smem + KST * Cfg::OP_BYTES;
  uint8_t* q_tiles = smem + Cfg::RING_BYTES;  // [hi|lo][NSUB][16][128]
  uint8_t* p_tiles = q_tiles + Cfg::Q_BYTES;  // [buf][hi|lo][16][128]
  uint8_t* misc = p_tiles + Cfg::P_BYTES;
  uint64_t* k_full = reinterpret_cast<uint64_t*>(misc);  // [KST]
  uint64_t* k_empty = k_full + KST;                      // [KST]
  uint64_t* v_full = k_empty + KST;                      // [VST]
  uint64_t* v_empty = v_full + VST;                      // [VST]
  uint64_t* s_full = v_empty + VST;                      // [2]
  uint64_t* s_empty = s_full + 2;                        // [2]
  uint64_t* p_full = s_empty + 2;                        // [2]
  uint64_t* o_full = p_full + 2;                         // [2]
  uint64_t* o_empty = o_full + 2;                        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);
  float* sc_l2 = reinterpret_cast<float*>(misc + 256);  // [16] exp2-domain logit scale per head
  float* sc_nat = sc_l2 + 16;                           // [16] natural-domain logit scale per head
  float* wmax = sc_nat + 16;                            // [group][parity][4 warps][16] tile max exchange
  float* lsum = wmax + 2 * 2 * 4 * 16;                  // [group][4 warps][16] final row-sum exchange

  if (threadIdx.x == 0) {
    for (int s = 0; s < KST; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
    }
    for (int s = 0; s < VST; ++s) {
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&s_full[b], 1);
      mbar_init(&s_empty[b], 4);
      mbar_init(&p_full[b], 4);
      mbar_init(&o_full[b], 1);
      mbar_init(&o_empty[b], 4);
    }
    fence_barrier_init();
    prefetch_tensormap(&tmap_k);
    prefetch_tensormap(&tmap_v);
  }
  if (warp == 1) umma::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);

  // ---- Q^T operand tiles: hi / lo e4m3 terms of q * 2^e(head), K-major rows of 128 B, 128-byte swizzle ----
  // zero everything first (dead head rows and the P^T rows of dead heads must be zero, not stale)
  for (int i = threadIdx.x; i < (Cfg::Q_BYTES + Cfg::P_BYTES) / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(q_tiles)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  {
    const float kscale = a.k_scale ? *a.k_scale : 1.f;
    const QT* qb = reinterpret_cast<const QT*>(a.query) + (int64_t)seq * a.q_stride_seq +
                   (int64_t)qh0 * a.q_stride_head;
    constexpr int PER = D / 32;  // elements per lane
    for (int h = warp; h < rows; h += kF8Threads / 32) {
      const QT* qr = qb + (int64_t)h * a.q_stride_head + lane * PER;
      float x[PER];
      float mx = 0.f;
#pragma unroll
      for (int i = 0; i < PER; i += 2) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(qr + i);  // 4-byte aligned (host check)
        const QT* e = reinterpret_cast<const QT*>(&w);
        x[i] = to_float(e[0]);
        x[i + 1] = to_float(e[1]);
        mx = fmaxf(mx, fmaxf(fabsf(x[i]), fabsf(x[i + 1])));
      }
      mx = warp_max(mx);
      float sq = 1.f, inv = 1.f;  // power-of-two scale: mx * sq in [64, 128)
      if (mx > 0.f && mx < INFINITY) {
        int ex;
        frexpf(mx, &ex);
        sq = ldexpf(1.f, 7 - ex);
        inv = ldexpf(1.f, ex - 7);
      }
      if (lane == 0) {
        sc_nat[h] = a.scale * kscale * inv;
        sc_l2[h] = a.scale * kscale * inv * kLog2e;
      }
#pragma unroll
      for (int i = 0; i < PER; i += 4) {
        float y[4], r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) y[j] = x[i + j] * sq;
        const uint32_t hi = pack4_e4m3(y[0], y[1], y[2], y[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) r[j] = y[j] - e4m3_to_float((hi >> (8 * j)) & 0xffu);
        const uint32_t lw = pack4_e4m3(r[0], r[1], r[2], r[3]);
        const int d = lane * PER + i;  // multiple of 4
        const int sub = d >> 7, c = d & 127;
        const uint32_t off = (uint32_t)(sub * 2048 + h * 128 + ((((c >> 4) ^ (h & 7)) << 4) | (c & 15)));
        *reinterpret_cast<uint32_t*>(q_tiles + off) = hi;
        *reinterpret_cast<uint32_t*>(q_tiles + NSUB * 2048 + off) = lw;
      }
    }
  }
  fence_async_smem();  // the UMMAs read the Q tiles through the async proxy
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) attn_stamp(TA_SETUP);
  const uint32_t tmem_s = tmem_base;                    // S^T buffers: cols [0,16) and [16,32)
  const uint32_t tmem_o = tmem_base + 2 * Cfg::S_COLS;  // O^T buffers: O_COLS each

  if (warp == 0 || warp == 2) {
    // ===== TMA producers (warp 0: K tiles, warp 2: V tiles): lane h < 8 owns the 16-token piece h of a tile =====
    const bool is_k = warp == 0;
    const CUtensorMap* tm = is_k ? &tmap_k : &tmap_v;
    uint64_t* fullb = is_k ? k_full : v_full;
    uint64_t* emptyb = is_k ? k_empty : v_empty;
    uint8_t* ringb = is_k ? k_ring : v_ring;
    const int nst = is_k ? KST : VST;
    const int32_t* bt_row =
        a.block_table + (int64_t)(a.token_to_seq ? a.token_to_seq[seq] : seq) * a.block_table_stride;
    int stage = 0;
    uint32_t phase = 0;
    WaitClock w_empty, w_issue, w_loop;
    const long long tp0 = clock64();
    // Block ids are fetched 4 tiles (32 pieces, one per lane) at a time and one batch ahead: a dependent
    // global load per tile in this loop caps the producer at ~1 tile/us (measured), below the HBM rate.
    auto load_ids = [&](int batch) -> int {
      const int pc = batch * 32 + lane, tile = pc >> 3;
      if (tile < n_tiles) {
        const int tok = (c0 + tile) * kTile + (pc & 7) * kChunk;
        if (tok < len) return bt_row[tok / a.block_size];
      }
      return 0;
    };
    int ids_cur = load_ids(0), ids_next = load_ids(1);
    for (int i = 0; i < n_tiles; ++i) {
      if ((i & 3) == 0 && i > 0) {
        ids_cur = ids_next;
        ids_next = load_ids((i >> 2) + 1);
      }
      const int tok0 = (c0 + i) * kTile;
      const int my_tok = tok0 + lane * kChunk;
      const bool mine = lane < kTile / kChunk && my_tok < len;  // pieces that start past the sequence are not loaded
      const int blk = __shfl_sync(0xffffffffu, ids_cur, (i & 3) * 8 + (lane & 7));
      const int n_valid = min(kTile / kChunk, (len - tok0 + kChunk - 1) / kChunk);
      if (lane == 0) {
        w_empty.wait(&emptyb[stage], phase ^ 1);
        if (i == kTraceTile && is_k) attn_stamp(TA_TMA_ISSUE);
        mbar_arrive_expect_tx(&fullb[stage], (uint32_t)n_valid * NSUB * (kChunk * 128));
      }
      __syncwarp();
      const long long ti0 = clock64();
      if (mine) {
        uint8_t* dst = ringb + stage * Cfg::OP_BYTES + lane * (kChunk * 128);
        const int tok_in_blk = my_tok % a.block_size;
#pragma unroll
        for (int sb = 0; sb < NSUB; ++sb)
          tma_load_4d(dst + sb * Cfg::SUB_TILE, tm, &fullb[stage], sb * 128, kvh, tok_in_blk, blk, kEvictFirst);
      }
      __syncwarp();
      w_issue.acc += clock64() - ti0;
      if (++stage == nst) {
        stage = 0;
        phase ^= 1;
      }
    }
    if (lane == 0) {
      w_empty.flush(is_k ? TW_K_EMPTY : TW_V_EMPTY);
      if (is_k) {
        w_issue.flush(27);
        w_loop.acc = clock64() - tp0;
        w_loop.flush(28);
      }
    }
  } else if (warp == 1) {
    // ===== QK^T issuer =====
    if (umma::elect_one()) {
      constexpr uint32_t idesc_qk = umma::make_idesc_f8(128, 16, kvfmt, umma::kFmtE4M3);
      const uint32_t q_addr = smem_u32(q_tiles);
      WaitClock w_kf, w_se;
      for (int i = 0; i < n_tiles; ++i) {
        const int st = i % KST, b = i & 1;
        w_kf.wait(&k_full[st], (uint32_t)(i / KST) & 1);
        if (i == kTraceTile) attn_stamp(TA_FULL);
        w_se.wait(&s_empty[b], (((uint32_t)i >> 1) & 1) ^ 1);
        umma::fence_after_sync();
        const uint32_t k_addr = smem_u32(k_ring + st * Cfg::OP_BYTES);
        uint32_t acc = 0;
#pragma unroll
        for (int sb = 0; sb < NSUB; ++sb) {
          const uint64_t dk = umma::make_desc_k_sw128(k_addr + sb * Cfg::SUB_TILE);
          const uint64_t dqh = umma::make_desc_k_sw128(q_addr + sb * 2048);
          const uint64_t dql = umma::make_desc_k_sw128(q_addr + (NSUB + sb) * 2048);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            umma::mma_f8f6f4(tmem_s + b * Cfg::S_COLS, umma::desc_advance(dk, ks * 32), umma::desc_advance(dqh, ks * 32),
                             idesc_qk, acc);
            acc = 1;
            umma::mma_f8f6f4(tmem_s + b * Cfg::S_COLS, umma::desc_advance(dk, ks * 32), umma::desc_advance(dql, ks * 32),
                             idesc_qk, 1);
          }
        }
        umma::commit(&s_full[b]);
        umma::commit(&k_empty[st]);  // the K tile is consumed once these MMAs have run
        if (i == kTraceTile) attn_stamp(TA_QK_ISSUED);
      }
      w_kf.flush(TW_K_FULL);
      w_se.flush(TW_S_EMPTY);
    }
  } else if (warp == 3) {
    // ===== PV issuer =====
    if (umma::elect_one()) {
      constexpr uint32_t idesc_pv = umma::make_idesc_f8(128, 16, kvfmt, umma::kFmtE4M3) | (1u << 15);  // A MN-major
      WaitClock w_vf, w_pf, w_oe;
      for (int i = 0; i < n_tiles; ++i) {
        const int st = i % VST, b = i & 1;
        if (i == kTraceTile) attn_stamp(TA_PV_START);
        w_vf.wait(&v_full[st], (uint32_t)(i / VST) & 1);
        w_pf.wait(&p_full[b], ((uint32_t)i >> 1) & 1);
        w_oe.wait(&o_empty[b], (((uint32_t)i >> 1) & 1) ^ 1);
        umma::fence_after_sync();
        const uint32_t v_addr = smem_u32(v_ring + st * Cfg::OP_BYTES);
        const uint32_t p_addr = smem_u32(p_tiles + b * 4096);
        const uint64_t dph = umma::make_desc_k_sw128(p_addr);
        const uint64_t dpl = umma::make_desc_k_sw128(p_addr + 2048);
#pragma unroll
        for (int sb = 0; sb < NSUB; ++sb) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {  // 32 tokens per step = 4 groups of 8 rows
            const uint64_t dv = make_desc_mn_sw128(v_addr + sb * Cfg::SUB_TILE + ks * 4096);
            umma::mma_f8f6f4(tmem_o + b * Cfg::O_COLS + sb * 16, dv, umma::desc_advance(dph, ks * 32), idesc_pv, ks != 0);
            umma::mma_f8f6f4(tmem_o + b * Cfg::O_COLS + sb * 16, dv, umma::desc_advance(dpl, ks * 32), idesc_pv, 1);
          }
        }
        umma::commit(&o_full[b]);
        umma::commit(&v_empty[st]);
        if (i == kTraceTile) attn_stamp(TA_PV_ISSUED);
      }
      w_vf.flush(TW_V_FULL);
      w_pf.flush(TW_P_FULL);
      w_oe.flush(TW_O_EMPTY);
    }
  } else {
    // ===== softmax / accumulate warps: thread = token (S^T lanes), then thread = d (O^T lanes) =====
    // Two groups of 4 warps run two independent online-softmax streams: group g owns the tiles i = g (mod 2)
    // and with them S / P / O buffer g, so the streams never wait for each other; their (m, l, O) states are
    // merged once at the end like split-KV partials.
    const int grp = (warp - kSvcWarps) >> 2;
    const int q4 = warp & 3;        // TMEM lane quarter this warp may access
    const int tl = q4 * 32 + lane;  // lane = token within the tile / d within a 128-wide span
    const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
    const bool use_cap = a.soft_cap > 0.f;
    const float cap_inv = use_cap ? 1.f / a.soft_cap : 0.f;
    const float cap_l2 = a.soft_cap * kLog2e;
    float scl[NH], scn[NH];
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      scl[h] = h < rows ? sc_l2[h] : 0.f;
      scn[h] = h < rows ? sc_nat[h] : 0.f;
    }
    float m_run[NH], l_thr[NH], alpha_prev[NH], o_reg[NSUB][NH];
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      m_run[h] = -INFINITY;
      l_thr[h] = 0.f;
      alpha_prev[h] = 1.f;
#pragma unroll
      for (int sb = 0; sb < NSUB; ++sb) o_reg[sb][h] = 0.f;
    }
    WaitClock w_sf, w_of, w_gb;
    const long long t_begin = clock64();
    const int b = grp;  // S / P / O buffer of this group
    auto fold_o = [&](int j) {  // o = o * alpha_j + O_tile_j   (thread = d)
      w_of.wait(&o_full[b], ((uint32_t)j >> 1) & 1);
      umma::fence_after_sync();
      if (j == kTraceTile && tl == 0) attn_stamp(TA_O_FULL);
#pragma unroll
      for (int sb = 0; sb < NSUB; ++sb) {
        uint32_t r[NH];
        tmem_ld_cols<NH>(tmem_o + lane_off + b * Cfg::O_COLS + sb * 16, r);
        tmem_ld_wait<NH>(r);
#pragma unroll
        for (int h = 0; h < NH; ++h) o_reg[sb][h] = o_reg[sb][h] * alpha_prev[h] + __uint_as_float(r[h]);
      }
      umma::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_empty[b]);
      if (j == kTraceTile && tl == 0) attn_stamp(TA_FOLDED);
    };

    for (int i = grp; i < n_tiles; i += 2) {
      const int par = (i >> 1) & 1;
      const int tok0 = (c0 + i) * kTile;
      const int tok = tok0 + tl;
      w_sf.wait(&s_full[b], (uint32_t)par);
      umma::fence_after_sync();
      const bool trc = (i == kTraceTile && tl == 0);
      if (trc) attn_stamp(TA_S_FULL);
      uint32_t sr[NH];
      tmem_ld_cols<NH>(tmem_s + lane_off + b * Cfg::S_COLS, sr);
      tmem_ld_wait<NH>(sr);
      umma::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_empty[b]);  // S buffer may be overwritten by QK^T of tile i+2
      if (trc) attn_stamp(TA_S_LOADED);
#ifdef B200_ATTN_EXP_NOSOFTMAX  // timing experiment only (wrong results): no softmax work between QK^T and PV
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[b]);
      if (i >= 2) fold_o(i - 2);
      continue;
#endif

      if (tok0 + kTile > len) {
        // boundary tile: V rows past the sequence were never loaded (or hold stale bytes); p is 0 there but
        // 0 * NaN would poison O, so clear this thread's V row.  K garbage only reaches masked logits.
        mbar_wait(&v_full[i % VST], (uint32_t)(i / VST) & 1);  // landed: TMA must not overwrite the zeros
        if (tok >= len) {
          uint8_t* vrow = v_ring + (i % VST) * Cfg::OP_BYTES + tl * 128;
#pragma unroll
          for (int sb = 0; sb < NSUB; ++sb)
#pragma unroll
            for (int c = 0; c < 8; ++c)
              *reinterpret_cast<uint4*>(vrow + sb * Cfg::SUB_TILE + c * 16) = make_uint4(0, 0, 0, 0);
        }
      }
      const bool dead = tok >= len || tok < lo;
      float v[NH], tmax[NH];
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        const float s = __uint_as_float(sr[h]);
        float x = use_cap ? cap_l2 * fast_tanh(s * scn[h] * cap_inv) : s * scl[h];
        if (dead || h >= rows) x = -INFINITY;
        v[h] = x;
        tmax[h] = warp_max_redux(x);
      }
      // every lane holds every head's warp max; lane h publishes head h
#pragma unroll
      for (int h = 0; h < NH; ++h)
        if (lane == h) wmax[((b * 2 + par) * 4 + q4) * 16 + h] = tmax[h];
      {
        const long long t0 = clock64();
        grp_bar(grp);
        w_gb.acc += clock64() - t0;
      }
      if (trc) attn_stamp(TA_MAX_DONE);
      float al[NH];
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        const float* wm = wmax + (b * 2 + par) * 64 + h;
        const float t4 = fmaxf(fmaxf(wm[0], wm[16]), fmaxf(wm[32], wm[48]));
        const float mn = fmaxf(m_run[h], t4);
        const float ms = (mn == -INFINITY) ? 0.f : mn;
        al[h] = fast_exp2(m_run[h] - ms);
        m_run[h] = mn;
        const float p = fast_exp2(v[h] - ms + kPShift);
        l_thr[h] = l_thr[h] * al[h] + p;
        v[h] = p;
      }
      // Fold this group's previous tile first: waiting for its O^T also guarantees that PV(i-2) has finished
      // reading P^T buffer b, which is overwritten next, and keeps p_full from lapping its consumer (a
      // barrier that completes two phases ahead of a waiter deadlocks it).
      if (i >= 2) fold_o(i - 2);
      // p * 2^8 = hi + lo: hi = p truncated to 3 mantissa bits (exactly an e4m3), lo rounded to e4m3
      {
        uint8_t* ph = p_tiles + b * 4096;
#pragma unroll
        for (int h = 0; h < NH; h += 2) {
          const float h0 = __uint_as_float(__float_as_uint(v[h]) & 0xfff00000u);
          const float h1 = __uint_as_float(__float_as_uint(v[h + 1]) & 0xfff00000u);
          const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(h0, h1), __NV_SATFINITE, __NV_E4M3);
          const uint32_t lw = __nv_cvt_float2_to_fp8x2(make_float2(v[h] - h0, v[h + 1] - h1), __NV_SATFINITE, __NV_E4M3);
          const uint32_t o0 = (uint32_t)(h * 128 + ((((tl >> 4) ^ (h & 7)) << 4) | (tl & 15)));
          const uint32_t o1 = (uint32_t)((h + 1) * 128 + ((((tl >> 4) ^ ((h + 1) & 7)) << 4) | (tl & 15)));
          ph[o0] = (uint8_t)(hi & 0xff);
          ph[o1] = (uint8_t)(hi >> 8);
          ph[2048 + o0] = (uint8_t)(lw & 0xff);
          ph[2048 + o1] = (uint8_t)(lw >> 8);
        }
      }
      fence_async_smem();  // P^T (and a cleared V row) must be visible to the async proxy before PV is issued
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[b]);
      if (trc) attn_stamp(TA_P_READY);
#pragma unroll
      for (int h = 0; h < NH; ++h) alpha_prev[h] = al[h];
    }
    if (n_tiles > grp) fold_o(((n_tiles - 1 - grp) & ~1) + grp);  // this group's last tile
    if (tl == 0 && grp == 0) {
      attn_stamp(TA_LOOP_END);
      w_sf.flush(TW_S_FULL);
      w_of.flush(TW_O_FULL);
      w_gb.flush(TW_GRP_BAR);
      WaitClock tot;
      tot.acc = clock64() - t_begin;
      tot.flush(TW_TOTAL);
    }
    // ---- row sums: thread-local partial sums -> group totals ----------------------------
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const float sgl = warp_sum(l_thr[h]);
      if (lane == h) lsum[(grp * 4 + q4) * 16 + h] = sgl;
    }
    grp_bar(grp);
    float l_grp[NH];
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const float* ls = lsum + grp * 64 + h;
      l_grp[h] = ls[0] + ls[16] + ls[32] + ls[48];
    }
    // ---- merge the two streams: group 1 parks (m, l, O) in the (idle) ring, group 0 finishes -------------
    all_softmax_bar();  // both groups have folded their last tile: every MMA has completed, the ring is free
    float* xm = reinterpret_cast<float*>(smem);  // [16] m, [16] l
    float* xo = xm + 32;                         // [NSUB][NH][128] O
    if (grp == 1) {
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        if (tl == 0) {
          xm[h] = m_run[h];
          xm[16 + h] = l_grp[h];
        }
#pragma unroll
        for (int sb = 0; sb < NSUB; ++sb) xo[(sb * NH + h) * 128 + tl] = o_reg[sb][h];
      }
    }
    all_softmax_bar();
    if (grp == 0) {
      constexpr float kUn = 1.f / 256.f;  // undo the 2^8 carried by p (numerator and denominator alike)
      const float vscale = a.v_scale ? *a.v_scale : 1.f;
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        if (h >= rows) continue;
        const float m1 = xm[h], l1 = xm[16 + h];
        const float mt = fmaxf(m_run[h], m1);
        const float w0 = (m_run[h] == -INFINITY) ? 0.f : fast_exp2(m_run[h] - mt);
        const float w1 = (m1 == -INFINITY) ? 0.f : fast_exp2(m1 - mt);
        const float lt = (l_grp[h] * w0 + l1 * w1) * kUn;
        const int qh = qh0 + h;
        if (a.num_splits > 1) {
          const int64_t pi = ((int64_t)seq * a.num_q_heads + qh) * a.num_splits + split;
#pragma unroll
          for (int sb = 0; sb < NSUB; ++sb)
            a.part_o[pi * D + sb * 128 + tl] = (o_reg[sb][h] * w0 + xo[(sb * NH + h) * 128 + tl] * w1) * kUn;
          if (tl == 0) a.part_ml[pi] = make_float2(mt, lt);
        } else {
          QT* op = reinterpret_cast<QT*>(a.out) + (int64_t)seq * a.out_stride_seq + (int64_t)qh * a.out_stride_head;
#pragma unroll
          for (int sb = 0; sb < NSUB; ++sb) {
            const float acc = (o_reg[sb][h] * w0 + xo[(sb * NH + h) * 128 + tl] * w1) * kUn;
            op[sb * 128 + tl] = from_float<QT>(lt > 0.f ? acc / lt * vscale : 0.f);
          }
          if (tl == 0 && a.out_lse)
            a.out_lse[(int64_t)seq * a.num_q_heads + qh] = lt > 0.f ? (mt + log2f(lt)) * kLn2 : -INFINITY;
        }
      }
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) attn_stamp(TA_EXIT);
  if (warp == 1) umma::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

template <typename QT, int D, int KV, int NH>
static int launch_fp8_nh(const CUtensorMap& tk, const CUtensorMap& tv, const DecodeArgs& a, int num_seqs,
                         cudaStream_t stream) {
  using Cfg = Fp8Cfg<D>;
  auto kern = paged_decode_fp8_kernel<QT, D, KV, NH>;
  if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc_attr;
  dim3 grid(a.num_splits, a.num_kv_heads * a.qgroups, num_seqs);
  launch_kernel(kern, grid, dim3(kF8Threads), Cfg::SMEM_BYTES, stream, 1, tk, tv, a);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

template <typename QT, int D, int KV>
int launch_decode_fp8(const CUtensorMap& tk, const CUtensorMap& tv, const DecodeArgs& a_in, int num_seqs,
                      cudaStream_t stream) {
  // Fewer, longer CTAs than the caller's split count when that pays: a CTA spends ~2 tile-times in its
  // prologue/merge, and the grid should still fill the 2-per-SM slots evenly.  Splits >= eff exit at once
  // with empty partials, so the caller's workspace layout and combine stay valid.
  DecodeArgs a = a_in;
  {
    static const int force = [] {
      const char* e = getenv("B200_ATTN_FP8_SPLITS");
      return (e && *e) ? atoi(e) : 0;
    }();
    const long base = (long)num_seqs * a.num_kv_heads * a.qgroups;
    const long slots = 2L * sm_count();
    const int total_tiles = a.max_seq_len > 0 ? (a.max_seq_len + kTile - 1) / kTile : 32 * a.num_splits;
    int best = 1;
    double best_eff = 0.0;
    for (int s = 1; s <= a.num_splits; ++s) {
      const double waves = (double)(base * s) / slots;
      const double fill = waves / (double)((long)(waves + 0.999999));
      const double per = (double)((total_tiles + s - 1) / s);  // tiles per CTA
      const double eff = fill * per / (per + 2.0);
      if (eff > best_eff * 1.02) {
        best_eff = eff;
        best = s;
      }
    }
    a.eff_splits = (force > 0 && force <= a.num_splits) ? force : best;
  }
  const int live = a.group < 16 ? a.group : 16;  // heads per CTA
  if (live <= 4) return launch_fp8_nh<QT, D, KV, 4>(tk, tv, a, num_seqs, stream);
  if (live <= 8) return launch_fp8_nh<QT, D, KV, 8>(tk, tv, a, num_seqs, stream);
  return launch_fp8_nh<QT, D, KV, 16>(tk, tv, a, num_seqs, stream);
}

#define B200_INST(QT, D, KV)                                                                                    \
  template int launch_decode_fp8<QT, D, KV>(const CUtensorMap&, const CUtensorMap&, const DecodeArgs&, int, \
                                            cudaStream_t);
B200_INST(__nv_bfloat16, 128, B200_KV_FP8_E4M3)
B200_INST(__nv_bfloat16, 256, B200_KV_FP8_E4M3)
B200_INST(__half, 128, B200_KV_FP8_E4M3)
B200_INST(__half, 256, B200_KV_FP8_E4M3)
B200_INST(__nv_bfloat16, 128, B200_KV_FP8_E5M2)
B200_INST(__nv_bfloat16, 256, B200_KV_FP8_E5M2)
B200_INST(__half, 128, B200_KV_FP8_E5M2)
B200_INST(__half, 256, B200_KV_FP8_E5M2)
#undef B200_INST

}  // namespace b200

extern "C" void b200_debug_attn_trace(unsigned long long* device_buffer) {
  cudaMemcpyToSymbol(b200::g_attn_trace, &device_buffer, sizeof(device_buffer));
}
