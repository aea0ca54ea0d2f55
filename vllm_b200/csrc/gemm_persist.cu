// Persistent 2-CTA (cta_group::2) FP8 GEMM for the compute-bound regime (M > 256):
//   out[M,N] = a_s[M|1] * b_s[N|1] * (A[M,K] @ B[N,K]^T) (+ bias[N]),  fp8 e4m3 operands, fp32 accumulate in TMEM.
// Reference semantics: cutlass_scaled_mm (scaled_mm_entry.cu:197-270); the reference's sm100 large-M config is
// c3x/scaled_mm_sm100_fp8_dispatch.cuh:100-116 (256x128x128 tile, 2x2 cluster, 2-SM MMA).
//
// B200 design
//  * a CLUSTER OF TWO CTAs (one TPC) computes a 256 x BN output tile with ONE tcgen05.mma.cta_group::2 stream issued
//    by the leader CTA: UMMA M = 256 (CTA r owns token rows 128r..128r+127 in its TMEM lanes), UMMA N = BN = 256;
//    each CTA TMA-loads only ITS 128 rows of A and ITS BN/2 rows of B per k-block (32 KB) - the tensor core reads
//    the other half of B out of the peer's shared memory, so per-SM shared-memory fill is half of what two
//    independent 128 x BN tiles would need (the measured per-SM ingest cap is ~125 GB/s, profiles/r02_*).
//  * PERSISTENT: one cluster per TPC, static round-robin over the output tiles (M fastest, so clusters running at
//    the same time share the weight tile in L2); the smem ring keeps streaming across tile boundaries.
//  * two TMEM accumulators (2 x 256 columns = all of TMEM): the epilogue of tile i (tcgen05.ld -> scales -> bf16 ->
//    global) runs while the MMAs of tile i+1 fill the other accumulator.
//  * warp roles per CTA: warp 0 TMA producer, warp 1 TMEM allocator + (leader only) MMA issuer, warps 2-9 epilogue
//    (two per TMEM lane quarter, each half of the tile's columns).  Barriers: full[] lives in the leader (both CTAs' TMA bytes land on it),
//    empty[] / tmem_full[] are per CTA and signalled by multicast tcgen05.commit, tmem_empty[] lives in the leader
//    and counts the epilogue warps of BOTH CTAs.
#include <cuda.h>

#include <cstdlib>

#include "umma.cuh"

namespace b200 {

int get_kmajor_tmap(CUtensorMap* out, const void* base, int rows, int k_bytes, int64_t ld_bytes, int box_rows);

unsigned long long* gemm_debug_trace();  // fp8_gemm_decode.cu (b200_debug_gemm_trace)

constexpr int kP2Threads = 320;  // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)
constexpr int kP2BK = 128;  // bytes of K per stage (one 128-byte swizzle span = 128 fp8)

// UM = UMMA M across the pair: 256 (each CTA owns 128 token rows; TMEM lane = row, BN columns per accumulator) or
// 128 (decode regime: each CTA owns 64 token rows; TMEM data-path "layout B": lanes 0-63 hold rows 0-63 x columns
// [0, BN/2), lanes 64-127 hold the SAME rows x columns [BN/2, BN), so an accumulator is BN/2 TMEM columns wide).
template <int UM, int BN> struct P2Cfg {
  static constexpr int A_ROWS = UM / 2;               // this CTA's token rows
  static constexpr int A_BYTES = A_ROWS * kP2BK;
  static constexpr int B_BYTES = (BN / 2) * kP2BK;    // this CTA's half of the weight tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int MAX_STAGES = UM == 128 ? 12 : 8;  // decode: latency-bound stream, keep more bytes in flight
  static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > MAX_STAGES ? MAX_STAGES : (200 * 1024) / STAGE_BYTES;
  static constexpr int ACC_COLS = UM == 256 ? BN : BN / 2;
  static constexpr int TMEM_COLS =
      2 * ACC_COLS <= 32 ? 32 : (2 * ACC_COLS <= 64 ? 64 : (2 * ACC_COLS <= 128 ? 128 : (2 * ACC_COLS <= 256 ? 256 : 512)));
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/;
  static_assert(UM == 128 || UM == 256, "UMMA M across the pair");
  static_assert(2 * ACC_COLS <= 512, "two accumulators must fit the 512 TMEM columns");
  static_assert((2 * STAGES + 4) * 8 + 8 <= 512, "barrier area");
};

struct P2Args {
  void* out;
  const float* a_scales;  // per token (M) or [1]
  const float* b_scales;  // per channel (N) or [1]
  const void* bias;       // [N] out dtype or null
  int64_t ldc;
  int M, N, K;
  int a_scales_numel, b_scales_numel;
  int tiles_m, tiles_n, num_kb;
  float* dump;  // debug only (b200_debug_gemm_dump), normally null
  unsigned long long* trace;  // debug only (b200_debug_gemm_trace): [gridDim.x][16] %globaltimer stamps
};
// phase stamps for scripts/prof_gemm.py --trace (slot names there); one thread per CTA writes each slot
__device__ __forceinline__ void p2_stamp(const P2Args& g, int slot) {
  if (g.trace) g.trace[(size_t)blockIdx.x * 16 + slot] = globaltimer_ns();
}

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // shared::cluster address of the same offset in CTA rank 0 of the pair

__device__ __forceinline__ uint32_t p2_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void p2_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Final rendezvous of the pair before TMEM dealloc / exit.  It only keeps each CTA's shared memory and barriers
// alive until the other is done with them (all such accesses were already observed complete through tmem_full);
// RELAXED, because the .release form first drains the epilogue warps' output stores (~0.5 us at the kernel's tail).
__device__ __forceinline__ void p2_cluster_sync_exit() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
}
// TMA load whose completion bytes are counted on the LEADER CTA's mbarrier (executed by both CTAs of the pair)
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                                uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void mma_f8f6f4_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void commit_2sm(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// arrive on the barrier at this offset in the LEADER CTA (local arrive when executed by the leader).  RELAXED on
// purpose: the epilogue warps use it to hand a TMEM accumulator back, and what has to be ordered before it - their
// tcgen05.ld reads - is complete (tcgen05.wait::ld) and fenced (tcgen05.fence::before_thread_sync).  A .release
// arrive compiles to MEMBAR.ALL.CTA and waits for the warp's outstanding OUTPUT STORES as well: measured 2.6 us per
// 256x128 tile at K = 1024 against a 0.8 us mainloop (profiles/r02_gemm2sm.md).
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask)
               : "memory");
}
template <int kCols> __device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols> __device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}

template <typename OutT> __device__ __forceinline__ uint32_t p2_pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t p2_pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}
template <> __device__ __forceinline__ uint32_t p2_pack2<__half>(float a, float b) {
  const __half2 t = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}

// 16 consecutive 16-bit outputs of one row = one full 32-byte sector in ONE store (sm_100 256-bit STG): a lane owns
// a row, so two 16-byte stores would each dirty half a sector
__device__ __forceinline__ void st_global_256(void* dst, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
               "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

// Epilogue of the 2-CTA kernels (both CTAs, warps 2-9): warp w owns TMEM lanes 32*(w%4).. and half of that lane
// quarter's columns; tcgen05.ld -> scales (+ bias) -> 16-bit -> one 256-bit store per 16 columns, then the
// accumulator is handed back to the leader's MMA warp.
template <int UM, int BN, int ACC_COLS, typename OutT>
__device__ __forceinline__ void p2_epilogue(const P2Args& g, uint32_t tmem_base, uint64_t* tmem_full_bar,
                                            uint64_t* tmem_empty_bar, int warp, int lane, int rank, int pair,
                                            int num_pairs) {
  const int num_tiles = g.tiles_m * g.tiles_n;
  const int q = warp & 3;
  OutT* outp = reinterpret_cast<OutT*>(g.out);
  const OutT* bias = reinterpret_cast<const OutT*>(g.bias);
  const bool vec_ok = ((g.ldc * (int64_t)sizeof(OutT)) % 32 == 0) && (reinterpret_cast<uintptr_t>(g.out) % 32 == 0);
  const float bs0 = g.b_scales[0];
  int it = 0;
  for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
    const int tm = t % g.tiles_m, tn = t / g.tiles_m;
    const int acc = it & 1;
    // UM = 256: lane quarter q = rows 32q.., all BN columns.  UM = 128: quarters 0,1 = rows 0..63 x columns
    // [0, BN/2), quarters 2,3 = the same rows x columns [BN/2, BN) (TMEM columns [0, BN/2) in both cases).
    const int m = UM == 256 ? tm * 256 + rank * 128 + q * 32 + lane : tm * 128 + rank * 64 + (q & 1) * 32 + lane;
    const int n0 = tn * BN + (UM == 256 ? 0 : (q >> 1) * (BN / 2));
    const float as = m < g.M ? g.a_scales[g.a_scales_numel > 1 ? m : 0] : 0.f;
    mbar_wait(&tmem_full_bar[acc], ((uint32_t)it >> 1) & 1);
    umma::fence_after_sync();
    if (warp == 2 && lane == 0 && t + num_pairs >= num_tiles) p2_stamp(g, 5);
    const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + acc * ACC_COLS;
    uint32_t rn[16];
    // the two warps of a lane quarter split that quarter's columns
    const int c_lo = ((warp - 2) >> 2) * (ACC_COLS / 2), c_hi = c_lo + ACC_COLS / 2;
    umma::ld_32x32b_x16(tacc + c_lo, rn);
#pragma unroll 1
    for (int c = c_lo; c < c_hi; c += 16) {
      umma::ld_wait_dep(rn);
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rn[j]);
      if (c + 16 < c_hi) umma::ld_32x32b_x16(tacc + c + 16, rn);  // next chunk in flight while this one is stored
      if (g.dump) {  // debug: raw TMEM image [cta][lane][column] (b200_debug_gemm_dump)
        float* d = g.dump + ((int64_t)(rank * 128 + q * 32 + lane)) * ACC_COLS + c;
#pragma unroll
        for (int j = 0; j < 16; ++j) d[j] = v[j];
      }
      const int n = n0 + c;
      if (m < g.M && n < g.N) {  // N % 16 == 0: a chunk is entirely inside or outside
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float bs = g.b_scales_numel > 1 ? g.b_scales[n + j] : bs0;
          v[j] = v[j] * as * bs + (bias ? to_float(bias[n + j]) : 0.f);
        }
        OutT* dst = outp + (int64_t)m * g.ldc + n;
        if (vec_ok) {
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = p2_pack2<OutT>(v[2 * i], v[2 * i + 1]);
          st_global_256(dst, w);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) dst[j] = from_float<OutT>(v[j]);
        }
      }
    }
    umma::fence_before_sync();
    __syncwarp();
    if (lane == 0) mbar_arrive_leader(&tmem_empty_bar[acc]);
    if (warp == 2 && lane == 0 && t + num_pairs >= num_tiles) p2_stamp(g, 8);
  }
}

template <int UM, int BN, typename OutT>
__global__ void __launch_bounds__(kP2Threads, 1)
    fp8_gemm_2sm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                        const P2Args g) {
  using Cfg = P2Cfg<UM, BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = (int)p2_cluster_ctarank();  // 0 = leader (issues the MMAs)
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_tiles = g.tiles_m * g.tiles_n;

  if (threadIdx.x == 0) {
    p2_stamp(g, 0);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);   // the leader's arrive.expect_tx covers BOTH CTAs' bytes (see the producer)
      mbar_init(&empty_bar[s], 1);  // multicast tcgen05.commit
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);   // multicast tcgen05.commit
      mbar_init(&tmem_empty_bar[b], 16);  // 8 epilogue warps x 2 CTAs (used in the leader only)
    }
    fence_barrier_init();
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
  }
  if (warp == 1) tmem_alloc_2sm<Cfg::TMEM_COLS>(tmem_slot);
  umma::fence_before_sync();
  p2_cluster_sync();  // both CTAs' barriers are initialised before any remote arrive / multicast commit
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  if (warp != 0) pdl_wait();  // (the producer waits after its weight prologue)

  if (warp == 0) {
    // ===== TMA producer (both CTAs): my UM/2 rows of A, my BN/2 rows of B =====
    if (umma::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      // The WEIGHTS do not depend on the preceding kernel: fill the ring with the first tile's weight halves while
      // the predecessor drains, only then wait for the activations (programmatic dependent launch).
      int pre = 0;
      p2_stamp(g, 1);
      if (pair < num_tiles) {
        pre = g.num_kb < STAGES ? g.num_kb : STAGES;
        const int b_row0 = (pair / g.tiles_m) * BN + rank * (BN / 2);
        for (int i = 0; i < pre; ++i) {
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[i], 2 * Cfg::STAGE_BYTES);
          tma_load_2d_2sm(smem + i * Cfg::STAGE_BYTES + Cfg::A_BYTES, &tmap_b, &full_bar[i], i * kP2BK, b_row0,
                          UM == 128 ? kEvictFirst : kEvictNormal);
        }
      }
      p2_stamp(g, 2);
      pdl_wait();
      p2_stamp(g, 6);
      for (int t = pair; t < num_tiles; t += num_pairs) {
        const int tm = t % g.tiles_m, tn = t / g.tiles_m;
        const int a_row = tm * UM + rank * Cfg::A_ROWS, b_row = tn * BN + rank * (BN / 2);
        for (int kb = 0; kb < g.num_kb; ++kb) {
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          if (pre > 0) {  // this slot's weights are already on their way
            --pre;
            tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], kb * kP2BK, a_row, UM == 128 ? kEvictLast : kEvictNormal);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
            continue;
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          // Byte counting alone synchronises the pair: the leader expects both CTAs' bytes; the peer only issues
          // its loads (complete_tx on the leader's barrier).  The peer cannot run ahead into the next phase of
          // this slot because its own empty[] barrier is signalled by the MMAs that consumed the current one.
          // (A remote mbarrier.arrive.release.cluster here compiled to MEMBAR.ALL.CTA + ERRBAR per k-block in the
          // one thread that issues the peer's loads: 0.8 us per k-block, 35 % tensor-pipe utilisation -
          // profiles/r02_gemm2sm_ncu.md.)
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
          tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], kb * kP2BK, a_row, UM == 128 ? kEvictLast : kEvictNormal);
          tma_load_2d_2sm(sa + Cfg::A_BYTES, &tmap_b, &full_bar[stage], kb * kP2BK, b_row,
                          UM == 128 ? kEvictFirst : kEvictNormal);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    if (rank == 0 && umma::elect_one()) {
      constexpr uint32_t idesc = umma::make_idesc_f8(UM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
        const int acc = it & 1;
        mbar_wait(&tmem_empty_bar[acc], (((uint32_t)it >> 1) & 1) ^ 1);  // both CTAs' epilogues drained this buffer
        umma::fence_after_sync();
        const uint32_t tacc = tmem_base + acc * Cfg::ACC_COLS;
        for (int kb = 0; kb < g.num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          umma::fence_after_sync();
          if (it == 0 && kb == 0) p2_stamp(g, 3);
          if (it == 0 && kb == STAGES) p2_stamp(g, 7);
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint64_t da = umma::make_desc_k_sw128(sa);
          const uint64_t db = umma::make_desc_k_sw128(sa + Cfg::A_BYTES);
#pragma unroll
          for (int kk = 0; kk < kP2BK / 32; ++kk)
            mma_f8f6f4_2sm(tacc, umma::desc_advance(da, kk * 32), umma::desc_advance(db, kk * 32), idesc,
                           (kb | kk) != 0);
          commit_2sm(&empty_bar[stage]);  // frees the slot in both CTAs
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        commit_2sm(&tmem_full_bar[acc]);  // accumulator complete: both CTAs' epilogues may drain their halves
      }
      p2_stamp(g, 4);
    }
  } else {
    p2_epilogue<UM, BN, Cfg::ACC_COLS, OutT>(g, tmem_base, tmem_full_bar, tmem_empty_bar, warp, lane, rank, pair, num_pairs);
  }
  umma::fence_before_sync();
  p2_cluster_sync_exit();  // the leader's MMAs read the peer's shared memory; multicast arrives target both CTAs
  if (warp == 1) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
  if (threadIdx.x == 0) p2_stamp(g, 10);
}

// (A decode-regime variant with SEPARATE token / weight rings - 4 token slots, 21 weight slots - was built and
// measured slower: the token tile's L2 round trip needs ring depth too.  Removed; see profiles/r02_gemm_decode_2sm.md
// and commit 9f01e8b.)

template <int UM, int BN, typename OutT>
static int launch_p2(const CUtensorMap& ta, const CUtensorMap& tb, const P2Args& g, int num_pairs, cudaStream_t stream) {
  using Cfg = P2Cfg<UM, BN>;
  auto kern = fp8_gemm_2sm_kernel<UM, BN, OutT>;
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc;
  launch_kernel(kern, dim3(2 * num_pairs), dim3(kP2Threads), Cfg::SMEM_BYTES, stream, 2, ta, tb, g);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

static float* g_p2_dump = nullptr;

template <int UM, int BN, typename OutT>
static int run_p2(P2Args g, const void* a, const void* b, int64_t lda, int64_t ldb, cudaStream_t stream) {
  g.tiles_m = (g.M + UM - 1) / UM;
  g.tiles_n = (g.N + BN - 1) / BN;
  g.dump = g_p2_dump;
  g.trace = gemm_debug_trace();
  CUtensorMap ta, tb;
  int rc;
  if ((rc = get_kmajor_tmap(&ta, a, g.M, g.K, lda, UM / 2))) return rc;
  if ((rc = get_kmajor_tmap(&tb, b, g.N, g.K, ldb, BN / 2))) return rc;
  const int tiles = g.tiles_m * g.tiles_n;
  int pairs = sm_count() / 2;
  if (pairs > tiles) pairs = tiles;
  return launch_p2<UM, BN, OutT>(ta, tb, g, pairs, stream);
}

// Returns 1 when the shape is not handled here (the caller falls back), 0 on success, a negative B200_ERR_*
// otherwise.  M <= 256 (decode regime) is only taken when `decode` is set.
template <typename OutT>
int run_fp8_gemm_2sm(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                     const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                     int b_scales_numel, cudaStream_t stream, bool decode) {
  if ((M <= 256 && !decode) || N < 64) return 1;
  P2Args g;
  g.out = out;
  g.a_scales = a_scales;
  g.b_scales = b_scales;
  g.bias = bias;
  g.ldc = ldc;
  g.M = M;
  g.N = N;
  g.K = K;
  g.a_scales_numel = a_scales_numel;
  g.b_scales_numel = b_scales_numel;
  g.tiles_m = 0;
  g.tiles_n = 0;
  g.num_kb = (K + kP2BK - 1) / kP2BK;
  g.dump = nullptr;
  g.trace = nullptr;
  const long pairs = sm_count() / 2;
  static const int force_bn = [] {
    const char* e = getenv("B200_GEMM_2SM_BN");  // tuning aid: 64 (decode only) / 128 / 256
    return (e && *e) ? atoi(e) : 0;
  }();
  if (M <= 256) {
    // Decode regime: the weights stream once.  One M tile (UMMA M = 128 across the pair up to 128 tokens, else 256).
    const long t128 = (N + 127) / 128;
    // 64-wide tiles only when the wide ones would occupy under a quarter of the pairs (measured: o_proj N = 4096
    // 9.07 us with 32 wide tiles against 9.25-9.9 with 64 narrow ones - the narrow stage is 2/3 token bytes)
    const bool narrow = force_bn ? force_bn == 64 : t128 * 4 <= pairs;
    if (M <= 128)
      return narrow ? run_p2<128, 64, OutT>(g, a, b, lda, ldb, stream) : run_p2<128, 128, OutT>(g, a, b, lda, ldb, stream);
    return narrow ? run_p2<256, 64, OutT>(g, a, b, lda, ldb, stream) : run_p2<256, 128, OutT>(g, a, b, lda, ldb, stream);
  }
  // Tile width.  256 x 256 is the balanced tile (a k-block's 32 KB per SM arrive in about the time its four
  // 256x256x32 MMAs take); 256 x 128 halves the MMA work per k-block but still needs 24 KB, so it costs ~0.7 of a
  // wide tile.  It wins when it does not add rounds of the persistent loop (few tiles: M = 512 o/down).
  const long tm = (M + 255) / 256;
  const long t256 = tm * ((N + 255) / 256), t128 = tm * ((N + 127) / 128);
  const double cost256 = (double)((t256 + pairs - 1) / pairs), cost128 = 0.7 * (double)((t128 + pairs - 1) / pairs);
  const bool narrow = force_bn ? force_bn == 128 : (cost128 < cost256 || N < 256);
  if (narrow) return run_p2<256, 128, OutT>(g, a, b, lda, ldb, stream);
  return run_p2<256, 256, OutT>(g, a, b, lda, ldb, stream);
}

template int run_fp8_gemm_2sm<__nv_bfloat16>(void*, const void*, const void*, const float*, const float*, const void*,
                                             int, int, int, int64_t, int64_t, int64_t, int, int, cudaStream_t, bool);
template int run_fp8_gemm_2sm<__half>(void*, const void*, const void*, const float*, const float*, const void*, int, int,
                                      int, int64_t, int64_t, int64_t, int, int, cudaStream_t, bool);
}  // namespace b200

extern "C" void b200_debug_gemm_dump(float* device_buffer) { b200::g_p2_dump = device_buffer; }

namespace b200 {


// ======================================================================================================
// NVFP4 block-scaled GEMM on the same persistent 2-CTA skeleton (M > 256):
//   out = alpha * (A4 . SFA) @ (B4 . SFB)^T, e2m1 operands packed two per byte, e4m3 scale factor per 16 elements.
// Reference: cutlass_scaled_fp4_mm (nvfp4_scaled_mm_kernels.cu:46-137, 2-SM Fp4GemmSm100 configs :214-248).
// Differences from the FP8 kernel above: a k-block is 128 BYTES = 256 elements = four K=64 MMAs
// (kind::mxf4nvf4.block_scale.block16); each stage also carries the scale factors - SFA for THIS CTA's 128 rows
// and SFB for ALL BN weight rows of the pair's tile (every CTA's tensor core multiplies its 128 rows by all BN
// columns, so it needs every column's scales in its own TMEM) - as 512-byte atoms ([32][4][4] e4m3, the tensor-core
// order they already have in global memory), moved smem -> TMEM by tcgen05.cp.cta_group::2 right before the MMA
// that reads them (two TMEM scale buffers, alternating per k-block).  TMEM: BN = 128 -> two accumulators
// (2 x 128 + 2 x 32 scale columns), BN = 256 -> one accumulator (256 + 2 x 48): the epilogue of a 256-wide
// tile is not overlapped, which is why the narrow tile is used when K is short.
constexpr int kF4AtomBytes = 512;

template <int BN> struct P4Cfg {
  static constexpr int YT = (BN + 127) / 128;                 // 128-row scale-factor tiles on the weight side
  static constexpr int A_BYTES = 128 * kP2BK;
  static constexpr int B_BYTES = (BN / 2) * kP2BK;
  static constexpr int SFA_BYTES = 4 * kF4AtomBytes;          // 4 K=64 atoms of this CTA's rows
  static constexpr int SFB_BYTES = YT * 4 * kF4AtomBytes;     // the same for all BN weight rows
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + SFA_BYTES + SFB_BYTES;
  static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > 8 ? 8 : (200 * 1024) / STAGE_BYTES;
  static constexpr int SF_COLS = 16 + 16 * YT;                // per scale buffer
  static constexpr int ACC_STAGES = (2 * BN + 2 * SF_COLS <= 512) ? 2 : 1;
  static constexpr int SF_BASE = ACC_STAGES * BN;
  static constexpr int TMEM_COLS = 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 512;
  static_assert(ACC_STAGES * BN + 2 * SF_COLS <= 512, "TMEM budget");
};

struct P4Args {
  void* out;
  const float* alpha;
  int64_t ldc;
  int M, N, K;
  int tiles_m, tiles_n, num_kb, num_k_atoms;  // ceil(K/256), ceil(K/64)
};

__device__ __forceinline__ void mma_nvf4_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate, uint32_t tmem_sfa, uint32_t tmem_sfb) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf4nvf4.block_scale.block16 [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(
          tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
// scale-factor atom smem -> TMEM in BOTH CTAs of the pair (each from its own shared memory), issued by the leader
__device__ __forceinline__ void cp_sf_2sm(uint32_t tmem_addr, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(tmem_addr), "l"(smem_desc) : "memory");
}
__device__ __forceinline__ uint64_t p4_sf_desc(uint32_t smem_addr) {  // SWIZZLE_NONE, 8-row x 16-byte core matrices
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(128 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

template <int BN, typename OutT>
__global__ void __launch_bounds__(kP2Threads, 1)
    nvfp4_gemm_2sm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                          const __grid_constant__ CUtensorMap tmap_sfa, const __grid_constant__ CUtensorMap tmap_sfb,
                          const P4Args g) {
  using Cfg = P4Cfg<BN>;
  constexpr int STAGES = Cfg::STAGES, NACC = Cfg::ACC_STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = (int)p2_cluster_ctarank();
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_tiles = g.tiles_m * g.tiles_n;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 16);
    }
    fence_barrier_init();
    prefetch_tensormap(&tmap_a);
    prefetch_tensormap(&tmap_b);
    prefetch_tensormap(&tmap_sfa);
    prefetch_tensormap(&tmap_sfb);
  }
  if (warp == 1) tmem_alloc_2sm<Cfg::TMEM_COLS>(tmem_slot);
  umma::fence_before_sync();
  p2_cluster_sync();
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  if (warp == 0) {
    if (umma::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += num_pairs) {
        const int tm = t % g.tiles_m, tn = t / g.tiles_m;
        const int a_row = tm * 256 + rank * 128, b_row = tn * BN + rank * (BN / 2);
        const int sfa_atom0 = (tm * 2 + rank) * g.num_k_atoms;  // scale-factor atoms are [row tile][k atom]
        for (int kb = 0; kb < g.num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* ssfa = sa + Cfg::A_BYTES + Cfg::B_BYTES;
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
          tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], kb * kP2BK, a_row, kEvictNormal);
          tma_load_2d_2sm(sa + Cfg::A_BYTES, &tmap_b, &full_bar[stage], kb * kP2BK, b_row, kEvictNormal);
          tma_load_2d_2sm(ssfa, &tmap_sfa, &full_bar[stage], 0, sfa_atom0 + 4 * kb, kEvictNormal);
#pragma unroll
          for (int y = 0; y < Cfg::YT; ++y)
            tma_load_2d_2sm(ssfa + Cfg::SFA_BYTES + y * 4 * kF4AtomBytes, &tmap_sfb, &full_bar[stage], 0,
                            (tn * Cfg::YT + y) * g.num_k_atoms + 4 * kb, kEvictNormal);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && umma::elect_one()) {
      constexpr uint32_t idesc = umma::make_idesc_nvf4(256, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      uint32_t kbg = 0;  // running k-block count: parity selects the TMEM scale buffer
      for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
        const int acc = NACC == 2 ? (it & 1) : 0;
        const uint32_t use = NACC == 2 ? ((uint32_t)it >> 1) : (uint32_t)it;
        mbar_wait(&tmem_empty_bar[acc], (use & 1) ^ 1);
        umma::fence_after_sync();
        const uint32_t tacc = tmem_base + acc * BN;
        for (int kb = 0; kb < g.num_kb; ++kb, ++kbg) {
          mbar_wait(&full_bar[stage], phase);
          umma::fence_after_sync();
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + Cfg::A_BYTES, ssfa = sb + Cfg::B_BYTES, ssfb = ssfa + Cfg::SFA_BYTES;
          const uint64_t da = umma::make_desc_k_sw128(sa);
          const uint64_t db = umma::make_desc_k_sw128(sb);
          const uint32_t t_sfa = tmem_base + Cfg::SF_BASE + (kbg & 1) * Cfg::SF_COLS;
          const uint32_t t_sfb = t_sfa + 16;
          const int atoms = min(4, g.num_k_atoms - 4 * kb);
          for (int kk = 0; kk < atoms; ++kk) {
            cp_sf_2sm(t_sfa + 4 * kk, p4_sf_desc(ssfa + kk * kF4AtomBytes));
#pragma unroll
            for (int y = 0; y < Cfg::YT; ++y)
              cp_sf_2sm(t_sfb + 4 * Cfg::YT * kk + 4 * y, p4_sf_desc(ssfb + (y * 4 + kk) * kF4AtomBytes));
            mma_nvf4_2sm(tacc, umma::desc_advance(da, kk * 32), umma::desc_advance(db, kk * 32), idesc, (kb | kk) != 0,
                         t_sfa + 4 * kk, t_sfb + 4 * Cfg::YT * kk);
          }
          commit_2sm(&empty_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        commit_2sm(&tmem_full_bar[acc]);
      }
    }
  } else {
    const int q = warp & 3;
    OutT* outp = reinterpret_cast<OutT*>(g.out);
    const bool vec_ok = ((g.ldc * (int64_t)sizeof(OutT)) % 32 == 0) && (reinterpret_cast<uintptr_t>(g.out) % 32 == 0);
    const float alpha = *g.alpha;
    int it = 0;
    for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
      const int tm = t % g.tiles_m, tn = t / g.tiles_m;
      const int acc = NACC == 2 ? (it & 1) : 0;
      const uint32_t use = NACC == 2 ? ((uint32_t)it >> 1) : (uint32_t)it;
      const int m = tm * 256 + rank * 128 + q * 32 + lane;
      const int n0 = tn * BN;
      mbar_wait(&tmem_full_bar[acc], use & 1);
      umma::fence_after_sync();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
      uint32_t rn[16];
      const int c_lo = ((warp - 2) >> 2) * (BN / 2), c_hi = c_lo + BN / 2;
      umma::ld_32x32b_x16(tacc + c_lo, rn);
#pragma unroll 1
      for (int c = c_lo; c < c_hi; c += 16) {
        umma::ld_wait_dep(rn);
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rn[j]) * alpha;
        if (c + 16 < c_hi) umma::ld_32x32b_x16(tacc + c + 16, rn);
        const int n = n0 + c;
        if (m < g.M && n < g.N) {
          OutT* dst = outp + (int64_t)m * g.ldc + n;
          if (vec_ok && n + 16 <= g.N) {
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = p2_pack2<OutT>(v[2 * i], v[2 * i + 1]);
            st_global_256(dst, w);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n + j < g.N) dst[j] = from_float<OutT>(v[j]);
          }
        }
      }
      umma::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty_bar[acc]);
    }
  }
  umma::fence_before_sync();
  p2_cluster_sync();
  if (warp == 1) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
}

// scale-factor array as a 2-D tensor: one 512-byte atom per row (128 x u32), box = 4 consecutive atoms
static int get_sf_tmap(CUtensorMap* out, const void* base, int64_t num_atoms) {
  uint64_t dims[2] = {128, (uint64_t)num_atoms};
  uint64_t strides[1] = {kF4AtomBytes};
  uint32_t box[2] = {128, 4};
  return encode_tmap(out, kTmaF32, 2, base, dims, strides, box, kSwizzleNone);
}

template <int BN, typename OutT>
static int run_p4(P4Args g, const void* a, const void* b, const void* sfa, const void* sfb, cudaStream_t stream) {
  using Cfg = P4Cfg<BN>;
  g.tiles_n = (g.N + BN - 1) / BN;
  const int kbytes = g.K / 2;
  CUtensorMap ta, tb, tsa, tsb;
  int rc;
  if ((rc = get_kmajor_tmap(&ta, a, g.M, kbytes, kbytes, 128))) return rc;
  if ((rc = get_kmajor_tmap(&tb, b, g.N, kbytes, kbytes, BN / 2))) return rc;
  if ((rc = get_sf_tmap(&tsa, sfa, (int64_t)((g.M + 127) / 128) * g.num_k_atoms))) return rc;
  if ((rc = get_sf_tmap(&tsb, sfb, (int64_t)((g.N + 127) / 128) * g.num_k_atoms))) return rc;
  auto kern = nvfp4_gemm_2sm_kernel<BN, OutT>;
  if ((rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES))) return rc;
  const int tiles = g.tiles_m * g.tiles_n;
  int pairs = sm_count() / 2;
  if (pairs > tiles) pairs = tiles;
  launch_kernel(kern, dim3(2 * pairs), dim3(kP2Threads), Cfg::SMEM_BYTES, stream, 2, ta, tb, tsa, tsb, g);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

// Returns 1 when the shape is not handled here (the caller uses nvfp4_gemm.cu), 0 on success, < 0 on error.
template <typename OutT>
int run_nvfp4_gemm_2sm(void* out, const void* a, const void* b, const void* sfa, const void* sfb, const float* alpha,
                       int M, int N, int K, int64_t ldc, cudaStream_t stream, int min_m) {
  if (M <= min_m || N < 128 || K % 32) return 1;
  P4Args g;
  g.out = out;
  g.alpha = alpha;
  g.ldc = ldc;
  g.M = M;
  g.N = N;
  g.K = K;
  g.tiles_m = (M + 255) / 256;
  g.tiles_n = 0;
  g.num_kb = (K / 2 + kP2BK - 1) / kP2BK;
  g.num_k_atoms = (K + 63) / 64;
  // 256-wide tiles have the better MMA : shared-memory-fill balance but a single TMEM accumulator (exposed
  // epilogue); the 128-wide tile double-buffers.  Narrow when K is short (epilogue-dominated) or tiles are few.
  const long pairs = sm_count() / 2;
  const long t256 = (long)g.tiles_m * ((N + 255) / 256), t128 = (long)g.tiles_m * ((N + 127) / 128);
  static const int force_bn = [] {
    const char* e = getenv("B200_F4_2SM_BN");  // tuning aid: 128 / 256
    return (e && *e) ? atoi(e) : 0;
  }();
  const bool narrow = force_bn ? force_bn == 128
                               : (g.num_kb <= 8 || N < 256 || 0.7 * (double)((t128 + pairs - 1) / pairs) < (double)((t256 + pairs - 1) / pairs));
  if (narrow) return run_p4<128, OutT>(g, a, b, sfa, sfb, stream);
  return run_p4<256, OutT>(g, a, b, sfa, sfb, stream);
}

template int run_nvfp4_gemm_2sm<__nv_bfloat16>(void*, const void*, const void*, const void*, const void*, const float*,
                                               int, int, int, int64_t, cudaStream_t, int);
template int run_nvfp4_gemm_2sm<__half>(void*, const void*, const void*, const void*, const void*, const float*, int, int,
                                        int, int64_t, cudaStream_t, int);

}  // namespace b200
