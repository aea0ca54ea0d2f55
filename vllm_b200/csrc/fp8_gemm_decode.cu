// FP8 W8A8 GEMM, decode regime (M <= 256 tokens): weight-bandwidth bound.
// Reference semantics: cutlass_scaled_mm (scaled_mm_entry.cu:197-270); the reference's sm100
// small-M configs are c3x/scaled_mm_sm100_fp8_dispatch.cuh:100-194 (swap-AB tiles, no K split).
//
// B200 design
//  * swap-AB: 128 weight rows per tile on the UMMA-M axis, the BN = 16..256 (padded) tokens on UMMA-N;
//    fp32 accumulators in TMEM, two buffers so an epilogue overlaps the next segment's stream.
//  * CTA pairs: two CTAs of a cluster work on ADJACENT weight tiles over the SAME k-blocks; each TMA-loads
//    half of the token tile and MULTICASTS it to both, halving the token operand's TMA requests.
//  * as many SMs as possible have to stream: one SM's mainloop ingests ~125 GB/s (32 KB per 0.26 us k-block,
//    %globaltimer trace profiles/r02_gemm_decode_trace.md; with ALL SMs pulling the aggregate L2->SM limit of
//    ~8 TB/s shows as ~57 GB/s per SM, scripts/tma_bench.cu), so HBM speed needs >= ~55-60 streaming SMs and
//    the fixed costs (start-up, split-K fix-up, exit) are what separates this kernel from the roofline.
//      mode 3 "cluster split-K" (default when tiles are few): S <= 4 splits x 2 tiles form ONE cluster of
//              2S CTAs; split s takes k-blocks [KB*s/S, KB*(s+1)/S).  Fix-up is a reduce-scatter over
//              distributed shared memory: each CTA parks the token chunks its peers will finish in its own
//              (idle) pipeline ring, signals their mbarriers, and PULLS its own chunks from the peers with
//              ld.shared::cluster.  No workspace, no global flags, no memory fences; the hardware
//              co-schedules the cluster so there is no residency assumption.
//      mode 0 whole super-tiles per pair (default when there are more super-tiles than pairs, or when a cluster
//              of 2S CTAs cannot be co-scheduled).
//    Both are deterministic: partial sums are added in a fixed order, no atomics on data.  (A split-K and a
//    stream-K partition through a global workspace with release/acquire flags were measured slower on every
//    Llama shape in round 1 - 3 us fix-up tail - and removed in round 2.)
//  * epilogue: 8 warps (two per TMEM lane quarter, alternate 16-token chunks), TMEM loads software-
//    pipelined one chunk ahead, token pairs packed to 16-bit x2 and transposed through shared memory so
//    a warp stores 256 contiguous bytes of an output row.
#include <cuda.h>

#include <cstdlib>

#include "umma.cuh"

namespace b200 {

int get_kmajor_tmap(CUtensorMap* out, const void* base, int rows, int k_bytes, int64_t ld_bytes, int box_rows);

constexpr int kDecThreads = 320;  // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)
constexpr int kDBK = 128;  // bytes of K per stage = one swizzle span

template <int BN> struct DecCfg {
  static constexpr int X_BYTES = 128 * kDBK;  // weights
  static constexpr int Y_BYTES = BN * kDBK;   // tokens (both halves)
  static constexpr int STAGE_BYTES = X_BYTES + Y_BYTES;
  static constexpr int STAGES = (196 * 1024) / STAGE_BYTES > 10 ? 10 : (196 * 1024) / STAGE_BYTES;
  static constexpr int ACC_COLS = BN < 32 ? 32 : BN;
  static constexpr int TMEM_COLS = 2 * ACC_COLS;
  static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  static constexpr int OUT_STAGE_BYTES = 2 * 2 * 8 * 128 * 4;  // per epilogue half: two [8 token pairs][128 rows] tiles
  static constexpr int SCALE_BYTES = BN * 4;                   // per-token scales
  static constexpr int SMEM_BYTES = RING_BYTES + OUT_STAGE_BYTES + SCALE_BYTES + 1024 /*align*/ + 512 /*barriers*/;
};

struct DecArgs {
  void* out;
  const float* a_scales;  // per token (M) or [1]
  const float* b_scales;  // per channel (N) or [1]
  const void* bias;       // [N] out dtype or null
  int64_t ldc;
  int M, N, K;
  int a_scales_numel, b_scales_numel;
  int num_tiles, num_super, num_kb;  // ceil(N/128), ceil(tiles/2), ceil(K/128)
  int mode, splits;                  // see header comment
  unsigned long long* trace;         // debug: [gridDim.x][16] %globaltimer stamps, or null
};

// phase stamps for scripts/prof_gemm.py --trace (one thread per CTA writes each slot)
enum { TR_ENTRY, TR_SETUP, TR_FILLED, TR_FIRST_FULL, TR_LAST_COMMIT, TR_ACC_READY, TR_PUBLISHED, TR_GATHERED,
       TR_OUT_DONE, TR_EPI_END, TR_EXIT };
__device__ __forceinline__ void trace_stamp(const DecArgs& g, int slot) {
  if (g.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                               uint16_t mask, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster.L2::cache_hint"
      " [%0], [%1, {%4, %5}], [%2], %3, %6;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "h"(mask), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void fence_acq_rel_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
  return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(map_to_cta(smem_u32(bar), cta))
               : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  unsigned long long t0 = 0;  // wall-time bound, see mbar_wait (common.cuh)
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (!ok && (++spins & 0xFFFFFu) == 0) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kMbarTimeoutNs) {
        printf("b200: cluster mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
        __trap();
      }
    }
  }
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t cluster_addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(cluster_addr)
               : "memory");
  return v;
}
template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  const __half2 t = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&t);
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// unit range of cluster `cid` (units are super-tile-major: u = st * KB + kb)
__device__ __forceinline__ void cluster_range(const DecArgs& g, int cid, int split, int NC, long long& u0,
                                              long long& u1) {
  const long long KB = g.num_kb;
  if (g.mode == 3) {
    u0 = cid * KB + KB * split / g.splits;
    u1 = cid * KB + KB * (split + 1) / g.splits;
  } else {
    u0 = ((long long)g.num_super * cid / NC) * KB;
    u1 = ((long long)g.num_super * (cid + 1) / NC) * KB;
  }
}

template <int BN, typename OutT>
__global__ void __launch_bounds__(kDecThreads, 1)
    fp8_gemm_decode_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_a,
                           const DecArgs g) {
  using Cfg = DecCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* out_stage = smem + Cfg::RING_BYTES;
  float* s_as = reinterpret_cast<float*>(out_stage + Cfg::OUT_STAGE_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES + Cfg::OUT_STAGE_BYTES + Cfg::SCALE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;  // [2]
  uint64_t* fix_bar = tmem_empty_bar + 2;        // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(fix_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int crank = (int)cluster_ctarank();
  const int rank = crank & 1, split = crank >> 1;  // CTA pair member / split of a cluster split-K (mode 3)
  const int csize = g.mode == 3 ? 2 * g.splits : 2;
  const int NC = gridDim.x / csize, cid = blockIdx.x / csize;
  const uint16_t pair_mask = (uint16_t)(3u << (2 * split));
  long long u0, u1;
  cluster_range(g, cid, split, NC, u0, u1);
  const int KB = g.num_kb;

  pdl_trigger();
  if (threadIdx.x == 0) {
    trace_stamp(g, TR_ENTRY);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 2);  // both CTAs of the pair must have consumed the slot
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 8);  // one arrive per epilogue warp
    }
    mbar_init(fix_bar, g.mode == 3 ? g.splits - 1 : 1);  // peers of a cluster split-K
    fence_barrier_init();
    prefetch_tensormap(&tmap_w);
    prefetch_tensormap(&tmap_a);
  }
  if (warp == 1) umma::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  umma::fence_before_sync();
  cluster_sync_all();  // barriers of BOTH CTAs are initialised before any multicast / remote arrive
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (umma::elect_one()) {
      trace_stamp(g, TR_SETUP);
      auto load_w = [&](long long u, int stage) {
        const int st = (int)(u / KB), kb = (int)(u - (long long)st * KB);
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
        tma_load_2d(smem + stage * Cfg::STAGE_BYTES, &tmap_w, &full_bar[stage], kb * kDBK, (2 * st + rank) * 128,
                    kEvictFirst);  // streamed once
      };
      auto load_a = [&](long long u, int stage) {
        const int kb = (int)(u % KB);
        // my half of the token tile, delivered to both CTAs of the pair
        tma_load_2d_mc(smem + stage * Cfg::STAGE_BYTES + Cfg::X_BYTES + rank * (Cfg::Y_BYTES / 2), &tmap_a,
                       &full_bar[stage], kb * kDBK, rank * (BN / 2), pair_mask, kEvictLast);
      };
      // The weights do not depend on the preceding kernel: fill the ring with them while it drains, and
      // only then wait for the activations (programmatic dependent launch).
      const int pre = (int)min((long long)STAGES, u1 - u0);
      for (int i = 0; i < pre; ++i) load_w(u0 + i, i);
      pdl_wait();
      for (int i = 0; i < pre; ++i) load_a(u0 + i, i);
      int stage = pre == STAGES ? 0 : pre;
      uint32_t phase = pre == STAGES ? 1 : 0;
      for (long long u = u0 + pre; u < u1; ++u) {
        if (u == u0 + STAGES) trace_stamp(g, TR_FILLED);
        mbar_wait(&empty_bar[stage], phase ^ 1);
        load_w(u, stage);
        load_a(u, stage);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (umma::elect_one()) {
      constexpr uint32_t idesc = umma::make_idesc_f8(128, BN);
      int stage = 0;
      uint32_t phase = 0;
      int seg = 0;
      for (long long u = u0; u < u1; ++seg) {
        const int kb0 = (int)(u % KB);
        const int nkb = (int)min((long long)(KB - kb0), u1 - u);
        const int buf = seg & 1;
        mbar_wait(&tmem_empty_bar[buf], (((uint32_t)seg >> 1) & 1) ^ 1);
        umma::fence_after_sync();
        const uint32_t tacc = tmem_base + buf * Cfg::ACC_COLS;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          umma::fence_after_sync();
          if (seg == 0 && kb == 0) trace_stamp(g, TR_FIRST_FULL);
          const uint32_t sx = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint64_t dx = umma::make_desc_k_sw128(sx);
          const uint64_t dy = umma::make_desc_k_sw128(sx + Cfg::X_BYTES);
#pragma unroll
          for (int kk = 0; kk < kDBK / 32; ++kk)
            umma::mma_f8f6f4(tacc, umma::desc_advance(dx, kk * 32), umma::desc_advance(dy, kk * 32), idesc,
                             (kb | kk) != 0);
          commit_mc(&empty_bar[stage], pair_mask);  // frees the slot in both CTAs of the pair
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma::commit(&tmem_full_bar[buf]);
        u += nkb;
      }
      trace_stamp(g, TR_LAST_COMMIT);
    }
  } else {
    // ===== epilogue warps: warp w owns TMEM lanes 32*(w%4) .. +31 (= weight rows of the tile); the two warps
    // of a lane quarter take alternate 16-token chunks.  One warp per scheduler is instruction-latency bound
    // (~600 cycles per chunk measured), so the epilogue gets two.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int et = q * 32 + lane;    // weight row within the tile
    const int eid = half * 128 + et;  // epilogue thread id 0..255
    const bool lead = eid == 0;
    OutT* outp = reinterpret_cast<OutT*>(g.out);
    const OutT* bias = reinterpret_cast<const OutT*>(g.bias);
    const int m_cols = min(BN, (g.M + 15) & ~15);  // token columns that can be non-zero (multiple of 16)
    const int nch = m_cols >> 4;
    uint32_t fix_phase = 0;
    pdl_wait();  // scales, bias, workspace and the output buffer belong to the stream order from here
    for (int j = eid; j < BN; j += 256) s_as[j] = j < g.M ? g.a_scales[g.a_scales_numel > 1 ? j : 0] : 0.f;
    epi_bar();
    // Scale, convert and store one finished chunk (16 tokens x this thread's weight row).  One warp per
    // scheduler makes the epilogue instruction-latency bound, so the instruction count per output matters:
    // token PAIRS are packed to bf16x2/half2 words and transposed through shared memory ([8 pairs][128 rows]),
    // then every thread splits one 16-byte piece into the two tokens' 8-byte row segments (a warp stores
    // 256 contiguous bytes of one output row).  ~75 instructions per chunk instead of ~130 for direct stores.
    const bool vec_out = sizeof(OutT) == 2 && ((g.ldc * (int64_t)sizeof(OutT)) % 8 == 0) &&
                         (reinterpret_cast<uintptr_t>(g.out) % 8 == 0);
    int out_par = 0;
    uint32_t* stage_half = reinterpret_cast<uint32_t*>(out_stage) + half * (2 * 8 * 128);
    auto half_bar = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory"); };
    auto store_chunk = [&](int ch, int tile, float (&v)[16], float sx_row, float bias_row) {
      float as[16];
#pragma unroll
      for (int j4 = 0; j4 < 4; ++j4) {
        const float4 t = *reinterpret_cast<const float4*>(s_as + ch * 16 + j4 * 4);
        as[j4 * 4 + 0] = t.x, as[j4 * 4 + 1] = t.y, as[j4 * 4 + 2] = t.z, as[j4 * 4 + 3] = t.w;
      }
      if (vec_out) {
        uint32_t* stg = stage_half + out_par * (8 * 128);
#pragma unroll
        for (int jp = 0; jp < 8; ++jp)
          stg[jp * 128 + et] = pack2<OutT>(v[2 * jp] * as[2 * jp] * sx_row + bias_row,
                                           v[2 * jp + 1] * as[2 * jp + 1] * sx_row + bias_row);
        half_bar();
        const int n0 = tile * 128 + lane * 4;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int jp = q + 4 * h;
          const uint4 w = *reinterpret_cast<const uint4*>(stg + jp * 128 + lane * 4);
          // w.{x,y,z,w} = rows n0..n0+3, each (token 2jp | token 2jp+1 << 16)
          uint2 lo, hi;
          lo.x = __byte_perm(w.x, w.y, 0x5410), lo.y = __byte_perm(w.z, w.w, 0x5410);
          hi.x = __byte_perm(w.x, w.y, 0x7632), hi.y = __byte_perm(w.z, w.w, 0x7632);
          const int m = ch * 16 + 2 * jp;
          if (n0 < g.N) {
            OutT* d0 = outp + (int64_t)m * g.ldc + n0;
            OutT* d1 = d0 + g.ldc;
            if (n0 + 4 <= g.N) {
              if (m < g.M) *reinterpret_cast<uint2*>(d0) = lo;
              if (m + 1 < g.M) *reinterpret_cast<uint2*>(d1) = hi;
            } else {
              const OutT* l = reinterpret_cast<const OutT*>(&lo);
              const OutT* u = reinterpret_cast<const OutT*>(&hi);
              for (int e = 0; e < 4 && n0 + e < g.N; ++e) {
                if (m < g.M) d0[e] = l[e];
                if (m + 1 < g.M) d1[e] = u[e];
              }
            }
          }
        }
        out_par ^= 1;
      } else if (tile * 128 + et < g.N) {
        OutT* dst = outp + (int64_t)(ch * 16) * g.ldc + tile * 128 + et;
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (ch * 16 + j < g.M) dst[(int64_t)j * g.ldc] = from_float<OutT>(v[j] * as[j] * sx_row + bias_row);
      }
    };

    int seg = 0;
    for (long long u = u0; u < u1; ++seg) {
      const int st = (int)(u / KB), kb0 = (int)(u - (long long)st * KB);
      const int nkb = (int)min((long long)(KB - kb0), u1 - u);
      const int buf = seg & 1;
      const int tile = 2 * st + rank;
      const bool tile_valid = tile < g.num_tiles;
      const int n = tile * 128 + et;
      // per-row scale / bias: issued before the wait so the (cold) loads overlap the weight stream
      float sx_row = 1.f, bias_row = 0.f;
      if (tile_valid && n < g.N) {
        sx_row = g.b_scales[g.b_scales_numel > 1 ? n : 0];
        if (bias) bias_row = to_float(bias[n]);
      }
      mbar_wait(&tmem_full_bar[buf], ((uint32_t)seg >> 1) & 1);
      umma::fence_after_sync();
      if (lead && u + nkb >= u1) trace_stamp(g, TR_ACC_READY);  // last segment
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + buf * Cfg::ACC_COLS;

      if (g.mode == 3) {
        // ---- cluster split-K: reduce-scatter over distributed shared memory (one segment per CTA).
        // The ring is idle (all of this CTA's MMAs have completed): park the chunks the peers will finish
        // there, tell the peers, then PULL the peers' copies of this CTA's own chunks with remote loads.
        const int S = g.splits;
        const int own_lo = nch * split / S, own_hi = nch * (split + 1) / S;
        float4* park = reinterpret_cast<float4*>(smem);  // [chunk][j/4][128 rows] float4
        if (tile_valid) {
          const int n_park = nch - (own_hi - own_lo);  // chunks other than own, visited in index order
          auto park_chunk = [&](int i) { return i < own_lo ? i : i + (own_hi - own_lo); };
          uint32_t rn[16];
          if (half < n_park) umma::ld_32x32b_x16(tacc + park_chunk(half) * 16, rn);
#pragma unroll 1
          for (int i = half; i < n_park; i += 2) {
            const int ch = park_chunk(i);
            umma::ld_wait_dep(rn);
            float4 t[4];
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4)
              t[j4] = make_float4(__uint_as_float(rn[j4 * 4]), __uint_as_float(rn[j4 * 4 + 1]),
                                  __uint_as_float(rn[j4 * 4 + 2]), __uint_as_float(rn[j4 * 4 + 3]));
            if (i + 2 < n_park) umma::ld_32x32b_x16(tacc + park_chunk(i + 2) * 16, rn);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) park[(ch * 4 + j4) * 128 + et] = t[j4];
          }
        }
        epi_bar();
        if (eid < S && eid != split)  // .release.cluster arrive after the CTA barrier: publishes every thread's parks
          mbar_arrive_remote(fix_bar, (uint32_t)(2 * eid + rank));
        if (lead) trace_stamp(g, TR_PUBLISHED);
        mbar_wait_cluster(fix_bar, fix_phase);
        fix_phase ^= 1;
        if (lead) trace_stamp(g, TR_GATHERED);
        if (tile_valid) {
          // software pipeline: chunk ch+1's TMEM load and remote (DSMEM) loads fly while chunk ch is stored
          uint32_t rn[16];
          float4 pn[3][4];  // S <= 4
          auto fetch = [&](int ch) {
            umma::ld_32x32b_x16(tacc + ch * 16, rn);
#pragma unroll
            for (int p = 0; p < 3; ++p) {
              const int sp = p + (p >= split);  // peers in increasing split order, skipping self
              if (sp < S) {
                const uint32_t base = map_to_cta(smem_u32(park + (ch * 4) * 128 + et), (uint32_t)(2 * sp + rank));
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4) pn[p][j4] = ld_dsmem_f4(base + j4 * 128 * 16);
              }
            }
          };
          if (own_lo + half < own_hi) fetch(own_lo + half);
#pragma unroll 1
          for (int ch = own_lo + half; ch < own_hi; ch += 2) {
            umma::ld_wait_dep(rn);
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rn[j]);
#pragma unroll
            for (int p = 0; p < 3; ++p) {
              const int sp = p + (p >= split);
              if (sp < S) {
#pragma unroll
                for (int j4 = 0; j4 < 4; ++j4) {
                  v[j4 * 4 + 0] += pn[p][j4].x, v[j4 * 4 + 1] += pn[p][j4].y;
                  v[j4 * 4 + 2] += pn[p][j4].z, v[j4 * 4 + 3] += pn[p][j4].w;
                }
              }
            }
            if (ch + 2 < own_hi) fetch(ch + 2);
            store_chunk(ch, tile, v, sx_row, bias_row);
          }
        }
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        u += nkb;
        if (lead) trace_stamp(g, TR_OUT_DONE);
        continue;
      }

      // ---- whole tiles (mode 0): scale, convert and store every chunk; the TMEM load of chunk ch+1 is in flight
      // while chunk ch is stored
      if (tile_valid) {
        uint32_t rn[16];
        if (half < nch) umma::ld_32x32b_x16(tacc + half * 16, rn);
#pragma unroll 1
        for (int ch = half; ch < nch; ch += 2) {
          umma::ld_wait_dep(rn);
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rn[j]);
          if (ch + 2 < nch) umma::ld_32x32b_x16(tacc + (ch + 2) * 16, rn);
          store_chunk(ch, tile, v, sx_row, bias_row);
        }
      }
      umma::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
      u += nkb;
      if (lead && u >= u1) trace_stamp(g, TR_OUT_DONE);
    }
    if (lead) trace_stamp(g, TR_EPI_END);
  }
  umma::fence_before_sync();
  cluster_sync_all();  // the peer may still multicast into / arrive on this CTA's shared memory
  if (threadIdx.x == 0) trace_stamp(g, TR_EXIT);
  if (warp == 1) umma::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

template <int BN, typename OutT>
static int launch_fp8_decode(const CUtensorMap& tw, const CUtensorMap& ta, const DecArgs& d, int num_clusters,
                             int cluster_size, cudaStream_t stream, int* max_active_out = nullptr) {
  using Cfg = DecCfg<BN>;
  auto kern = fp8_gemm_decode_kernel<BN, OutT>;
  if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc_attr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster_size * num_clusters, 1, 1);
  cfg.blockDim = dim3(kDecThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster_size;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (pdl_enabled() && !max_active_out) {
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.numAttrs = 2;
  }
  if (max_active_out) {  // query only: how many clusters of this size are co-resident
    int n = 0;
    cudaError_t qe = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    if (qe != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
    *max_active_out = n;
    return 0;
  }
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tw, ta, d);
  if (e != cudaSuccess) {
    set_error("fp8 decode gemm launch failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  count_launch();
  return 0;
}

static unsigned long long* g_gemm_trace = nullptr;
static int g_gemm_mode = -1;  // b200_debug_gemm_mode
int gemm_debug_mode() { return g_gemm_mode; }
unsigned long long* gemm_debug_trace() { return g_gemm_trace; }

static int decode_bn(int M) { return M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256))); }

template <typename OutT>
int run_fp8_decode(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                   const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                   int b_scales_numel, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  const int bn = decode_bn(M);
  CUtensorMap tw, ta;
  int rc;
  if ((rc = get_kmajor_tmap(&tw, b, N, K, ldb, 128))) return rc;
  if ((rc = get_kmajor_tmap(&ta, a, M, K, lda, bn / 2))) return rc;  // each CTA of a pair loads half the tokens
  DecArgs d;
  d.out = out;
  d.a_scales = a_scales;
  d.b_scales = b_scales;
  d.bias = bias;
  d.ldc = ldc;
  d.M = M;
  d.N = N;
  d.K = K;
  d.a_scales_numel = a_scales_numel;
  d.b_scales_numel = b_scales_numel;
  d.num_tiles = (N + 127) / 128;
  d.num_super = (d.num_tiles + 1) / 2;
  d.num_kb = (K + kDBK - 1) / kDBK;
  const int max_clusters = sm_count() / 2;
  (void)workspace;  // the decode regime needs no scratch: the split-K fix-up runs over distributed shared memory
  (void)workspace_bytes;
  int NC;
  d.mode = 0;
  d.splits = 1;
  static const int env_mode = [] {
    const char* e = getenv("B200_GEMM_MODE");  // tuning aid: 0 whole tiles / 3 cluster split-K
    return (e && *e) ? atoi(e) : -1;
  }();
  const int force_mode = g_gemm_mode >= 0 ? g_gemm_mode : env_mode;
  d.trace = g_gemm_trace;
  // Cluster split-K (mode 3) needs no workspace: S splits x 2 tiles form one cluster of 2S <= 8 CTAs.
  auto dispatch = [&](int nc, int csize, int* query) -> int {
    switch (bn) {
      case 16: return launch_fp8_decode<16, OutT>(tw, ta, d, nc, csize, stream, query);
      case 32: return launch_fp8_decode<32, OutT>(tw, ta, d, nc, csize, stream, query);
      case 64: return launch_fp8_decode<64, OutT>(tw, ta, d, nc, csize, stream, query);
      case 128: return launch_fp8_decode<128, OutT>(tw, ta, d, nc, csize, stream, query);
      default: return launch_fp8_decode<256, OutT>(tw, ta, d, nc, csize, stream, query);
    }
  };
  // co-resident clusters per cluster size, per token-tile width (queried once)
  static int max_active[5][5] = {};  // [bn index][S] ; 0 = not yet queried, -1 = unsupported
  const int bi = bn == 16 ? 0 : bn == 32 ? 1 : bn == 64 ? 2 : bn == 128 ? 3 : 4;
  auto active_clusters = [&](int S) -> int {
    if (max_active[bi][S] == 0) {
      int n = 0;
      if ((rc = dispatch(1, 2 * S, &n))) return 0;
      max_active[bi][S] = n > 0 ? n : -1;
    }
    return max_active[bi][S] > 0 ? max_active[bi][S] : 0;
  };
  int S3 = 0;
  if (force_mode < 0 || force_mode == 3) {
    static const int s_max = [] {
      const char* e = getenv("B200_GEMM_SMAX");  // tuning aid: largest cluster split
      return (e && *e) ? atoi(e) : 4;
    }();
    for (int S = s_max; S >= 2; --S) {
      if (d.num_kb / S < 2) continue;  // >= 2 k-blocks per split
      if (active_clusters(S) >= d.num_super) {
        S3 = S;
        break;
      }
    }
  }
  if (S3 >= 2) {
    d.mode = 3;
    d.splits = S3;
    return dispatch(d.num_super, 2 * S3, nullptr);
  }
  // whole super-tiles per pair (mode 0): at most one cluster per TPC
  NC = d.num_super < max_clusters ? d.num_super : max_clusters;
  return dispatch(NC, 2, nullptr);
}

template int run_fp8_decode<__nv_bfloat16>(void*, const void*, const void*, const float*, const float*, const void*,
                                           int, int, int, int64_t, int64_t, int64_t, int, int, void*, size_t,
                                           cudaStream_t);
template int run_fp8_decode<__half>(void*, const void*, const void*, const float*, const float*, const void*, int, int,
                                    int, int64_t, int64_t, int64_t, int, int, void*, size_t, cudaStream_t);

}  // namespace b200

extern "C" size_t b200_scaled_mm_workspace_bytes(int M, int N) {
  // No kernel of b200_scaled_mm_fp8 needs scratch any more (the decode regime's split-K fix-up is a reduce-scatter
  // over distributed shared memory; the workspace-based split-K / stream-K partitions were removed: slower on every
  // measured shape).  Kept in the ABI so callers that size a workspace keep working.
  (void)M;
  (void)N;
  return 0;
}

extern "C" void b200_debug_gemm_trace(unsigned long long* device_buffer) { b200::g_gemm_trace = device_buffer; }
extern "C" void b200_debug_gemm_mode(int mode) { b200::g_gemm_mode = mode; }
