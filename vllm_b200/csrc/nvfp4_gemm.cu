// NVFP4 block-scaled GEMM on tcgen05:  out = alpha * (A4 . SFA) @ (B4 . SFB)^T.
// Reference semantics: cutlass_scaled_fp4_mm, csrc/libtorch_stable/quantization/fp4/
// nvfp4_scaled_mm_entry.cu:41-69, nvfp4_scaled_mm_kernels.cu:46-137 (Fp4GemmSm100 configs),
// :214-248 (M dispatch), :279-347 (shape checks); SF layout nvfp4_utils.cuh:168-199.
//
// B200 design: same warp-specialised skeleton as fp8_gemm.cu.  Operands are packed
// e2m1 (2 per byte), K-major, one 128-byte swizzle span = 256 K-elements per stage;
// tcgen05.mma.kind::mxf4nvf4.block_scale.block16 consumes K=64 per instruction.  The
// e4m3 scale factors already sit in global memory in the tensor-core 128x4 atom order
// ([rowTile][kTile][32][4][4] = 512 B per atom), so a stage's 4 atoms are one
// contiguous 2 KB cp.async.bulk; tcgen05.cp.32x128b.warpx4 moves each atom into 4 TMEM
// columns (double-buffered by stage parity) right before the MMA that reads it.
//
// Decode regime (M <= 256, weight-bandwidth bound; the reference's small-M configs are
// nvfp4_scaled_mm_kernels.cu:46-74,214-248): swap-AB - 128 weight rows on UMMA-M, the (padded) tokens on
// UMMA-N = 16..256 - and, because N/128 tiles alone leave most SMs idle (70B TP8 qkv: 10 tiles), CLUSTER SPLIT-K:
// the S <= 8 CTAs of a cluster take consecutive K ranges of one tile; every CTA parks its fp32 partial in its own
// (now idle) pipeline ring, the cluster synchronises, and CTA s finishes the s-th share of the token chunks by
// pulling the S partials over distributed shared memory and adding them in split order (deterministic, no
// workspace, no atomics).
#include <cuda.h>

#include <cstdlib>

#include "umma.cuh"

namespace b200 {

int get_kmajor_tmap(CUtensorMap* out, const void* base, int rows, int k_bytes, int64_t ld_bytes, int box_rows);

constexpr int kF4Threads = 192;
constexpr int kF4BK = 128;       // bytes per stage along K (256 e2m1)
constexpr int kAtomBytes = 512;  // 128 rows x 4 scale factors

template <int BN> struct F4Cfg {
  static constexpr int X_BYTES = 128 * kF4BK;
  static constexpr int Y_BYTES = BN * kF4BK;
  static constexpr int YT = (BN + 127) / 128;           // 128-row SF tiles on the Y side
  static constexpr int SFX_BYTES = 4 * kAtomBytes;      // 4 k-atoms
  static constexpr int SFY_BYTES = YT * 4 * kAtomBytes;
  static constexpr int STAGE_BYTES = X_BYTES + Y_BYTES + SFX_BYTES + SFY_BYTES;
  static constexpr int STAGES = (196 * 1024) / STAGE_BYTES > 8 ? 8 : (196 * 1024) / STAGE_BYTES;
  static constexpr int ACC_COLS = BN < 32 ? 32 : BN;
  static constexpr int SF_COLS = 16 + 16 * YT;  // per parity buffer
  static constexpr int NEED_COLS = ACC_COLS + 2 * SF_COLS;
  static constexpr int TMEM_COLS = NEED_COLS <= 128 ? 128 : (NEED_COLS <= 256 ? 256 : 512);
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
};

struct F4Args {
  void* out;
  const uint8_t* sfx;  // SF of the X-side operand (swizzled atoms)
  const uint8_t* sfy;
  const float* alpha;
  int64_t ldc;
  int M, N, K;
  int num_k_atoms;  // ceil(K / 64)
  int splits;       // cluster split-K factor (swap-AB decode regime), 1 otherwise
};

__device__ __forceinline__ uint32_t f4_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void f4_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 f4_ld_dsmem(uint32_t local_addr, uint32_t cta) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(cta));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(ra)
               : "memory");
  return v;
}

__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// SWIZZLE_NONE K-major descriptor over one 512-byte SF atom: 8-row x 16-byte core
// matrices stacked every 128 bytes (SBO), single core matrix along K (LBO unused).
__device__ __forceinline__ uint64_t make_sf_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(128 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

// 16 consecutive outputs of one row: two 16-byte stores when the chunk is whole and aligned
template <typename OutT>
__device__ __forceinline__ void f4_store_row16(OutT* dst, const float (&v)[16], float alpha, int valid) {
  if (valid >= 16 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if constexpr (std::is_same<OutT, __nv_bfloat16>::value) {
        __nv_bfloat162 p = __floats2bfloat162_rn(v[2 * i] * alpha, v[2 * i + 1] * alpha);
        w[i] = *reinterpret_cast<uint32_t*>(&p);
      } else {
        __half2 p = __floats2half2_rn(v[2 * i] * alpha, v[2 * i + 1] * alpha);
        w[i] = *reinterpret_cast<uint32_t*>(&p);
      }
    }
    reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
    reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < valid) dst[j] = from_float<OutT>(v[j] * alpha);
  }
}

template <int BN, bool kSwap, typename OutT>
__global__ void __launch_bounds__(kF4Threads, 1)
    nvfp4_gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                      const F4Args g) {
  using Cfg = F4Cfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = g.splits;
  const int split = S > 1 ? (int)f4_cluster_ctarank() : 0;
  const int xt = S > 1 ? (int)blockIdx.x / S : (int)blockIdx.x;  // X tile index
  const int x0 = xt * 128, y0 = blockIdx.y * BN;
  const int k_bytes = g.K / 2;
  const int nkb_total = (k_bytes + kF4BK - 1) / kF4BK;
  const int kb_begin = nkb_total * split / S, kb_end = nkb_total * (split + 1) / S;
  const int y_rows = kSwap ? g.M : g.N;
  const int y_tiles_total = (y_rows + 127) / 128;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
    prefetch_tensormap(&tmap_x);
    prefetch_tensormap(&tmap_y);
  }
  if (warp == 1) umma::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();  // setup above overlapped the previous kernel's tail

  if (warp == 0) {
    if (umma::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint64_t hint_x = kSwap ? kEvictFirst : kEvictLast;
      const uint64_t hint_y = kSwap ? kEvictLast : kEvictFirst;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sx = smem + stage * Cfg::STAGE_BYTES;
        uint8_t* sy = sx + Cfg::X_BYTES;
        uint8_t* ssfx = sy + Cfg::Y_BYTES;
        uint8_t* ssfy = ssfx + Cfg::SFX_BYTES;
        const int atoms = min(4, g.num_k_atoms - 4 * kb);
        int yt_live = 0;
#pragma unroll
        for (int t = 0; t < Cfg::YT; ++t) yt_live += (blockIdx.y * Cfg::YT + t) < y_tiles_total ? 1 : 0;
        const uint32_t bytes = Cfg::X_BYTES + Cfg::Y_BYTES + (uint32_t)atoms * kAtomBytes * (1 + yt_live);
        mbar_arrive_expect_tx(&full_bar[stage], bytes);
        tma_load_2d(sx, &tmap_x, &full_bar[stage], kb * kF4BK, x0, hint_x);
        tma_load_2d(sy, &tmap_y, &full_bar[stage], kb * kF4BK, y0, hint_y);
        bulk_load_1d(ssfx, g.sfx + ((int64_t)xt * g.num_k_atoms + 4 * kb) * kAtomBytes, atoms * kAtomBytes,
                     &full_bar[stage]);
#pragma unroll
        for (int t = 0; t < Cfg::YT; ++t) {
          const int yt = blockIdx.y * Cfg::YT + t;
          if (yt < y_tiles_total)
            bulk_load_1d(ssfy + t * 4 * kAtomBytes, g.sfy + ((int64_t)yt * g.num_k_atoms + 4 * kb) * kAtomBytes,
                         atoms * kAtomBytes, &full_bar[stage]);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (umma::elect_one()) {
      constexpr uint32_t idesc = umma::make_idesc_nvf4(128, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        umma::fence_after_sync();
        const uint32_t sx = smem_u32(smem + stage * Cfg::STAGE_BYTES);
        const uint32_t sy = sx + Cfg::X_BYTES;
        const uint32_t ssfx = sy + Cfg::Y_BYTES;
        const uint32_t ssfy = ssfx + Cfg::SFX_BYTES;
        const uint64_t dx = umma::make_desc_k_sw128(sx);
        const uint64_t dy = umma::make_desc_k_sw128(sy);
        const uint32_t t_sfx = tmem_base + Cfg::ACC_COLS + (kb & 1) * Cfg::SF_COLS;
        const uint32_t t_sfy = t_sfx + 16;
        const int atoms = min(4, g.num_k_atoms - 4 * kb);
        for (int kk = 0; kk < atoms; ++kk) {
          umma::cp_32x128b_warpx4(t_sfx + 4 * kk, make_sf_desc(ssfx + kk * kAtomBytes));
#pragma unroll
          for (int t = 0; t < Cfg::YT; ++t)
            umma::cp_32x128b_warpx4(t_sfy + 4 * Cfg::YT * kk + 4 * t,
                                    make_sf_desc(ssfy + t * 4 * kAtomBytes + kk * kAtomBytes));
          umma::mma_nvf4_block16(tmem_base, umma::desc_advance(dx, kk * 32), umma::desc_advance(dy, kk * 32), idesc,
                                 (kb != kb_begin) || (kk != 0), t_sfx + 4 * kk, t_sfy + 4 * Cfg::YT * kk);
        }
        umma::commit(&empty_bar[stage]);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma::commit(tmem_full_bar);
    }
  } else {
    const int q = warp & 3;
    const int xi = x0 + q * 32 + lane;
    mbar_wait(tmem_full_bar, 0);
    umma::fence_after_sync();
    const int x_lim = kSwap ? g.N : g.M;
    const int y_lim = kSwap ? g.M : g.N;
    OutT* outp = reinterpret_cast<OutT*>(g.out);
    const float alpha = *g.alpha;
    if (S > 1) {
      // cluster split-K: park this CTA's fp32 partial in the idle ring, [chunk][j/4][128 rows] float4 so that a
      // warp's stores (and the peers' remote loads) are 512 contiguous bytes
      float4* park = reinterpret_cast<float4*>(smem);
      const int et = q * 32 + lane;
      const int nch = (min(BN, y_lim - y0) + 15) >> 4;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
        uint32_t r[16];
        umma::ld_32x32b_x16(tmem_base + ((uint32_t)(q * 32) << 16) + ch * 16, r);
        umma::ld_wait();
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4)
          park[(ch * 4 + j4) * 128 + et] = make_float4(__uint_as_float(r[j4 * 4]), __uint_as_float(r[j4 * 4 + 1]),
                                                       __uint_as_float(r[j4 * 4 + 2]), __uint_as_float(r[j4 * 4 + 3]));
      }
    } else
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      uint32_t r[16];
      umma::ld_32x32b_x16(tmem_base + ((uint32_t)(q * 32) << 16) + c, r);
      umma::ld_wait();
      if (xi >= x_lim || y0 + c >= y_lim) continue;
      if constexpr (kSwap) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int m = y0 + c + j;
          if (m < g.M) outp[(int64_t)m * g.ldc + xi] = from_float<OutT>(__uint_as_float(r[j]) * alpha);
        }
      } else {
        OutT* dst = outp + (int64_t)xi * g.ldc + y0 + c;
        if (y0 + c + 16 <= y_lim && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if constexpr (std::is_same<OutT, __nv_bfloat16>::value) {
              __nv_bfloat162 p =
                  __floats2bfloat162_rn(__uint_as_float(r[2 * i]) * alpha, __uint_as_float(r[2 * i + 1]) * alpha);
              w[i] = *reinterpret_cast<uint32_t*>(&p);
            } else {
              __half2 p = __floats2half2_rn(__uint_as_float(r[2 * i]) * alpha, __uint_as_float(r[2 * i + 1]) * alpha);
              w[i] = *reinterpret_cast<uint32_t*>(&p);
            }
          }
          reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
          reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (y0 + c + j < y_lim) dst[j] = from_float<OutT>(__uint_as_float(r[j]) * alpha);
        }
      }
    }
  }
  if (S > 1) {
    f4_cluster_sync();  // every partial of the cluster is parked and visible cluster-wide
    if (warp >= 2) {
      const int q = warp & 3, et = q * 32 + lane, xi = x0 + et;
      const int x_lim = kSwap ? g.N : g.M, y_lim = kSwap ? g.M : g.N;
      const int nch = (min(BN, y_lim - y0) + 15) >> 4;
      const int own_lo = nch * split / S, own_hi = nch * (split + 1) / S;
      OutT* outp = reinterpret_cast<OutT*>(g.out);
      const float alpha = *g.alpha;
      const uint32_t park0 = smem_u32(smem);
#pragma unroll 1
      for (int ch = own_lo; ch < own_hi; ++ch) {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
#pragma unroll 1
        for (int sp = 0; sp < S; ++sp) {  // split order: the sum does not depend on which CTA finishes the chunk
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const float4 t = f4_ld_dsmem(park0 + (uint32_t)(((ch * 4 + j4) * 128 + et) * 16), (uint32_t)sp);
            v[j4 * 4 + 0] += t.x, v[j4 * 4 + 1] += t.y, v[j4 * 4 + 2] += t.z, v[j4 * 4 + 3] += t.w;
          }
        }
        if (xi >= x_lim) continue;
        if constexpr (kSwap) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int m = y0 + ch * 16 + j;
            if (m < g.M) outp[(int64_t)m * g.ldc + xi] = from_float<OutT>(v[j] * alpha);
          }
        } else {
          f4_store_row16<OutT>(outp + (int64_t)xi * g.ldc + y0 + ch * 16, v, alpha, y_lim - (y0 + ch * 16));
        }
      }
    }
    f4_cluster_sync();  // peers may still be reading this CTA's ring
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 1) umma::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

template <int BN, bool kSwap, typename OutT>
static int launch_f4(const CUtensorMap& tx, const CUtensorMap& ty, const F4Args& g, dim3 grid, cudaStream_t stream,
                     int* max_clusters_query = nullptr) {
  using Cfg = F4Cfg<BN>;
  auto kern = nvfp4_gemm_kernel<BN, kSwap, OutT>;
  if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc_attr;
  if (max_clusters_query) {  // how many clusters of g.splits CTAs are co-resident on this device
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(g.splits, 1, 1);
    cfg.blockDim = dim3(kF4Threads, 1, 1);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = g.splits;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
    *max_clusters_query = n;
    return 0;
  }
  launch_kernel(kern, dim3(grid), dim3(kF4Threads), Cfg::SMEM_BYTES, stream, g.splits, tx, ty, g);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

// large M: persistent 2-CTA kernel (gemm_persist.cu); returns 1 when it does not take the shape
template <typename OutT>
int run_nvfp4_gemm_2sm(void* out, const void* a, const void* b, const void* sfa, const void* sfb, const float* alpha,
                       int M, int N, int K, int64_t ldc, cudaStream_t stream, int min_m);

template <typename OutT>
static int run_f4(void* out, const void* a, const void* b, const void* sfa, const void* sfb, const float* alpha,
                  int M, int N, int K, int64_t ldc, cudaStream_t stream) {
  F4Args g;
  g.out = out;
  g.alpha = alpha;
  g.ldc = ldc;
  g.M = M;
  g.N = N;
  g.K = K;
  g.num_k_atoms = (K + 63) / 64;
  static const int decode_2sm = [] {
    const char* e = getenv("B200_F4_DECODE_2SM");  // tuning aid: M threshold above which M <= 256 also runs on the
    return (e && *e) ? atoi(e) : 0;                // persistent 2-CTA kernel (0 = never)
  }();
  if (decode_2sm > 0 && M > decode_2sm && M <= 256) {
    const int r2 = run_nvfp4_gemm_2sm<OutT>(out, a, b, sfa, sfb, alpha, M, N, K, ldc, stream, decode_2sm);
    if (r2 <= 0) return r2;
  }
  CUtensorMap tx, ty;
  int rc;
  const int kb = K / 2;
  g.splits = 1;
  const int nkb = (kb + kF4BK - 1) / kF4BK;
  static const int s_cap = [] {
    const char* e = getenv("B200_F4_SMAX");  // tuning aid: largest cluster split-K factor (1 = off)
    return (e && *e) ? atoi(e) : 8;
  }();
  // cluster split-K factor: the largest S <= 8 that keeps base x S CTAs co-resident in ONE wave with >= 4 k-blocks
  // per split (measured: with fewer the two cluster barriers + the DSMEM reduce cost more than the shorter K loop
  // saves - o_proj K = 1024 went 5.3 -> 6.3 us with S = 2)
  auto pick_splits = [&](long base_ctas, auto&& query_max_clusters) -> int {
    for (int s = s_cap < 8 ? s_cap : 8; s >= 2; --s) {
      if (nkb / s < 4 || base_ctas * s > sm_count()) continue;
      if (query_max_clusters(s) >= base_ctas) return s;
    }
    return 1;
  };
  if (M <= 128) {
    const int bn = M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : 128));
    g.sfx = reinterpret_cast<const uint8_t*>(sfb);
    g.sfy = reinterpret_cast<const uint8_t*>(sfa);
    if ((rc = get_kmajor_tmap(&tx, b, N, kb, kb, 128))) return rc;
    if ((rc = get_kmajor_tmap(&ty, a, M, kb, kb, bn))) return rc;
    const int tiles = (N + 127) / 128;
    auto dispatch = [&](dim3 grid, int* query) -> int {
      switch (bn) {
        case 16: return launch_f4<16, true, OutT>(tx, ty, g, grid, stream, query);
        case 32: return launch_f4<32, true, OutT>(tx, ty, g, grid, stream, query);
        case 64: return launch_f4<64, true, OutT>(tx, ty, g, grid, stream, query);
        default: return launch_f4<128, true, OutT>(tx, ty, g, grid, stream, query);
      }
    };
    static int max_active[4][9] = {};  // [bn index][S]: 0 not queried, -1 unsupported
    const int bi = bn == 16 ? 0 : bn == 32 ? 1 : bn == 64 ? 2 : 3;
    const int S = pick_splits(tiles, [&](int s) {
      if (max_active[bi][s] == 0) {
        int n = 0;
        g.splits = s;
        if (dispatch(dim3(1, 1, 1), &n)) n = 0;
        max_active[bi][s] = n > 0 ? n : -1;
      }
      return max_active[bi][s];
    });
    g.splits = S;
    return dispatch(dim3(tiles * S, 1, 1), nullptr);
  }
  static const bool use_2sm = [] {
    const char* e = getenv("B200_F4_2SM");  // tuning aid: 0 = the one-tile-per-CTA kernel below
    return !(e && *e && atoi(e) == 0);
  }();
  if (use_2sm && M > 256) {
    const int r2 = run_nvfp4_gemm_2sm<OutT>(out, a, b, sfa, sfb, alpha, M, N, K, ldc, stream, 256);
    if (r2 <= 0) return r2;
  }
  g.sfx = reinterpret_cast<const uint8_t*>(sfa);
  g.sfy = reinterpret_cast<const uint8_t*>(sfb);
  // 128x256 tiles unless that leaves SMs idle (few token rows): then twice as many 128x128 tiles
  int bn = N >= 2048 ? 256 : 128;
  if (bn == 256 && (long)((M + 127) / 128) * ((N + 255) / 256) < sm_count()) bn = 128;
  if ((rc = get_kmajor_tmap(&tx, a, M, kb, kb, 128))) return rc;
  if ((rc = get_kmajor_tmap(&ty, b, N, kb, kb, bn))) return rc;
  const int mt = (M + 127) / 128, nt = (N + bn - 1) / bn;
  if (bn == 128) {  // few tiles (e.g. 70B TP8 qkv at M = 256: 20): cluster split-K along K, clusters laid along x
    static int max_active128[9] = {};
    g.splits = pick_splits((long)mt * nt, [&](int s) {
      if (max_active128[s] == 0) {
        int n = 0;
        g.splits = s;
        if (launch_f4<128, false, OutT>(tx, ty, g, dim3(1, 1, 1), stream, &n)) n = 0;
        max_active128[s] = n > 0 ? n : -1;
      }
      return max_active128[s];
    });
  }
  dim3 grid(mt * g.splits, nt, 1);
  if (bn == 256) return launch_f4<256, false, OutT>(tx, ty, g, grid, stream);
  return launch_f4<128, false, OutT>(tx, ty, g, grid, stream);
}

}  // namespace b200

extern "C" int b200_scaled_mm_nvfp4(void* out, const void* a, const void* b, const void* sfa, const void* sfb,
                                    const float* alpha, int M, int N, int K, int64_t ldc, int out_dtype,
                                    b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID_ARG, "scaled_fp4_mm: bad sizes M=%d N=%d K=%d", M, N, K);
  B200_CHECK_ARG(K % 32 == 0, B200_ERR_INVALID_ARG, "Expected k to be divisible by 32, but got a shape: (%d, %d)", M,
                 K / 2);
  B200_CHECK_ARG(N % 32 == 0, B200_ERR_INVALID_ARG, "Expected n to be divisible by 32, but got b shape: (%d, %d)", N,
                 K / 2);
  B200_CHECK_ARG(out_dtype == B200_DT_BF16 || out_dtype == B200_DT_F16, B200_ERR_INVALID_ARG,
                 "Unsupported output data type of nvfp4 mm");
  if (M == 0) return 0;
  B200_CHECK_ARG(out && a && b && sfa && sfb && alpha, B200_ERR_INVALID_ARG, "scaled_fp4_mm: null pointer");
  B200_CHECK_ARG(ldc >= N, B200_ERR_INVALID_ARG, "scaled_fp4_mm: ldc too small");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(a) % 16 == 0 && reinterpret_cast<uintptr_t>(b) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(sfa) % 16 == 0 && reinterpret_cast<uintptr_t>(sfb) % 16 == 0,
                 B200_ERR_ALIGNMENT, "scaled_fp4_mm: operands and scale factors must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  if (out_dtype == B200_DT_BF16) return run_f4<__nv_bfloat16>(out, a, b, sfa, sfb, alpha, M, N, K, ldc, s);
  return run_f4<__half>(out, a, b, sfa, sfb, alpha, M, N, K, ldc, s);
}
