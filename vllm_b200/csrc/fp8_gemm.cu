// FP8 W8A8 GEMM on tcgen05 / TMEM / TMA:  out = a_s * b_s * (A @ B^T) (+ bias).
// Reference semantics: cutlass_scaled_mm, csrc/libtorch_stable/quantization/w8a8/cutlass/
// scaled_mm_entry.cu:197-270 (checks), c3x/scaled_mm_sm100_fp8_dispatch.cuh:100-296
// (tile / swap-AB dispatch by M), epilogue scaled_mm_epilogues_c3x.hpp:150-183.
//
// B200 design.  Both operands are K-major (A [M][K]; B stored [N][K]), so either can
// sit on the 128-row UMMA "M" side.  The CTA computes a 128 x BN fp32 tile in TMEM:
//   warp 0   TMA producer   (X tile 128 x 128B, Y tile BN x 128B per stage, SW128)
//   warp 1   TMEM alloc + single-thread tcgen05.mma.kind::f8f6f4 issue (4 x K32 per stage)
//   warps 2-5 epilogue      (tcgen05.ld 32x32b, scales, bias, cast, store)
// Decode regime (M <= 256, weight-bandwidth bound): fp8_gemm_decode_kernel below -- swap-AB
// (128 weight rows per tile on the UMMA-M axis, tokens on UMMA-N = 16..256), PERSISTENT
// one CTA per SM, stream-K: the (tile, k-block) units are cut into 148 equal contiguous
// ranges so every SM streams the same number of weight bytes whatever N is; two TMEM
// accumulators so the epilogue of one segment overlaps the TMA/MMA stream of the next;
// a tile whose K range spans several CTAs is fixed up through an fp32 workspace (the tail
// owners emit partials as their FIRST segment, the head owner adds them in its LAST).
// Large M: tokens on X, weights on Y, BN = 256 (fp8_gemm_kernel).
#include <cuda.h>

#include <mutex>
#include <unordered_map>

#include "umma.cuh"

namespace b200 {

constexpr int kGemmThreads = 192;
constexpr int kBK = 128;  // bytes of K per stage = one swizzle span (128 fp8 elements)

template <int BN> struct GemmCfg {
  static constexpr int X_BYTES = 128 * kBK;
  static constexpr int Y_BYTES = BN * kBK;
  static constexpr int STAGE_BYTES = X_BYTES + Y_BYTES;
  static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > 8 ? 8 : (200 * 1024) / STAGE_BYTES;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

struct GemmArgs {
  void* out;
  const float* a_scales;  // per token (M) or [1]
  const float* b_scales;  // per channel (N) or [1]
  const void* bias;       // [N] out dtype or null
  float* splitk_ws;       // fp32 [M][N] zero-initialised (split-K only)
  int* splitk_cnt;        // per output tile arrival counters, zero-initialised
  int64_t ldc;
  int M, N, K;
  int a_scales_numel, b_scales_numel;
  int k_blocks_per_split;
};

template <typename OutT> __device__ __forceinline__ void store16(OutT* dst, const float (&v)[16], bool vec_ok);
template <> __device__ __forceinline__ void store16<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[16], bool vec_ok) {
  if (vec_ok) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      __nv_bfloat162 p = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&p);
    }
    reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
    reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) dst[i] = __float2bfloat16_rn(v[i]);
  }
}
template <> __device__ __forceinline__ void store16<__half>(__half* dst, const float (&v)[16], bool vec_ok) {
  if (vec_ok) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      __half2 p = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&p);
    }
    reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
    reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) dst[i] = __float2half_rn(v[i]);
  }
}

// X rows -> TMEM lanes (128 per CTA), Y rows -> TMEM columns (BN per CTA).
// kSwap == false: X = A (tokens, M), Y = B (weights, N).  kSwap == true: X = B, Y = A.
template <int BN, bool kSwap, typename OutT>
__global__ void __launch_bounds__(kGemmThreads, 1)
    fp8_gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                    const GemmArgs g) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * 128, y0 = blockIdx.y * BN;
  const int total_kb = (g.K + kBK - 1) / kBK;
  const int kb_begin = blockIdx.z * g.k_blocks_per_split;
  const int kb_end = min(total_kb, kb_begin + g.k_blocks_per_split);
  const int nkb = kb_end - kb_begin;  // host guarantees > 0

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

  if (warp == 0) {
    // ===== TMA producer =====
    if (umma::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      // weights are streamed once (evict-first); the small token operand is re-read by every CTA (keep)
      const uint64_t hint_x = kSwap ? kEvictFirst : kEvictLast;
      const uint64_t hint_y = kSwap ? kEvictLast : kEvictFirst;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sx = smem + stage * Cfg::STAGE_BYTES;
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
        tma_load_2d(sx, &tmap_x, &full_bar[stage], kb * kBK, x0, hint_x);
        tma_load_2d(sx + Cfg::X_BYTES, &tmap_y, &full_bar[stage], kb * kBK, y0, hint_y);
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
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        umma::fence_after_sync();
        const uint32_t sx = smem_u32(smem + stage * Cfg::STAGE_BYTES);
        const uint64_t dx = umma::make_desc_k_sw128(sx);
        const uint64_t dy = umma::make_desc_k_sw128(sx + Cfg::X_BYTES);
#pragma unroll
        for (int kk = 0; kk < kBK / 32; ++kk)
          umma::mma_f8f6f4(tmem_base, umma::desc_advance(dx, kk * 32), umma::desc_advance(dy, kk * 32), idesc,
                           (kb | kk) != 0);
        umma::commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma::commit(tmem_full_bar);
    }
  } else {
    // ===== epilogue: warp w owns TMEM lanes 32*(w%4) .. +31 =====
    const int q = warp & 3;
    const int xi = x0 + q * 32 + lane;  // X row of this thread
    mbar_wait(tmem_full_bar, 0);
    umma::fence_after_sync();
    const int x_lim = kSwap ? g.N : g.M;
    const int y_lim = kSwap ? g.M : g.N;
    OutT* outp = reinterpret_cast<OutT*>(g.out);
    const OutT* bias = reinterpret_cast<const OutT*>(g.bias);
    const bool split = gridDim.z > 1;
    // per-row (X side) scale and bias
    float sx_row = 1.f, bias_row = 0.f;
    if (xi < x_lim) {
      if (kSwap) {
        sx_row = g.b_scales[g.b_scales_numel > 1 ? xi : 0];
        if (bias) bias_row = to_float(bias[xi]);
      } else {
        sx_row = g.a_scales[g.a_scales_numel > 1 ? xi : 0];
      }
    }
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      uint32_t r[16];
      umma::ld_32x32b_x16(tmem_base + ((uint32_t)(q * 32) << 16) + c, r);
      umma::ld_wait();
      if (xi >= x_lim || y0 + c >= y_lim) continue;
      if (split) {
        // accumulate raw fp32 partials; scaling happens in the finishing pass
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int yj = y0 + c + j;
          if (yj < y_lim) {
            const int64_t o = kSwap ? (int64_t)yj * g.N + xi : (int64_t)xi * g.N + yj;
            atomicAdd(g.splitk_ws + o, __uint_as_float(r[j]));
          }
        }
        continue;
      }
      float v[16];
      if constexpr (kSwap) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int m = y0 + c + j;
          if (m < g.M) {
            const float as = g.a_scales[g.a_scales_numel > 1 ? m : 0];
            outp[(int64_t)m * g.ldc + xi] = from_float<OutT>(__uint_as_float(r[j]) * as * sx_row + bias_row);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int n = y0 + c + j;  // N % 16 == 0 -> the whole chunk is in range
          const float bs = g.b_scales[g.b_scales_numel > 1 ? n : 0];
          v[j] = __uint_as_float(r[j]) * sx_row * bs + (bias ? to_float(bias[n]) : 0.f);
        }
        OutT* dst = outp + (int64_t)xi * g.ldc + y0 + c;
        store16<OutT>(dst, v, (reinterpret_cast<uintptr_t>(dst) & 15) == 0);
      }
    }
    if (split) {
      // last CTA of this output tile converts the accumulated workspace and re-zeroes it
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");  // epilogue warps only
      __shared__ int s_last;
      if (threadIdx.x == 64) {
        const int tile = blockIdx.y * gridDim.x + blockIdx.x;
        const int prev = atomicAdd(g.splitk_cnt + tile, 1);
        s_last = (prev == (int)gridDim.z - 1);
        if (s_last) g.splitk_cnt[tile] = 0;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (s_last) {
        __threadfence();
        if (xi < x_lim) {
          for (int j = 0; j < BN; ++j) {
            const int yj = y0 + j;
            if (yj >= y_lim) break;
            const int m = kSwap ? yj : xi, n = kSwap ? xi : yj;
            float* w = g.splitk_ws + (int64_t)m * g.N + n;
            const float acc = __ldcg(w);
            *w = 0.f;
            const float as = g.a_scales[g.a_scales_numel > 1 ? m : 0];
            const float bs = g.b_scales[g.b_scales_numel > 1 ? n : 0];
            outp[(int64_t)m * g.ldc + n] = from_float<OutT>(acc * as * bs + (bias ? to_float(bias[n]) : 0.f));
          }
        }
      }
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 1) umma::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}


// ===================================================================================
// Decode-regime kernel: persistent, stream-K, swap-AB.
// ===================================================================================
template <int BN> struct DecCfg {
  static constexpr int X_BYTES = 128 * kBK;  // weights
  static constexpr int Y_BYTES = BN * kBK;   // tokens
  static constexpr int STAGE_BYTES = X_BYTES + Y_BYTES;
  static constexpr int STAGES = (196 * 1024) / STAGE_BYTES > 10 ? 10 : (196 * 1024) / STAGE_BYTES;
  static constexpr int ACC_COLS = BN < 32 ? 32 : BN;  // columns per accumulator buffer
  static constexpr int TMEM_COLS = 2 * ACC_COLS;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/;
};

struct DecArgs {
  void* out;
  const float* a_scales;  // per token (M) or [1]
  const float* b_scales;  // per channel (N) or [1]
  const void* bias;       // [N] out dtype or null
  float* partials;        // [gridDim.x][BN][128] fp32 stream-K partials (null: whole tiles only)
  int* flags;             // [gridDim.x], zero between launches
  int64_t ldc;
  int M, N, K;
  int a_scales_numel, b_scales_numel;
  int num_tiles, num_kb;  // ceil(N/128), ceil(K/128)
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// unit range of CTA c: units are (tile-major) u = tile * num_kb + kb
__device__ __forceinline__ void cta_range(const DecArgs& g, int c, int P, long long& u0, long long& u1) {
  if (g.partials) {
    const long long U = (long long)g.num_tiles * g.num_kb;
    u0 = U * c / P;
    u1 = U * (c + 1) / P;
  } else {  // no workspace: whole tiles only
    u0 = ((long long)g.num_tiles * c / P) * g.num_kb;
    u1 = ((long long)g.num_tiles * (c + 1) / P) * g.num_kb;
  }
}

template <int BN, typename OutT>
__global__ void __launch_bounds__(kGemmThreads, 1)
    fp8_gemm_decode_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_a,
                           const DecArgs g) {
  using Cfg = DecCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = gridDim.x, c = blockIdx.x;
  long long u0, u1;
  cta_range(g, c, P, u0, u1);
  const int KB = g.num_kb;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
    prefetch_tensormap(&tmap_w);
    prefetch_tensormap(&tmap_a);
  }
  if (warp == 1) umma::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: one continuous stream of (weights, tokens) stages across segments =====
    if (umma::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long u = u0; u < u1; ++u) {
        const int tile = (int)(u / KB), kb = (int)(u - (long long)tile * KB);
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sx = smem + stage * Cfg::STAGE_BYTES;
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
        tma_load_2d(sx, &tmap_w, &full_bar[stage], kb * kBK, tile * 128, kEvictFirst);       // streamed once
        tma_load_2d(sx + Cfg::X_BYTES, &tmap_a, &full_bar[stage], kb * kBK, 0, kEvictLast);  // re-read by every CTA
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
        mbar_wait(&tmem_empty_bar[buf], (((uint32_t)seg >> 1) & 1) ^ 1);  // epilogue drained this accumulator
        umma::fence_after_sync();
        const uint32_t tacc = tmem_base + buf * Cfg::ACC_COLS;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          umma::fence_after_sync();
          const uint32_t sx = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint64_t dx = umma::make_desc_k_sw128(sx);
          const uint64_t dy = umma::make_desc_k_sw128(sx + Cfg::X_BYTES);
#pragma unroll
          for (int kk = 0; kk < kBK / 32; ++kk)
            umma::mma_f8f6f4(tacc, umma::desc_advance(dx, kk * 32), umma::desc_advance(dy, kk * 32), idesc,
                             (kb | kk) != 0);
          umma::commit(&empty_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma::commit(&tmem_full_bar[buf]);
        u += nkb;
      }
    }
  } else {
    // ===== epilogue warps: warp w owns TMEM lanes 32*(w%4) .. +31 (= weight rows of the tile) =====
    const int q = warp & 3;
    const int et = q * 32 + lane;  // 0..127 within the epilogue group == TMEM lane == weight row in tile
    OutT* outp = reinterpret_cast<OutT*>(g.out);
    const OutT* bias = reinterpret_cast<const OutT*>(g.bias);
    int seg = 0;
    for (long long u = u0; u < u1; ++seg) {
      const int tile = (int)(u / KB), kb0 = (int)(u - (long long)tile * KB);
      const int nkb = (int)min((long long)(KB - kb0), u1 - u);
      const int buf = seg & 1;
      const bool is_head = (kb0 == 0), is_full = is_head && nkb == KB;
      const int n = tile * 128 + et;  // weight row = output column
      mbar_wait(&tmem_full_bar[buf], ((uint32_t)seg >> 1) & 1);
      umma::fence_after_sync();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + buf * Cfg::ACC_COLS;
      if (!is_head) {
        // ---- tail part of a tile: emit the raw fp32 partial, column-major so every store is coalesced
        float* ws = g.partials + ((size_t)c * BN) * 128 + et;
#pragma unroll 1
        for (int j0 = 0; j0 < BN; j0 += 16) {
          uint32_t r[16];
          umma::ld_32x32b_x16(tacc + j0, r);
          umma::ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j0 + j < g.M) ws[(size_t)(j0 + j) * 128] = __uint_as_float(r[j]);
        }
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        __threadfence();
        epi_bar();
        if (et == 0) st_release(g.flags + c, 1);
      } else {
        // ---- head part (or the whole K range): gather the tail owners' partials, scale, store
        int c_last = c;
        if (!is_full) {
          const long long U = (long long)g.num_tiles * KB;
          const long long last_unit = (long long)(tile + 1) * KB - 1;
          // owner of unit x is the largest cc with U*cc/P <= x
          c_last = (int)(((last_unit + 1) * P - 1) / U);
          while (U * (c_last + 1) / P <= last_unit) ++c_last;
          while (U * c_last / P > last_unit) --c_last;
          if (et == 0) {
            for (int cc = c + 1; cc <= c_last; ++cc) {
              uint32_t spins = 0;
              while (ld_acquire(g.flags + cc) == 0) {
                if (++spins > (1u << 26)) {
                  printf("b200: stream-K fixup wait timed out (cta %d waits for %d)\n", c, cc);
                  __trap();
                }
              }
            }
          }
          epi_bar();
        }
        float sx_row = 1.f, bias_row = 0.f;
        if (n < g.N) {
          sx_row = g.b_scales[g.b_scales_numel > 1 ? n : 0];
          if (bias) bias_row = to_float(bias[n]);
        }
#pragma unroll 1
        for (int j0 = 0; j0 < BN; j0 += 16) {
          uint32_t r[16];
          umma::ld_32x32b_x16(tacc + j0, r);
          umma::ld_wait();
          if (j0 >= g.M) continue;
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          for (int cc = c + 1; cc <= c_last; ++cc) {
            const float* ws = g.partials + ((size_t)cc * BN + j0) * 128 + et;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j0 + j < g.M) v[j] += __ldcg(ws + (size_t)j * 128);
          }
          if (n < g.N) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = j0 + j;
              if (m < g.M) {
                const float as = g.a_scales[g.a_scales_numel > 1 ? m : 0];
                outp[(int64_t)m * g.ldc + n] = from_float<OutT>(v[j] * as * sx_row + bias_row);
              }
            }
          }
        }
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
        if (!is_full) {
          epi_bar();  // every thread has consumed the partials
          if (et == 0)
            for (int cc = c + 1; cc <= c_last; ++cc) g.flags[cc] = 0;
        }
      }
      u += nkb;
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 1) umma::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// ---- host side ------------------------------------------------------------------
struct Tmap2Key {
  const void* base;
  int64_t ld;
  int rows, k, box_rows, dt;
  bool operator==(const Tmap2Key& o) const {
    return base == o.base && ld == o.ld && rows == o.rows && k == o.k && box_rows == o.box_rows && dt == o.dt;
  }
};
struct Tmap2Hash {
  size_t operator()(const Tmap2Key& k) const {
    size_t x = reinterpret_cast<size_t>(k.base);
    auto mix = [&](uint64_t v) { x ^= v + 0x9e3779b97f4a7c15ull + (x << 6) + (x >> 2); };
    mix(k.ld); mix(k.rows); mix(k.k); mix(k.box_rows); mix(k.dt);
    return x;
  }
};
// 2-D K-major operand [rows][k_bytes] (row pitch ld bytes), box = 128 bytes x box_rows, SW128.
int get_kmajor_tmap(CUtensorMap* out, const void* base, int rows, int k_bytes, int64_t ld_bytes, int box_rows) {
  static std::mutex mu;
  static std::unordered_map<Tmap2Key, CUtensorMap, Tmap2Hash> cache;
  Tmap2Key key{base, ld_bytes, rows, k_bytes, box_rows, 0};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return 0;
  }
  uint64_t dims[2] = {(uint64_t)k_bytes, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {128, (uint32_t)box_rows};
  CUtensorMap tm;
  int rc = encode_tmap(&tm, kTmaU8, 2, base, dims, strides, box, kSwizzle128);
  if (rc) return rc;
  if (cache.size() > 8192) cache.clear();
  cache.emplace(key, tm);
  *out = tm;
  return 0;
}

template <int BN, bool kSwap, typename OutT>
static int launch_fp8_gemm(const CUtensorMap& tx, const CUtensorMap& ty, const GemmArgs& g, dim3 grid,
                           cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = fp8_gemm_kernel<BN, kSwap, OutT>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("fp8 gemm: cudaFuncSetAttribute(%d) failed: %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    attr = true;
  }
  kern<<<grid, kGemmThreads, Cfg::SMEM_BYTES, stream>>>(tx, ty, g);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

template <int BN, typename OutT>
static int launch_fp8_decode(const CUtensorMap& tw, const CUtensorMap& ta, const DecArgs& d, dim3 grid,
                             cudaStream_t stream) {
  using Cfg = DecCfg<BN>;
  auto kern = fp8_gemm_decode_kernel<BN, OutT>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("fp8 decode gemm: cudaFuncSetAttribute(%d) failed: %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    attr = true;
  }
  kern<<<grid, kGemmThreads, Cfg::SMEM_BYTES, stream>>>(tw, ta, d);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

template <typename OutT>
static int run_fp8_gemm(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                        const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc,
                        int a_scales_numel, int b_scales_numel, void* workspace, size_t workspace_bytes,
                        cudaStream_t stream) {
  GemmArgs g;
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
  g.splitk_ws = nullptr;
  g.splitk_cnt = nullptr;
  const int total_kb = (K + kBK - 1) / kBK;
  g.k_blocks_per_split = total_kb;
  if (M <= 256) {
    // ---- decode regime: persistent stream-K, swap-AB
    const int bn = M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256)));
    CUtensorMap tw, ta;
    int rc;
    if ((rc = get_kmajor_tmap(&tw, b, N, K, ldb, 128))) return rc;
    if ((rc = get_kmajor_tmap(&ta, a, M, K, lda, bn))) return rc;
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
    d.num_kb = total_kb;
    const long long units = (long long)d.num_tiles * d.num_kb;
    int P = sm_count();
    const size_t need = b200_scaled_mm_workspace_bytes(M, N);
    const bool have_ws = workspace && workspace_bytes >= need && reinterpret_cast<uintptr_t>(workspace) % 16 == 0;
    if (have_ws) {
      // at least 4 k-blocks per CTA so the fixed prologue/epilogue cost is amortised
      if (units / 4 < P) P = (int)(units / 4 > 0 ? units / 4 : 1);
      d.flags = reinterpret_cast<int*>(workspace);
      d.partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + 1024);
    } else {
      if (d.num_tiles < P) P = d.num_tiles;
      d.flags = nullptr;
      d.partials = nullptr;
    }
    dim3 grid(P, 1, 1);
    switch (bn) {
      case 16: return launch_fp8_decode<16, OutT>(tw, ta, d, grid, stream);
      case 32: return launch_fp8_decode<32, OutT>(tw, ta, d, grid, stream);
      case 64: return launch_fp8_decode<64, OutT>(tw, ta, d, grid, stream);
      case 128: return launch_fp8_decode<128, OutT>(tw, ta, d, grid, stream);
      default: return launch_fp8_decode<256, OutT>(tw, ta, d, grid, stream);
    }
  }
  CUtensorMap tx, ty;
  int rc;
  const int bn = (N % 256 == 0 || N > 2048) ? 256 : 128;
  if ((rc = get_kmajor_tmap(&tx, a, M, K, lda, 128))) return rc;
  if ((rc = get_kmajor_tmap(&ty, b, N, K, ldb, bn))) return rc;
  dim3 grid((M + 127) / 128, (N + bn - 1) / bn, 1);
  if (bn == 256) return launch_fp8_gemm<256, false, OutT>(tx, ty, g, grid, stream);
  return launch_fp8_gemm<128, false, OutT>(tx, ty, g, grid, stream);
}

}  // namespace b200

extern "C" size_t b200_scaled_mm_workspace_bytes(int M, int N) {
  // stream-K fixup scratch of the decode regime (M <= 256): 1 KB of per-CTA flags, then one fp32
  // [BN][128] partial per CTA.  Zero the first 1 KB ONCE after allocation; the kernels leave the
  // flags zeroed.  One workspace must not be shared by GEMMs running concurrently on two streams.
  (void)N;
  if (M > 256 || M <= 0) return 0;
  const int bn = M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256)));
  int P = b200::sm_count();
  if (P > 256) P = 256;
  return 1024 + (size_t)P * bn * 128 * 4;
}

extern "C" int b200_scaled_mm_fp8(void* out, const void* a, const void* b, const float* a_scales,
                                  const float* b_scales, const void* bias, int M, int N, int K, int64_t lda,
                                  int64_t ldb, int64_t ldc, int a_scales_numel, int b_scales_numel, int out_dtype,
                                  void* workspace, size_t workspace_bytes, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID_ARG, "scaled_mm: bad sizes M=%d N=%d K=%d", M, N, K);
  B200_CHECK_ARG(K % 16 == 0 && N % 16 == 0, B200_ERR_INVALID_ARG,
                 "scaled_mm: b.shape[0] (K=%d) and b.shape[1] (N=%d) must be multiples of 16", K, N);
  B200_CHECK_ARG(a_scales_numel == 1 || a_scales_numel == M, B200_ERR_INVALID_ARG,
                 "scaled_mm: a_scales must have 1 or M=%d elements (got %d)", M, a_scales_numel);
  B200_CHECK_ARG(b_scales_numel == 1 || b_scales_numel == N, B200_ERR_INVALID_ARG,
                 "scaled_mm: b_scales must have 1 or N=%d elements (got %d)", N, b_scales_numel);
  B200_CHECK_ARG(out_dtype == B200_DT_BF16 || out_dtype == B200_DT_F16, B200_ERR_INVALID_ARG,
                 "scaled_mm: out dtype must be bf16 or fp16");
  if (M == 0) return 0;
  B200_CHECK_ARG(out && a && b && a_scales && b_scales, B200_ERR_INVALID_ARG, "scaled_mm: null pointer");
  B200_CHECK_ARG(lda >= K && ldb >= K && ldc >= N, B200_ERR_INVALID_ARG, "scaled_mm: leading dimensions too small");
  B200_CHECK_ARG(lda % 16 == 0 && ldb % 16 == 0 && reinterpret_cast<uintptr_t>(a) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(b) % 16 == 0,
                 B200_ERR_ALIGNMENT, "scaled_mm: a and b must be 16-byte aligned with row pitches multiple of 16");
  cudaStream_t s = (cudaStream_t)stream;
  if (out_dtype == B200_DT_BF16)
    return run_fp8_gemm<__nv_bfloat16>(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                                       b_scales_numel, workspace, workspace_bytes, s);
  return run_fp8_gemm<__half>(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                              b_scales_numel, workspace, workspace_bytes, s);
}
