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
// This file: the one-tile-per-CTA kernel (tokens on X, weights on Y, BN = 128/256) and the C entry point.
// The decode regime (M <= 256, weight-bandwidth bound) lives in fp8_gemm_decode.cu; large M goes to the
// persistent 2-CTA kernel in gemm_persist.cu when the shape allows, else to the kernel below.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "umma.cuh"

namespace b200 {

constexpr int kGemmThreads = 192;
constexpr int kBK = 128;  // bytes of K per stage = one swizzle span (128 fp8 elements)

// kTwo: ring sized for TWO resident CTAs per SM (one CTA's epilogue overlaps the other's mainloop; the kernel is
// not persistent).  Measured better for K <= 8192, worse for longer K where the deeper single-CTA ring wins.
template <int BN, bool kTwo = false> struct GemmCfg {
  static constexpr int X_BYTES = 128 * kBK;
  static constexpr int Y_BYTES = BN * kBK;
  static constexpr int STAGE_BYTES = X_BYTES + Y_BYTES;
  static constexpr int RING_KB = kTwo ? 108 : 200;
  static constexpr int STAGES = (RING_KB * 1024) / STAGE_BYTES > 8 ? 8 : (RING_KB * 1024) / STAGE_BYTES;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

struct GemmArgs {
  void* out;
  const float* a_scales;  // per token (M) or [1]
  const float* b_scales;  // per channel (N) or [1]
  const void* bias;       // [N] out dtype or null
  int64_t ldc;
  int M, N, K;
  int a_scales_numel, b_scales_numel;
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
template <int BN, bool kSwap, typename OutT, bool kTwo = false>
__global__ void __launch_bounds__(kGemmThreads, kTwo ? 2 : 1)
    fp8_gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                    const GemmArgs g) {
  using Cfg = GemmCfg<BN, kTwo>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * 128, y0 = blockIdx.y * BN;
  const int nkb = (g.K + kBK - 1) / kBK;
  const int kb_begin = 0, kb_end = nkb;

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

template <int BN, bool kSwap, typename OutT, bool kTwo = false>
static int launch_fp8_gemm(const CUtensorMap& tx, const CUtensorMap& ty, const GemmArgs& g, dim3 grid,
                           cudaStream_t stream) {
  using Cfg = GemmCfg<BN, kTwo>;
  auto kern = fp8_gemm_kernel<BN, kSwap, OutT, kTwo>;
  if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(kern), Cfg::SMEM_BYTES)) return rc_attr;
  launch_kernel(kern, dim3(grid), dim3(kGemmThreads), Cfg::SMEM_BYTES, stream, 1, tx, ty, g);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

// decode regime (M <= 256): fp8_gemm_decode.cu
template <typename OutT>
int run_fp8_decode(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                   const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                   int b_scales_numel, void* workspace, size_t workspace_bytes, cudaStream_t stream);

int gemm_debug_mode();  // fp8_gemm_decode.cu

// large M: persistent 2-CTA kernel (gemm_persist.cu); returns 1 when it does not take the shape
template <typename OutT>
int run_fp8_gemm_2sm(void* out, const void* a, const void* b, const float* a_scales, const float* b_scales,
                     const void* bias, int M, int N, int K, int64_t lda, int64_t ldb, int64_t ldc, int a_scales_numel,
                     int b_scales_numel, cudaStream_t stream, bool decode);

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
  // Decode regime, which kernel (same-box A/B on B200, profiles/r02_gemm_decode_2sm.md): the 2-CTA kernel with UMMA
  // M = 128 / 256 across the pair (tokens on M, no K split, no fix-up) wins from 65 tokens up when its 128-wide
  // tiles fit ONE wave of SM pairs and K <= 8192 (qkv 10.7 -> 9.6 us at M = 128, 14.1 -> 10.9 at M = 256; o 12.8 ->
  // 10.0 at M = 256).  Otherwise the swap-AB kernel of fp8_gemm_decode.cu stays: long K with few tiles (down_proj)
  // needs its cluster split-K to keep enough weight bytes in flight, gate_up's 224 tiles are a 3.03-wave tail
  // here, and up to 64 tokens its token tile is 2-8 KB per k-block instead of 16 KB, so its ring is nearly all
  // weights.
  static const int decode_2sm = [] {
    const char* e = getenv("B200_GEMM_DECODE_2SM");  // 0 = never, 1 = the policy above (default), 2 = always
    return (e && *e) ? atoi(e) : 1;
  }();
  const int dbg = gemm_debug_mode();  // 4 forces the 2-CTA kernel, 0 / 3 force a partition of the decode kernel
  const bool policy = decode_2sm == 2 ||
                      (decode_2sm == 1 && M > 64 && K <= 8192 && (N + 127) / 128 <= sm_count() / 2);
  if (M <= 256 && (dbg == 4 || (dbg == -1 && policy))) {
    const int r2 = run_fp8_gemm_2sm<OutT>(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                                          b_scales_numel, stream, true);
    if (r2 <= 0) return r2;
  }
  if (M <= 256)
    return run_fp8_decode<OutT>(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                                b_scales_numel, workspace, workspace_bytes, stream);
  static const bool use_2sm = [] {
    const char* e = getenv("B200_GEMM_2SM");  // tuning aid: 0 = one-tile-per-CTA kernel below
    return !(e && *e && atoi(e) == 0);
  }();
  if (use_2sm) {
    const int r2 = run_fp8_gemm_2sm<OutT>(out, a, b, a_scales, b_scales, bias, M, N, K, lda, ldb, ldc, a_scales_numel,
                                          b_scales_numel, stream, false);
    if (r2 <= 0) return r2;
  }
  CUtensorMap tx, ty;
  int rc;
  int bn = (N % 256 == 0 || N > 2048) ? 256 : 128;
  {  // do not leave SMs idle: twice as many 128x128 tiles when they still fit in one wave of resident CTAs
    const long mt = (M + 127) / 128, t256 = mt * ((N + 255) / 256), t128 = mt * ((N + 127) / 128);
    const long resident = (long)sm_count() * (K <= 8192 ? 2 : 1);
    if (bn == 256 && t256 < sm_count() && t128 <= resident) bn = 128;
  }
  if ((rc = get_kmajor_tmap(&tx, a, M, K, lda, 128))) return rc;
  if ((rc = get_kmajor_tmap(&ty, b, N, K, ldb, bn))) return rc;
  dim3 grid((M + 127) / 128, (N + bn - 1) / bn, 1);
  const bool two = K <= 8192;
  if (bn == 256)
    return two ? launch_fp8_gemm<256, false, OutT, true>(tx, ty, g, grid, stream)
               : launch_fp8_gemm<256, false, OutT, false>(tx, ty, g, grid, stream);
  return two ? launch_fp8_gemm<128, false, OutT, true>(tx, ty, g, grid, stream)
             : launch_fp8_gemm<128, false, OutT, false>(tx, ty, g, grid, stream);
}

}  // namespace b200

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
