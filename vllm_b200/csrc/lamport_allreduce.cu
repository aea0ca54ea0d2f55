// Barrier-free all-reduce over NVLink peer memory ("Lamport" protocol), optionally fused with the residual add +
// RMSNorm + per-token FP8 quant that follows a row-parallel linear in the decoder layer.
//
// Reference semantics: tensor_model_parallel_all_reduce after RowParallelLinear (vllm/model_executor/layers/linear.py
// :1657-1660); the reference reaches the same idea through FlashInfer's trtllm_allreduce_fusion
// (vllm/distributed/device_communicators/flashinfer_all_reduce.py, vllm/compilation/passes/fusion/allreduce_rms_fusion.py).
// This is a from-scratch kernel.
//
// Why: profiles/r02_allreduce.md - a two-shot all-reduce with flag barriers costs ~16 us for 1 MB at 8 ranks however the
// bytes travel (peer loads, peer stores or the NVSwitch's in-switch sum), because each barrier is a st.release.sys that
// first collects the acknowledgements of every earlier store (one NVLink round trip) and then sends a flag (another
// one-way trip).  Here a datum IS its own flag: receive buffers are pre-filled with a sentinel bit pattern that no
// pushed word can have (0x80008000 = two 16-bit negative zeros; pushed words equal to it are rewritten as +0, +0), the
// producer pushes plain 16-byte stores into the consumers' buffers, the consumer polls the words it needs until none
// of them is the sentinel.  One one-way NVLink latency per shot, no fences, no acknowledgements.
//
// Buffers: three generations per rank (call k uses generation k mod 3, and re-fills the generation call k-1 used
// with sentinels; a peer can push into that generation only in its call k+2, which it cannot enter before it has
// consumed this rank's pushes of call k+1, which follow this kernel in stream order).  The call counter and the extent
// each generation was last used with live in DEVICE memory, so a captured CUDA graph replays correctly whatever the
// number of calls per replay is.
//   one-shot  (world <= 4): rank r pushes its whole [rows][cols] message into slot r of every peer; every rank sums
//             the world slots (its own contribution straight from its input) in rank order with fp32 accumulation.
//   two-shot  (world > 4):  rank r pushes column slice j of every row into slot r of rank j's scatter area; rank j
//             sums the slots (rank order, fp32), rounds to T and pushes the reduced slice into every rank's gather
//             area; every rank then reads whole rows from its gather area.  Two one-way latencies.
// CTA b owns rows b, b + grid, ...; it only ever waits for the CTAs that own the same rows on the other ranks, and a
// grid never exceeds one co-resident wave.  Sums are bit-identical to b200_peer_all_reduce(_out) (same order, same
// single rounding) except that a sum whose every addend word was the sentinel pattern loses the sign of its zeros.
#include <atomic>

#include "nvfp4_quant.cuh"
#include "peer_sync.cuh"

namespace b200 {

constexpr uint32_t kLpSentinel = 0x80008000u;
constexpr int kLpStateBytes = 256;  // u32 state[64]: [0] call counter, [1..3] extent (16-byte vectors) per generation, [8] ticket

struct LpArgs {
  uint8_t* ws[kArMaxWorld];  // every rank's workspace: state, then 3 generations of gen_vecs 16-byte vectors
  int rank, world;
  int emu;                   // single-GPU emulation: rank = blockIdx.y (see ArArgs)
  int two_shot;
  int64_t gen_vecs;          // capacity of one generation, in 16-byte vectors
  int rows, cols;            // message shape, T elements; cols % (8 * world) == 0
  unsigned long long spin_limit;
};

__device__ __forceinline__ uint4 lp_ld_volatile(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void lp_st_volatile(uint4* p, const uint4& v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t lp_clean(uint32_t w) { return w == kLpSentinel ? 0u : w; }
__device__ __forceinline__ uint4 lp_clean4(uint4 v) {
  return make_uint4(lp_clean(v.x), lp_clean(v.y), lp_clean(v.z), lp_clean(v.w));
}
__device__ __forceinline__ bool lp_arrived(const uint4& v) {
  return v.x != kLpSentinel && v.y != kLpSentinel && v.z != kLpSentinel && v.w != kLpSentinel;
}
// poll one 16-byte vector of a receive buffer until every word of it has been written
__device__ __forceinline__ uint4 lp_wait(const uint4* p, unsigned long long limit) {
  uint4 v = lp_ld_volatile(p);
  unsigned long long spins = 0;
  while (!lp_arrived(v)) {
    if (limit && ++spins > limit) {
      printf("b200: lamport all-reduce timed out (block %d,%d thread %d waits for %p)\n", blockIdx.x, blockIdx.y,
             threadIdx.x, (const void*)p);
      __trap();
    }
    v = lp_ld_volatile(p);
  }
  return v;
}

template <typename T> __device__ __forceinline__ void lp_accumulate(float (&acc)[8], const uint4& v) {
  const T* e = reinterpret_cast<const T*>(&v);
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] += to_float(e[i]);
}
template <typename T> __device__ __forceinline__ uint4 lp_round(const float (&acc)[8]) {
  uint4 o;
  T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int i = 0; i < 8; ++i) oe[i] = from_float<T>(acc[i]);
  return o;
}

// The reduction proper, for the rows this CTA owns.  One code path for all epilogues: every summed 16-byte vector
// (8 values of T: fp32 accumulation in rank order, rounded once) is handed to `sink(row, v, sum)` by the thread
// that owns vector v of the row (v = tid, tid + blockDim, ...), and `row_done(row)` is then called by ALL threads of
// the CTA (it may contain __syncthreads).  The one-shot form sums while it polls; the two-shot form reads the
// gather area.
template <typename T, typename Sink, typename RowDone>
__device__ __forceinline__ void lp_all_reduce_rows(const LpArgs& a, const T* __restrict__ in, int rank, Sink&& sink,
                                                   RowDone&& row_done) {
  const int nv = a.cols / 8;  // 16-byte vectors per row
  uint32_t* state = reinterpret_cast<uint32_t*>(a.ws[rank]);
  const uint32_t call = *reinterpret_cast<volatile uint32_t*>(state);
  const int g = call % 3, gclr = (g + 2) % 3;
  const int64_t clr_vecs = *reinterpret_cast<volatile uint32_t*>(state + 1 + gclr);
  const int64_t msg_vecs = (int64_t)a.rows * nv;
  auto gen_base = [&](int r, int gen) { return reinterpret_cast<uint4*>(a.ws[r] + kLpStateBytes) + (int64_t)gen * a.gen_vecs; };
  uint4* mine = gen_base(rank, g);
  const uint4* inv = reinterpret_cast<const uint4*>(in);
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint4 sentinel4 = make_uint4(kLpSentinel, kLpSentinel, kLpSentinel, kLpSentinel);

  if (!a.two_shot) {
    // ---- push my rows into slot `rank` of every peer ----
    for (int row = blockIdx.x; row < a.rows; row += gridDim.x)
      for (int v = tid; v < nv; v += nt) {
        const uint4 x = lp_clean4(inv[(int64_t)row * nv + v]);
#pragma unroll
        for (int q = 0; q < kArMaxWorld; ++q)
          if (q < a.world && q != rank) lp_st_volatile(gen_base(q, g) + (int64_t)rank * msg_vecs + (int64_t)row * nv + v, x);
      }
    // ---- while the pushes fly: re-arm the generation the previous call used ----
    {
      uint4* clr = gen_base(rank, gclr);
      for (int64_t i = (int64_t)blockIdx.x * nt + tid; i < clr_vecs; i += (int64_t)gridDim.x * nt) clr[i] = sentinel4;
    }
    // ---- sum the slots in rank order ----
    for (int row = blockIdx.x; row < a.rows; row += gridDim.x) {
      for (int v = tid; v < nv; v += nt) {
        const int64_t idx = (int64_t)row * nv + v;
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
        for (int q = 0; q < kArMaxWorld; ++q)
          if (q < a.world) {
            const uint4 x = q == rank ? lp_clean4(inv[idx]) : lp_wait(mine + (int64_t)q * msg_vecs + idx, a.spin_limit);
            lp_accumulate<T>(acc, x);
          }
        sink(row, v, lp_round<T>(acc));
      }
      row_done(row);
    }
    return;
  }
  // ------------------------------------------------ two-shot ------------------------------------------------
  const int cpr = nv / a.world;                       // vectors of one row per rank
  const int64_t sc_vecs = (int64_t)a.rows * cpr;      // one scatter slot
  // scatter area: world slots of sc_vecs (= msg_vecs in total), gather area: msg_vecs behind it
  for (int row = blockIdx.x; row < a.rows; row += gridDim.x)
    for (int v = tid; v < nv; v += nt) {
      const int j = v / cpr;
      if (j == rank) continue;
      const uint4 x = lp_clean4(inv[(int64_t)row * nv + v]);
      lp_st_volatile(gen_base(j, g) + (int64_t)rank * sc_vecs + (int64_t)row * cpr + (v - j * cpr), x);
    }
  {
    uint4* clr = gen_base(rank, gclr);
    for (int64_t i = (int64_t)blockIdx.x * nt + tid; i < clr_vecs; i += (int64_t)gridDim.x * nt) clr[i] = sentinel4;
  }
  for (int row = blockIdx.x; row < a.rows; row += gridDim.x)
    for (int c = tid; c < cpr; c += nt) {
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
      for (int q = 0; q < kArMaxWorld; ++q)
        if (q < a.world) {
          const uint4 x = q == rank ? lp_clean4(inv[(int64_t)row * nv + rank * cpr + c])
                                    : lp_wait(mine + (int64_t)q * sc_vecs + (int64_t)row * cpr + c, a.spin_limit);
          lp_accumulate<T>(acc, x);
        }
      const uint4 s = lp_clean4(lp_round<T>(acc));
      const int64_t dst = msg_vecs + (int64_t)row * nv + rank * cpr + c;
#pragma unroll
      for (int q = 0; q < kArMaxWorld; ++q)
        if (q < a.world) lp_st_volatile(gen_base(q, g) + dst, s);  // own copy included: one read path below
    }
  for (int row = blockIdx.x; row < a.rows; row += gridDim.x) {
    for (int v = tid; v < nv; v += nt) sink(row, v, lp_wait(mine + msg_vecs + (int64_t)row * nv + v, a.spin_limit));
    row_done(row);
  }
}

// last CTA of the rank: record the extent this generation now holds and advance the call counter
__device__ __forceinline__ void lp_finish(const LpArgs& a, int rank) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* state = reinterpret_cast<uint32_t*>(a.ws[rank]);
    __threadfence();
    if (atomicAdd(state + 8, 1u) == gridDim.x - 1) {
      const uint32_t call = *reinterpret_cast<volatile uint32_t*>(state);
      const int64_t msg_vecs = (int64_t)a.rows * (a.cols / 8);
      state[1 + call % 3] = (uint32_t)(a.two_shot ? 2 * msg_vecs : (int64_t)a.world * msg_vecs);
      state[8] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(state) = call + 1;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(1024) lamport_all_reduce_kernel(const LpArgs a, const T* in, T* out, int64_t emu_stride) {
  pdl_trigger();
  pdl_wait();
  const int rank = a.emu ? (int)blockIdx.y : a.rank;
  if (a.emu) {
    in += (int64_t)rank * emu_stride;
    out += (int64_t)rank * emu_stride;
  }
  const int nv = a.cols / 8;
  lp_all_reduce_rows<T>(
      a, in, rank, [&](int row, int v, const uint4& s) { reinterpret_cast<uint4*>(out)[(int64_t)row * nv + v] = s; },
      [](int) {});
  lp_finish(a, rank);
}

// ---- fused with the consumer: residual += sum; (result, scales) = per-token FP8 quant of RMSNorm(residual) * w ----
// Arithmetic of peer_allreduce_rms_norm_quant_kernel (fused_quant.cu), i.e. of b200_rms_norm_dynamic_per_token_quant
// applied to the all-reduced tensor: bit-identical to those.
constexpr float kLpFp8Max = 448.f;
constexpr float kLpMinScale = 1.0f / (448.f * 512.f);
__device__ __forceinline__ uint8_t lp_e4m3(float x) {
  const float r = fmaxf(-kLpFp8Max, fminf(x, kLpFp8Max));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}
template <bool kMax> __device__ __forceinline__ float lp_block_reduce(float v, float* sh) {
  v = kMax ? warp_max(v) : warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = lane < nw ? sh[lane] : 0.f;
  return kMax ? warp_max(r) : warp_sum(r);
}

template <typename T>
__global__ void __launch_bounds__(1024) lamport_all_reduce_rms_norm_quant_kernel(
    const LpArgs a, const T* in, uint8_t* out, float* scales, const T* __restrict__ w, T* residual, float eps,
    int64_t emu_stride) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sh[32];
  extern __shared__ __align__(16) uint8_t row_buf[];
  float* rowf = reinterpret_cast<float*>(row_buf);
  const int rank = a.emu ? (int)blockIdx.y : a.rank;
  if (a.emu) {
    in += (int64_t)rank * emu_stride;
    out += (int64_t)rank * emu_stride;
    residual += (int64_t)rank * emu_stride;
    scales += (int64_t)rank * a.rows;
  }
  const int hidden = a.cols, nv = hidden / 8;
  float ss = 0.f;
  lp_all_reduce_rows<T>(
      a, in, rank,
      [&](int row, int v, const uint4& s) {
        const T* e = reinterpret_cast<const T*>(&s);
        T* res = residual + (int64_t)row * hidden;
        const uint4 rr = *reinterpret_cast<const uint4*>(res + v * 8);
        const T* rv = reinterpret_cast<const T*>(&rr);
        uint4 o;
        T* ov = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = to_float(e[i]) + to_float(rv[i]);  // all-reduce output is T; then the fused add
          ov[i] = from_float<T>(x);
          rowf[v * 8 + i] = x;
          ss += x * x;
        }
        *reinterpret_cast<uint4*>(res + v * 8) = o;
      },
      [&](int row) {
        const float tot = lp_block_reduce<false>(ss, sh);
        ss = 0.f;
        const float rms = rsqrtf(tot / hidden + eps);
        float amax = 0.f;
        for (int v = threadIdx.x; v < nv; v += blockDim.x) {
          const uint4 wr = *reinterpret_cast<const uint4*>(w + v * 8);
          const T* wv = reinterpret_cast<const T*>(&wr);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float y = to_float(from_float<T>(to_float(from_float<T>(rowf[v * 8 + i] * rms)) * to_float(wv[i])));
            rowf[v * 8 + i] = y;
            amax = fmaxf(amax, fabsf(y));
          }
        }
        amax = lp_block_reduce<true>(amax, sh);
        const float sc = fmaxf(amax / kLpFp8Max, kLpMinScale);
        if (threadIdx.x == 0) scales[row] = sc;
        uint8_t* dst = out + (int64_t)row * hidden;
        for (int v = threadIdx.x; v < hidden / 4; v += blockDim.x) {
          const float4 y = *reinterpret_cast<const float4*>(rowf + v * 4);
          const uint32_t qv = lp_e4m3(__fdiv_rn(y.x, sc)) | (lp_e4m3(__fdiv_rn(y.y, sc)) << 8) |
                              (lp_e4m3(__fdiv_rn(y.z, sc)) << 16) | (lp_e4m3(__fdiv_rn(y.w, sc)) << 24);
          *reinterpret_cast<uint32_t*>(dst + v * 4) = qv;
        }
        __syncthreads();  // rowf is reused by the next row
      });
  lp_finish(a, rank);
}

// ---- fused with the NVFP4 consumer: residual += sum; (codes, swizzled scales) = NVFP4 quant of RMSNorm(residual) * w ----
// Bit-identical to the all-reduce followed by b200_rms_norm_nvfp4_quant (= fused_add_rms_norm + scaled_fp4_quant):
// the sum and the residual add are each rounded to T, and the variance is summed in rms_norm_kernel's element order,
// which is why only threads 0..255 take part in that sum (pushes, polls and the quantiser use the whole block).
// Scale rows of the 128-row padding are zeroed here too.
template <typename T>
__global__ void __launch_bounds__(1024) lamport_all_reduce_rms_norm_nvfp4_quant_kernel(
    const LpArgs a, const T* in, uint32_t* out, uint8_t* sf_out, const T* __restrict__ w, T* residual,
    const float* __restrict__ global_scale, float eps, int64_t emu_stride, int64_t emu_sf_stride) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sh[8];
  extern __shared__ __align__(16) uint8_t row_buf[];
  float* rowf = reinterpret_cast<float*>(row_buf);
  const int rank = a.emu ? (int)blockIdx.y : a.rank;
  if (a.emu) {
    in += (int64_t)rank * emu_stride;
    out += (int64_t)rank * (emu_stride / 8);
    residual += (int64_t)rank * emu_stride;
    sf_out += (int64_t)rank * emu_sf_stride;
  }
  const int hidden = a.cols, nv = hidden / 8;
  const int num_k_tiles = (hidden + 63) / 64;
  const int vecs_total = ((hidden / 16 + 3) / 4) * 4 * 2;  // 8-element vectors incl. the padded scale columns
  const int vecs_rounded = (vecs_total + 31) & ~31;
  const float gs = global_scale ? global_scale[0] : 1.0f;
  const float rcp6 = rcp_approx_ftz(6.0f);
  const float rcp_gs = rcp_approx_ftz(gs);
  const int rows_total = (a.rows + 127) / 128 * 128;
  for (int row = a.rows + blockIdx.x; row < rows_total; row += gridDim.x)
    for (int cv = threadIdx.x; cv < vecs_total; cv += blockDim.x)
      if ((cv & 1) == 0) sf_out[sf_swizzled_offset(row, cv >> 1, num_k_tiles)] = 0;
  lp_all_reduce_rows<T>(
      a, in, rank,
      [&](int row, int v, const uint4& s) {
        const T* e = reinterpret_cast<const T*>(&s);
        T* res = residual + (int64_t)row * hidden;
        const uint4 rr = *reinterpret_cast<const uint4*>(res + v * 8);
        const T* rv = reinterpret_cast<const T*>(&rr);
        uint4 o;
        T* ov = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          ov[i] = from_float<T>(to_float(e[i]) + to_float(rv[i]));  // rounded to T before the norm
          rowf[v * 8 + i] = to_float(ov[i]);
        }
        *reinterpret_cast<uint4*>(res + v * 8) = o;
      },
      [&](int row) {
        __syncthreads();
        float ss = 0.f;
        if (threadIdx.x < 256)  // rms_norm_kernel's element-to-thread map (256 threads, stride 256)
          for (int e = threadIdx.x; e < hidden; e += 256) ss += rowf[e] * rowf[e];
        ss = warp_sum(ss);
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0 && warp < 8) sh[warp] = ss;
        __syncthreads();
        ss = warp_sum(lane < 8 ? sh[lane] : 0.f);
        const float inv = rsqrtf(ss / hidden + eps);
        for (int cv = threadIdx.x; cv < vecs_rounded; cv += blockDim.x) {
          float f[8];
          if (cv < nv) {
            const uint4 wr = *reinterpret_cast<const uint4*>(w + cv * 8);
            const T* wv = reinterpret_cast<const T*>(&wr);
#pragma unroll
            for (int i = 0; i < 8; ++i)
              f[i] = to_float(from_float<T>(to_float(from_float<T>(rowf[cv * 8 + i] * inv)) * to_float(wv[i])));
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = 0.f;
          }
          uint8_t sf8;
          const uint32_t codes = nvfp4_quant8(f, gs, rcp6, rcp_gs, &sf8);
          if (cv < vecs_total) {
            if ((cv & 1) == 0) sf_out[sf_swizzled_offset(row, cv >> 1, num_k_tiles)] = sf8;
            if (cv < nv) out[((int64_t)row * hidden >> 3) + cv] = codes;
          }
        }
        __syncthreads();  // rowf / sh are reused by the next row
      });
  lp_finish(a, rank);
}

}  // namespace b200


namespace b200 {
unsigned long long ar_spin_limit();  // peer_allreduce.cu (b200_peer_set_barrier_timeout)

__global__ void lamport_fill_kernel(uint4* p, int64_t vecs) {
  const uint4 sv = make_uint4(kLpSentinel, kLpSentinel, kLpSentinel, kLpSentinel);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < vecs; i += (int64_t)gridDim.x * blockDim.x) p[i] = sv;
}

static int lp_fill_args(LpArgs& a, const char* what, const uint64_t* peer_workspaces, int64_t workspace_bytes, int rank,
                        int world_size, int64_t rows, int cols, int dtype, int* grid_out) {
  B200_CHECK_ARG(peer_workspaces != nullptr, B200_ERR_INVALID_ARG, "%s: null workspace table", what);
  const bool emu = rank == -1;
  B200_CHECK_ARG(world_size >= 2 && world_size <= kArMaxWorld && (emu || (rank >= 0 && rank < world_size)),
                 B200_ERR_INVALID_ARG, "%s: world size %d / rank %d not supported (2..8 ranks)", what, world_size, rank);
  B200_CHECK_ARG(dtype == B200_DT_BF16 || dtype == B200_DT_F16, B200_ERR_UNSUPPORTED, "%s: only 16-bit floating point tensors",
                 what);
  B200_CHECK_ARG(rows > 0 && rows <= (1 << 20) && cols > 0 && cols % (8 * world_size) == 0 && cols <= 49152, B200_ERR_INVALID_ARG,
                 "%s: cols (%d) must be a multiple of 8 * world_size, at most 49152", what, cols);
  const int64_t gen_vecs = (workspace_bytes - kLpStateBytes) / 3 / 16;
  a.two_shot = world_size > 4 ? 1 : 0;
  const int64_t msg_vecs = rows * (cols / 8);
  const int64_t need = a.two_shot ? 2 * msg_vecs : (int64_t)world_size * msg_vecs;
  B200_CHECK_ARG(workspace_bytes >= kLpStateBytes && need <= gen_vecs && need < (1ll << 32), B200_ERR_WORKSPACE,
                 "%s: workspace too small (%lld bytes; this message needs %lld, see b200_lamport_workspace_bytes)", what,
                 (long long)workspace_bytes, (long long)(kLpStateBytes + 3 * need * 16));
  for (int q = 0; q < kArMaxWorld; ++q) {
    a.ws[q] = q < world_size ? reinterpret_cast<uint8_t*>(peer_workspaces[q]) : nullptr;
    B200_CHECK_ARG(q >= world_size || (peer_workspaces[q] && peer_workspaces[q] % 16 == 0), B200_ERR_ALIGNMENT,
                   "%s: workspaces must be 16-byte aligned", what);
  }
  a.rank = emu ? 0 : rank;
  a.world = world_size;
  a.emu = emu ? 1 : 0;
  a.gen_vecs = gen_vecs;
  a.rows = (int)rows;
  a.cols = cols;
  a.spin_limit = ar_spin_limit();
  // one row per CTA up to 128 CTAs; a grid is one co-resident wave (the CTAs of a rank wait for their peers' CTAs)
  int64_t grid = rows < 128 ? rows : 128;
  if (emu && grid > sm_count() / world_size) grid = sm_count() / world_size;
  if (grid < 1) grid = 1;
  *grid_out = (int)grid;
  return 0;
}
static int lp_threads(int vecs) {
  int t = (vecs + 31) / 32 * 32;
  return t < 64 ? 64 : (t > 1024 ? 1024 : t);
}
}  // namespace b200

extern "C" int64_t b200_lamport_workspace_bytes(int world_size, int64_t max_numel) {
  if (world_size < 2 || world_size > b200::kArMaxWorld || max_numel <= 0) return -1;
  const int64_t per_gen = (world_size > 4 ? 2 : world_size) * max_numel * 2;  // bytes, 16-bit elements
  return b200::kLpStateBytes + 3 * ((per_gen + 15) / 16 * 16);
}

extern "C" int b200_lamport_workspace_init(void* workspace, int64_t workspace_bytes, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(workspace && workspace_bytes >= kLpStateBytes && reinterpret_cast<uintptr_t>(workspace) % 16 == 0,
                 B200_ERR_INVALID_ARG, "lamport_workspace_init: bad workspace");
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(workspace, 0, kLpStateBytes, s);
  if (e != cudaSuccess) {
    set_error("lamport_workspace_init: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  const int64_t vecs = (workspace_bytes - kLpStateBytes) / 16;
  if (vecs > 0) {
    lamport_fill_kernel<<<dim3((unsigned)((vecs + 255) / 256 < 1184 ? (vecs + 255) / 256 : 1184)), dim3(256), 0, s>>>(
        reinterpret_cast<uint4*>(static_cast<uint8_t*>(workspace) + kLpStateBytes), vecs);
    B200_CHECK_LAUNCH();
    count_launch();
  }
  return 0;
}

extern "C" int b200_lamport_all_reduce(void* out, const void* input, const uint64_t* peer_workspaces,
                                       int64_t workspace_bytes, int rank, int world_size, int64_t rows, int cols,
                                       int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(out && input && reinterpret_cast<uintptr_t>(out) % 16 == 0 && reinterpret_cast<uintptr_t>(input) % 16 == 0,
                 B200_ERR_INVALID_ARG, "lamport_all_reduce: null or misaligned tensor");
  LpArgs a;
  int grid = 0;
  if (int rc = lp_fill_args(a, "lamport_all_reduce", peer_workspaces, workspace_bytes, rank, world_size, rows, cols, dtype, &grid))
    return rc;
  const dim3 g(grid, a.emu ? world_size : 1);
  const int threads = lp_threads(cols / 8);
  const int64_t stride = rows * cols;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200_DT_BF16)
    launch_kernel_ex(lamport_all_reduce_kernel<__nv_bfloat16>, g, dim3(threads), 0, s, 1, a.emu != 0, a,
                     (const __nv_bfloat16*)input, (__nv_bfloat16*)out, stride);
  else
    launch_kernel_ex(lamport_all_reduce_kernel<__half>, g, dim3(threads), 0, s, 1, a.emu != 0, a, (const __half*)input,
                     (__half*)out, stride);
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

extern "C" int b200_lamport_all_reduce_rms_norm_quant(void* result, float* scales, const void* input,
                                                      const uint64_t* peer_workspaces, int64_t workspace_bytes, int rank,
                                                      int world_size, const void* weight, void* residual, float epsilon,
                                                      int64_t num_tokens, int hidden, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(result && scales && input && weight && residual, B200_ERR_INVALID_ARG,
                 "lamport_all_reduce_rms_norm_quant: null pointer");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(weight) % 16 == 0 && reinterpret_cast<uintptr_t>(residual) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(input) % 16 == 0 && reinterpret_cast<uintptr_t>(result) % 4 == 0,
                 B200_ERR_ALIGNMENT, "lamport_all_reduce_rms_norm_quant: input/weight/residual must be 16-byte aligned");
  LpArgs a;
  int grid = 0;
  if (int rc = lp_fill_args(a, "lamport_all_reduce_rms_norm_quant", peer_workspaces, workspace_bytes, rank, world_size,
                            num_tokens, hidden, dtype, &grid))
    return rc;
  const dim3 g(grid, a.emu ? world_size : 1);
  const int threads = lp_threads(hidden / 8);
  const int64_t stride = num_tokens * hidden;
  cudaStream_t s = (cudaStream_t)stream;
  auto go = [&](auto tag) -> int {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(lamport_all_reduce_rms_norm_quant_kernel<T>), 49152 * 4))
      return rc_attr;
    launch_kernel_ex(lamport_all_reduce_rms_norm_quant_kernel<T>, g, dim3(threads), (size_t)hidden * 4, s, 1, a.emu != 0, a,
                     (const T*)input, (uint8_t*)result, scales, (const T*)weight, (T*)residual, epsilon, stride);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  };
  return dtype == B200_DT_BF16 ? go(__nv_bfloat16()) : go(__half());
}

extern "C" int b200_lamport_all_reduce_rms_norm_nvfp4_quant(void* result, void* result_block_scale, const void* input,
                                                            const uint64_t* peer_workspaces, int64_t workspace_bytes,
                                                            int rank, int world_size, const void* weight, void* residual,
                                                            const float* input_global_scale, float epsilon,
                                                            int64_t num_tokens, int hidden, int dtype, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK_ARG(result && result_block_scale && input && weight && residual, B200_ERR_INVALID_ARG,
                 "lamport_all_reduce_rms_norm_nvfp4_quant: null pointer");
  B200_CHECK_ARG(hidden % 16 == 0 && hidden <= 32768, B200_ERR_INVALID_ARG, "The N dimension must be multiple of 16.");
  B200_CHECK_ARG(reinterpret_cast<uintptr_t>(weight) % 16 == 0 && reinterpret_cast<uintptr_t>(residual) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(input) % 16 == 0 && reinterpret_cast<uintptr_t>(result) % 4 == 0,
                 B200_ERR_ALIGNMENT, "lamport_all_reduce_rms_norm_nvfp4_quant: input/weight/residual must be 16-byte aligned");
  LpArgs a;
  int grid = 0;
  if (int rc = lp_fill_args(a, "lamport_all_reduce_rms_norm_nvfp4_quant", peer_workspaces, workspace_bytes, rank, world_size,
                            num_tokens, hidden, dtype, &grid))
    return rc;
  const dim3 g(grid, a.emu ? world_size : 1);
  const int64_t stride = num_tokens * hidden;
  const int64_t sf_stride = ((num_tokens + 127) / 128 * 128) * (((hidden / 16 + 3) / 4) * 4);
  cudaStream_t s = (cudaStream_t)stream;
  auto go = [&](auto tag) -> int {
    using T = decltype(tag);
    if (int rc_attr = ensure_dyn_smem(reinterpret_cast<const void*>(lamport_all_reduce_rms_norm_nvfp4_quant_kernel<T>), 32768 * 4))
      return rc_attr;
    const int threads = lp_threads(hidden / 8) < 256 ? 256 : lp_threads(hidden / 8);
    launch_kernel_ex(lamport_all_reduce_rms_norm_nvfp4_quant_kernel<T>, g, dim3(threads), (size_t)hidden * 4, s, 1, a.emu != 0, a,
                     (const T*)input, (uint32_t*)result, (uint8_t*)result_block_scale, (const T*)weight, (T*)residual,
                     input_global_scale, epsilon, stride, sf_stride);
    B200_CHECK_LAUNCH();
    count_launch();
    return 0;
  };
  return dtype == B200_DT_BF16 ? go(__nv_bfloat16()) : go(__half());
}
