// Two-shot all-reduce over NVLink peer memory for the row-parallel linears (o_proj / down_proj) of a
// tensor-parallel decode step.
//
// Reference semantics: `tensor_model_parallel_all_reduce` after RowParallelLinear
// (vllm/model_executor/layers/linear.py, vllm/distributed/communication_op.py); the reference's own fast path for
// small messages is csrc/custom_all_reduce.cuh (one-/two-shot over IPC buffers).  This is a from-scratch kernel on
// the same idea, over buffers that the host maps with torch's symmetric memory (one allocation per rank, every
// rank holds device pointers to all of them plus a small signal pad per rank).
//
// Why not NCCL here: the message is 1 MB (128 tokens x 4096 bf16) and latency-bound; inside the captured decode
// step NCCL's all-reduce cost ~30 us per call at TP8 (64 calls per step: a third of the step).
//
// Algorithm (in place, result ends up in every rank's own buffer):
//   barrier            every rank's producer GEMM has written its buffer (the kernel follows it in stream order)
//   reduce-scatter     rank r sums slice r of all buffers (16-byte peer loads, fp32 accumulation in rank order ->
//                      bit-identical results on every rank and run to run)
//   all-gather         rank r stores the reduced slice r into every rank's buffer (peer stores)
//   barrier            all stores have landed: the local consumer may read the whole buffer
// CTA b of every rank only ever synchronises with CTA b of the other ranks (signal-pad slots [b][peer]), so the
// grid never needs more co-residency than "the same CTA index is running on every GPU".
#include <atomic>

#include "peer_sync.cuh"

namespace b200 {


template <typename T>
__global__ void __launch_bounds__(kArThreads) peer_all_reduce_kernel(const ArArgs a) {
  pdl_trigger();
  pdl_wait();  // the producer of the local buffer has finished
  __shared__ uint32_t* s_pads[kArMaxWorld];
  if ((int)threadIdx.x < a.world) s_pads[threadIdx.x] = a.pads[threadIdx.x];
  const int rank = ar_rank(a);
  uint32_t* my_epoch = a.pads[rank] + blockIdx.x * kArSlotWords + 2 * kArMaxWorld;
  const uint32_t epoch = *my_epoch + 1;  // every rank runs the same sequence of calls: the epochs agree
  ar_barrier(s_pads, rank, a.world, 0, epoch, a.spin_limit);

  const int64_t vec_total = a.numel / 8;            // 16-byte vectors
  const int64_t per_rank = vec_total / a.world;     // my slice
  const int64_t v0 = per_rank * rank;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < per_rank; v += (int64_t)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    uint4 in[kArMaxWorld];
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < a.world) in[q] = reinterpret_cast<const uint4*>(a.bufs[q])[v0 + v];
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < a.world) {
        const T* e = reinterpret_cast<const T*>(&in[q]);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += to_float(e[i]);
      }
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int i = 0; i < 8; ++i) oe[i] = from_float<T>(acc[i]);
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < a.world) reinterpret_cast<uint4*>(a.bufs[q])[v0 + v] = o;
  }
  ar_barrier(s_pads, rank, a.world, 1, epoch, a.spin_limit);
  if (threadIdx.x == 0) *my_epoch = epoch;
}

// One-shot variant for small worlds: every rank reads ALL buffers and writes the sum to a local `out`.  Only the
// start barrier is needed provided the caller alternates between at least two buffers: a rank can rewrite
// buffer X (two calls later) only after passing the start barrier of the call in between, which every peer
// reaches only after it has finished reading X.
template <typename T>
__global__ void __launch_bounds__(kArThreads) peer_all_reduce_oneshot_kernel(const ArArgs a, T* out) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t* s_pads[kArMaxWorld];
  if ((int)threadIdx.x < a.world) s_pads[threadIdx.x] = a.pads[threadIdx.x];
  const int rank = ar_rank(a);
  if (a.emu) out += rank * a.emu_stride;
  uint32_t* my_epoch = a.pads[rank] + blockIdx.x * kArSlotWords + 2 * kArMaxWorld;
  const uint32_t epoch = *my_epoch + 1;
  ar_barrier(s_pads, rank, a.world, 0, epoch, a.spin_limit);
  const int64_t vec_total = a.numel / 8;
  if (a.mc) {
    // NVLS: the switch adds the ranks' copies (fp32 accumulation) and hands back the rounded sum
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < vec_total; v += (int64_t)gridDim.x * blockDim.x)
      reinterpret_cast<uint4*>(out)[v] = multimem_ld_reduce_add<T>(reinterpret_cast<const uint4*>(a.mc) + v);
    __syncthreads();
    if (threadIdx.x == 0) *my_epoch = epoch;
    return;
  }
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < vec_total; v += (int64_t)gridDim.x * blockDim.x) {
    uint4 in[kArMaxWorld];
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < a.world) in[q] = reinterpret_cast<const uint4*>(a.bufs[q])[v];
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int q = 0; q < kArMaxWorld; ++q)
      if (q < a.world) {
        const T* e = reinterpret_cast<const T*>(&in[q]);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += to_float(e[i]);
      }
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int i = 0; i < 8; ++i) oe[i] = from_float<T>(acc[i]);
    reinterpret_cast<uint4*>(out)[v] = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) *my_epoch = epoch;
}

// Two-shot THROUGH THE SWITCH (NVLS), in place: rank r asks the switch for the sum of slice r over all ranks
// (multimem.ld_reduce: one 16-byte load returns 8 summed values, fp32 accumulation, rounded once) and stores it
// with ONE multicast store that the switch replicates into every rank's buffer.  Per GPU the links carry
// ~numel*2 bytes each way whatever the world size (the one-shot forms pull world x that), and neither the loads
// nor the stores are issued once per peer.  Same barriers as peer_all_reduce_kernel.
template <typename T>
__global__ void __launch_bounds__(kArThreads) nvls_all_reduce_kernel(const ArArgs a) {
  pdl_trigger();
  pdl_wait();
  __shared__ uint32_t* s_pads[kArMaxWorld];
  if ((int)threadIdx.x < a.world) s_pads[threadIdx.x] = a.pads[threadIdx.x];
  const int rank = a.rank;
  uint32_t* my_epoch = a.pads[rank] + blockIdx.x * kArSlotWords + 2 * kArMaxWorld;
  const uint32_t epoch = *my_epoch + 1;
  ar_barrier(s_pads, rank, a.world, 0, epoch, a.spin_limit);
  const int64_t per_rank = a.numel / 8 / a.world;
  uint4* mc = reinterpret_cast<uint4*>(const_cast<void*>(a.mc)) + per_rank * rank;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < per_rank; v += (int64_t)gridDim.x * blockDim.x) {
    const uint4 sum = multimem_ld_reduce_add<T>(mc + v);
    multimem_st<T>(mc + v, sum);
  }
  ar_barrier(s_pads, rank, a.world, 1, epoch, a.spin_limit);
  if (threadIdx.x == 0) *my_epoch = epoch;
}

}  // namespace b200

namespace b200 {
static std::atomic<unsigned long long> g_ar_spin_limit{1ull << 28};
unsigned long long ar_spin_limit() { return g_ar_spin_limit.load(std::memory_order_relaxed); }
}  // namespace b200

extern "C" void b200_peer_set_barrier_timeout(unsigned long long max_polls) {
  b200::g_ar_spin_limit.store(max_polls, std::memory_order_relaxed);
}

static int peer_all_reduce_impl(void* out, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads, int rank,
                                int world_size, int64_t numel, int dtype, int64_t byte_offset,
                                int64_t signal_pad_bytes, b200_stream_t stream, uint64_t multicast_ptr = 0) {
  using namespace b200;
  B200_CHECK_ARG(peer_buffers && peer_signal_pads, B200_ERR_INVALID_ARG, "peer_all_reduce: null pointer table");
  const bool emu = rank == -1;  // single-GPU emulation of all ranks (see include/b200_vllm.h)
  B200_CHECK_ARG(world_size >= 2 && world_size <= kArMaxWorld && (emu || (rank >= 0 && rank < world_size)),
                 B200_ERR_INVALID_ARG, "peer_all_reduce: world size %d / rank %d not supported (2..8 ranks)", world_size,
                 rank);
  B200_CHECK_ARG(dtype == B200_DT_BF16 || dtype == B200_DT_F16, B200_ERR_UNSUPPORTED,
                 "peer_all_reduce: only 16-bit floating point tensors");
  B200_CHECK_ARG(numel > 0 && numel % (8 * world_size) == 0 && byte_offset % 16 == 0, B200_ERR_INVALID_ARG,
                 "peer_all_reduce: numel (%lld) must be a multiple of 8 * world_size and the offset 16-byte aligned",
                 (long long)numel);
  // one CTA per ~8 KB of the local slice, bounded by the signal pad (32 words per CTA) and by 32
  int64_t per_rank_vec = numel / 8 / (out ? 1 : world_size);
  int grid = (int)((per_rank_vec + kArThreads - 1) / kArThreads);
  const int64_t max_grid = signal_pad_bytes / (kArSlotWords * 4);
  if (grid > 64) grid = 64;
  if (emu && grid > sm_count() / world_size) grid = sm_count() / world_size;  // one cooperative wave
  if (grid > max_grid) grid = (int)max_grid;
  B200_CHECK_ARG(grid >= 1, B200_ERR_INVALID_ARG, "peer_all_reduce: signal pad too small (%lld bytes)",
                 (long long)signal_pad_bytes);
  ArArgs a;
  for (int q = 0; q < kArMaxWorld; ++q) {
    a.bufs[q] = q < world_size ? reinterpret_cast<void*>(peer_buffers[q] + byte_offset) : nullptr;
    a.pads[q] = q < world_size ? reinterpret_cast<uint32_t*>(peer_signal_pads[q]) : nullptr;
  }
  a.rank = emu ? 0 : rank;
  a.world = world_size;
  a.numel = numel;
  a.spin_limit = ar_spin_limit();
  a.emu = emu ? 1 : 0;
  a.emu_stride = numel;
  a.mc = (multicast_ptr && !emu) ? reinterpret_cast<const void*>(multicast_ptr + byte_offset) : nullptr;
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 g(grid, emu ? world_size : 1);
  if (!out && a.mc) {
    if (dtype == B200_DT_BF16) launch_kernel_ex(nvls_all_reduce_kernel<__nv_bfloat16>, g, dim3(kArThreads), 0, s, 1, false, a);
    else launch_kernel_ex(nvls_all_reduce_kernel<__half>, g, dim3(kArThreads), 0, s, 1, false, a);
  } else if (out) {
    B200_CHECK_ARG(reinterpret_cast<uintptr_t>(out) % 16 == 0, B200_ERR_ALIGNMENT, "peer_all_reduce: out must be 16-byte aligned");
    if (dtype == B200_DT_BF16)
      launch_kernel_ex(peer_all_reduce_oneshot_kernel<__nv_bfloat16>, g, dim3(kArThreads), 0, s, 1, emu, a,
                       reinterpret_cast<__nv_bfloat16*>(out));
    else
      launch_kernel_ex(peer_all_reduce_oneshot_kernel<__half>, g, dim3(kArThreads), 0, s, 1, emu, a,
                       reinterpret_cast<__half*>(out));
  } else if (dtype == B200_DT_BF16) {
    launch_kernel_ex(peer_all_reduce_kernel<__nv_bfloat16>, g, dim3(kArThreads), 0, s, 1, emu, a);
  } else {
    launch_kernel_ex(peer_all_reduce_kernel<__half>, g, dim3(kArThreads), 0, s, 1, emu, a);
  }
  B200_CHECK_LAUNCH();
  count_launch();
  return 0;
}

extern "C" int b200_peer_all_reduce(const uint64_t* peer_buffers, const uint64_t* peer_signal_pads, int rank,
                                    int world_size, int64_t numel, int dtype, int64_t byte_offset,
                                    int64_t signal_pad_bytes, b200_stream_t stream) {
  return peer_all_reduce_impl(nullptr, peer_buffers, peer_signal_pads, rank, world_size, numel, dtype, byte_offset,
                              signal_pad_bytes, stream);
}

extern "C" int b200_peer_all_reduce_out(void* out, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                                        int rank, int world_size, int64_t numel, int dtype, int64_t byte_offset,
                                        int64_t signal_pad_bytes, b200_stream_t stream) {
  B200_CHECK_ARG(out != nullptr, B200_ERR_INVALID_ARG, "peer_all_reduce_out: null output");
  return peer_all_reduce_impl(out, peer_buffers, peer_signal_pads, rank, world_size, numel, dtype, byte_offset,
                              signal_pad_bytes, stream);
}

extern "C" int b200_nvls_all_reduce_out(void* out, uint64_t multicast_ptr, const uint64_t* peer_buffers,
                                        const uint64_t* peer_signal_pads, int rank, int world_size, int64_t numel,
                                        int dtype, int64_t byte_offset, int64_t signal_pad_bytes, b200_stream_t stream) {
  B200_CHECK_ARG(out != nullptr && multicast_ptr != 0, B200_ERR_INVALID_ARG,
                 "nvls_all_reduce_out: null output or no multicast mapping (NVLS unavailable)");
  B200_CHECK_ARG((multicast_ptr + byte_offset) % 16 == 0, B200_ERR_ALIGNMENT, "nvls_all_reduce_out: multicast address must be 16-byte aligned");
  return peer_all_reduce_impl(out, peer_buffers, peer_signal_pads, rank, world_size, numel, dtype, byte_offset,
                              signal_pad_bytes, stream, multicast_ptr);
}

extern "C" int b200_nvls_all_reduce(uint64_t multicast_ptr, const uint64_t* peer_buffers,
                                    const uint64_t* peer_signal_pads, int rank, int world_size, int64_t numel,
                                    int dtype, int64_t byte_offset, int64_t signal_pad_bytes, b200_stream_t stream) {
  B200_CHECK_ARG(multicast_ptr != 0 && rank >= 0, B200_ERR_INVALID_ARG,
                 "nvls_all_reduce: no multicast mapping (NVLS unavailable) or emulated rank");
  B200_CHECK_ARG((multicast_ptr + byte_offset) % 16 == 0, B200_ERR_ALIGNMENT, "nvls_all_reduce: multicast address must be 16-byte aligned");
  return peer_all_reduce_impl(nullptr, peer_buffers, peer_signal_pads, rank, world_size, numel, dtype, byte_offset,
                              signal_pad_bytes, stream, multicast_ptr);
}
