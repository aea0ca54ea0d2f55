// Host-side runtime glue of libb200vllm: error text, launch counter, SM count,
// and the cuTensorMapEncodeTiled wrapper (resolved through the runtime so the
// library has no link-time dependency on libcuda).
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <utility>
#include <vector>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";
static std::atomic<int> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static std::atomic<int> g_pdl{-1};
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200_PDL");  // default on; B200_PDL=0 restores plain stream order
    v = (e && *e) ? (atoi(e) != 0) : 1;
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}
void set_pdl(int on);
void set_pdl(int on) { g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int ensure_dyn_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::vector<std::pair<const void*, int>> done;  // (kernel, device) pairs already opted in
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  std::lock_guard<std::mutex> lk(mu);
  for (const auto& p : done)
    if (p.first == kernel && p.second == dev) return 0;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(MaxDynamicSharedMemorySize=%d) failed on device %d: %s", bytes, dev,
              cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  done.emplace_back(kernel, dev);
  return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int encode_tmap(void* tmap_out, TmaDtype dt, int rank, const void* base, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, TmaSwizzle sw, bool l2_promote_256) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return B200_ERR_CUDA;
  }
  CUtensorMapDataType cdt;
  switch (dt) {
    case kTmaU8: cdt = CU_TENSOR_MAP_DATA_TYPE_UINT8; break;
    case kTmaU16: cdt = CU_TENSOR_MAP_DATA_TYPE_UINT16; break;
    case kTmaBf16: cdt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; break;
    case kTmaF16: cdt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16; break;
    case kTmaF32: cdt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; break;
    case kTmaU4x16Align8: cdt = CU_TENSOR_MAP_DATA_TYPE_16U4_ALIGN8B; break;
    default: set_error("bad tma dtype"); return B200_ERR_INVALID_ARG;
  }
  CUtensorMapSwizzle csw = sw == kSwizzle128  ? CU_TENSOR_MAP_SWIZZLE_128B
                           : sw == kSwizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B
                           : sw == kSwizzle32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                              : CU_TENSOR_MAP_SWIZZLE_NONE;
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i + 1 < rank) gstr[i] = strides_bytes[i];
  }
  CUresult r = fn(reinterpret_cast<CUtensorMap*>(tmap_out), cdt, (cuuint32_t)rank, const_cast<void*>(base), gdims,
                  gstr, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, csw,
                  l2_promote_256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rank=%d dims=[%llu,%llu,%llu,%llu] box=[%u,%u,%u,%u] base=%p",
              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
              rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0, base);
    return B200_ERR_CUDA;
  }
  return 0;
}

}  // namespace b200

extern "C" {
const char* b200_version(void) { return "b200vllm 1 sm_100a"; }
const char* b200_last_error(void) { return b200::g_err; }
int b200_num_kernels_launched(void) { return b200::g_launches.load(std::memory_order_relaxed); }
void b200_set_pdl(int enabled) { b200::set_pdl(enabled); }
int b200_get_pdl(void) { return b200::pdl_enabled() ? 1 : 0; }
}
