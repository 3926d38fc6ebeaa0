// Host-side helpers shared by every launcher: error plumbing for the C ABI, the driver entry point for
// cuTensorMapEncodeTiled (resolved at run time so the library links without libcuda), device queries.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <cuda.h>
#include <cuda_runtime.h>

namespace dpt {

// Last error text per thread, readable through dpt_last_error().
inline std::string& last_error_ref() {
  thread_local std::string s;
  return s;
}
inline int fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return -1;
}

#define DPT_CUDA_OK(expr)                                                                              \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess) return ::dpt::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                              __FILE__, __LINE__);                                     \
  } while (0)

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

inline CUtensorMapSwizzle tma_swizzle(int row_bytes) {
  switch (row_bytes) {
    case 128: return CU_TENSOR_MAP_SWIZZLE_128B;
    case 64: return CU_TENSOR_MAP_SWIZZLE_64B;
    case 32: return CU_TENSOR_MAP_SWIZZLE_32B;
    default: return CU_TENSOR_MAP_SWIZZLE_NONE;
  }
}

// bf16 tiled tensor map of rank `rank` (<= 5).  dims/box in elements (dim 0 innermost, contiguous);
// strides in BYTES for dims 1..rank-1.  Out-of-bounds reads fill with zero.
inline int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes, const uint32_t* box, int swizzle_row_bytes) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail("cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bdim[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(base),
                  gdim, gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, tma_swizzle(swizzle_row_bytes),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return fail("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu,%llu box %u,%u,%u swz %d)",
                (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0,
                swizzle_row_bytes);
  }
  return 0;
}

inline int sm_count_real() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
// SMs the persistent kernels leave free (dpt_set_sm_reserve / DPT_SM_RESERVE).  The tcgen05 kernels run ONE CTA per SM
// (they own ~220 KB of its shared memory) with a static tile -> CTA assignment, so a communication kernel that sits on
// an SM (NCCL's send/recv kernels in pipeline mode) stalls one CTA and with it the whole launch.  With a reserve the
// grids are sized to the remaining SMs and the two kinds of kernels never queue behind each other.
inline int& sm_reserve_ref() {
  static int r = [] {
    const char* e = getenv("DPT_SM_RESERVE");
    return e ? atoi(e) : 0;
  }();
  return r;
}
inline int sm_count() {
  const int n = sm_count_real() - sm_reserve_ref();
  return n < 8 ? 8 : n;
}

// Launch with programmatic stream serialization: the kernel may become resident while the previous kernel of the stream
// is still draining, and orders itself behind it with grid_dep_wait() (ptx.cuh) - EVERY kernel launched through here must
// call it before its first global-memory access.  Hides the launch gap and the prologue of ~100 launches per training step.
// DPT_PDL=0 turns the attribute off (plain stream order).
inline bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("DPT_PDL");
    return e == nullptr || atoi(e) != 0;
  }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace dpt
