// Symmetric device buffers for multi-process teams: what ProcessGroupNCCL's registered buffers / c10d
// symmetric memory provide to the reference's DDP path (train.py:58, utils/train_utils.py:196), reduced to what
// the fused collectives need.  Each rank cudaMalloc's its own zeroed buffer, exports a CUDA IPC handle, and
// imports every peer's handle; the resulting pointers go straight into dpt_team / dpt_fused_ar_desc.  Handles
// are 64 opaque bytes and are exchanged by the caller (torch.distributed all_gather in the Python package).
#include <cstring>

#include "../../include/dpt_b200.h"
#include "../common/host.h"

using namespace dpt;

extern "C" int dpt_symm_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) return fail("symm_alloc: bad arguments");
  void* p = nullptr;
  DPT_CUDA_OK(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail("symm_alloc: cudaMemset failed: %s", cudaGetErrorString(e));
  }
  *out = p;
  return 0;
}

extern "C" int dpt_symm_free(void* p) {
  if (p) DPT_CUDA_OK(cudaFree(p));
  return 0;
}

extern "C" int dpt_ipc_export(void* p, unsigned char handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  cudaIpcMemHandle_t h;
  DPT_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle, &h, 64);
  return 0;
}

extern "C" int dpt_ipc_import(const unsigned char handle[64], void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  DPT_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out = p;
  return 0;
}

extern "C" int dpt_ipc_close(void* p) {
  if (p) DPT_CUDA_OK(cudaIpcCloseMemHandle(p));
  return 0;
}
