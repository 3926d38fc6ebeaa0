// Symmetric memory on the CUDA virtual-memory-management API, plus NVLS multicast objects: the pieces of
// ProcessGroupNCCL / c10d symmetric memory that the reference's DDP path reaches through torch (train.py:58,
// utils/train_utils.py:196), owned here so the fused collective can use `multimem.ld_reduce` / `multimem.st`
// (csrc/comm/fused_allreduce.cu, DPT_AR_MULTIMEM): the NVSwitch adds the ranks' gradients in the fabric and
// replicates the updated weights, instead of every GPU pulling (world-1) peer copies over its own links.
//
// Life of a multicast buffer (one GPU per process, or several GPUs in one process):
//   every rank : dpt_vmm_alloc        cuMemCreate (POSIX-fd shareable) + reserve + map + access   -> ptr, handle, fd
//   every rank : dpt_vmm_import       map every peer's allocation from its fd                      -> peer ptr
//   rank 0     : dpt_mc_create        cuMulticastCreate for `world` devices                        -> mc handle, fd
//   other ranks: dpt_mc_import        the same object from rank 0's fd
//   every rank : dpt_mc_add_device    (all ranks must have added their device before anyone binds)
//   every rank : dpt_mc_bind          bind the local allocation at offset 0 of the multicast object
//   every rank : dpt_mc_map           map the multicast object                                      -> multicast ptr
// File descriptors travel between processes over AF_UNIX sockets (SCM_RIGHTS); the Python package does that
// (distributedpytorch_b200/parallel/comm.py).  The driver API is reached through cudaGetDriverEntryPoint, so the
// library links against nothing but the CUDA runtime and still builds on machines without a driver.
#include <mutex>

#include <unistd.h>

#include "../../include/dpt_b200.h"
#include "../common/host.h"

using namespace dpt;

namespace {

struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
};

template <typename F>
bool load(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(p);
  return true;
}

const DriverApi& drv() {
  static DriverApi d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= load("cuMemCreate", &d.MemCreate);
    ok &= load("cuMemRelease", &d.MemRelease);
    ok &= load("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
    ok &= load("cuMemAddressReserve", &d.MemAddressReserve);
    ok &= load("cuMemAddressFree", &d.MemAddressFree);
    ok &= load("cuMemMap", &d.MemMap);
    ok &= load("cuMemUnmap", &d.MemUnmap);
    ok &= load("cuMemSetAccess", &d.MemSetAccess);
    ok &= load("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
    ok &= load("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
    ok &= load("cuMulticastCreate", &d.MulticastCreate);
    ok &= load("cuMulticastAddDevice", &d.MulticastAddDevice);
    ok &= load("cuMulticastBindMem", &d.MulticastBindMem);
    ok &= load("cuMulticastGetGranularity", &d.MulticastGetGranularity);
    ok &= load("cuDeviceGet", &d.DeviceGet);
    ok &= load("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    ok &= load("cuGetErrorString", &d.GetErrorString);
    d.ok = ok;
  });
  return d;
}

int cu_fail(const char* what, CUresult r) {
  const char* s = nullptr;
  if (drv().GetErrorString) drv().GetErrorString(r, &s);
  return fail("%s failed: %s (CUresult %d)", what, s ? s : "?", static_cast<int>(r));
}
#define DPT_CU_OK(call)                                   \
  do {                                                    \
    const CUresult r_ = (call);                           \
    if (r_ != CUDA_SUCCESS) return cu_fail(#call, r_);    \
  } while (0)

CUmemAllocationProp device_prop(int device) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

CUmulticastObjectProp mc_prop(int world, size_t bytes) {
  CUmulticastObjectProp prop = {};
  prop.numDevices = static_cast<unsigned int>(world);
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int device, void** out) {
  const DriverApi& d = drv();
  CUdeviceptr va = 0;
  DPT_CU_OK(d.MemAddressReserve(&va, bytes, align, 0, 0));
  CUresult r = d.MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(va, bytes);
    return cu_fail("cuMemMap", r);
  }
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.MemUnmap(va, bytes);
    d.MemAddressFree(va, bytes);
    return cu_fail("cuMemSetAccess", r);
  }
  *out = reinterpret_cast<void*>(va);
  return 0;
}

int enter_device(int device) {
  DPT_CUDA_OK(cudaSetDevice(device));
  DPT_CUDA_OK(cudaFree(nullptr));  // the runtime creates the primary context lazily; the driver calls below need it
  return 0;
}

// Makes `device` current (with its primary context) for the rest of the calling function and puts the caller's device
// back on return: a process that drives several GPUs must not find its current device changed by a library call.
struct DeviceScope {
  int prev = -1;
  int rc;
  explicit DeviceScope(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    rc = enter_device(device);
  }
  ~DeviceScope() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
#define DPT_ENTER_DEVICE(device) \
  DeviceScope dpt_scope_(device); \
  if (dpt_scope_.rc) return -1

}  // namespace

// 1 if the device can do switch multicast / in-fabric reduction (NVLS), 0 if not.  Needs VMM and POSIX-fd export too.
extern "C" int dpt_vmm_query(int device, int* vmm_supported, int* multicast_supported) {
  if (vmm_supported) *vmm_supported = 0;
  if (multicast_supported) *multicast_supported = 0;
  const DriverApi& d = drv();
  if (!d.ok) return 0;  // no driver (build machine) or an old one: both answers stay 0
  DPT_ENTER_DEVICE(device);
  CUdevice dev;
  DPT_CU_OK(d.DeviceGet(&dev, device));
  int vmm = 0, fd = 0, mc = 0;
  DPT_CU_OK(d.DeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev));
  DPT_CU_OK(d.DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  DPT_CU_OK(d.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  if (vmm_supported) *vmm_supported = vmm && fd;
  if (multicast_supported) *multicast_supported = vmm && fd && mc;
  return 0;
}

// Size every rank must allocate for a request of `bytes`: a multiple of the device's allocation granularity and, when
// `multicast_world` > 1, of the multicast granularity for that many devices.
extern "C" int dpt_vmm_round_size(int device, size_t bytes, int multicast_world, size_t* rounded) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!rounded || bytes == 0) return fail("vmm_round_size: bad arguments");
  DPT_ENTER_DEVICE(device);
  const CUmemAllocationProp prop = device_prop(device);
  size_t gran = 0;
  DPT_CU_OK(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (multicast_world > 1) {
    const CUmulticastObjectProp mp = mc_prop(multicast_world, bytes);
    size_t mg = 0;
    DPT_CU_OK(d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    if (mg > gran) gran = mg;
  }
  *rounded = (bytes + gran - 1) / gran * gran;
  return 0;
}

// Allocate `bytes` (already rounded) on `device`, map it read-write, zero it, export a POSIX fd for the peers.
extern "C" int dpt_vmm_alloc(int device, size_t bytes, void** ptr, unsigned long long* handle, int* fd) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!ptr || !handle || !fd || bytes == 0) return fail("vmm_alloc: bad arguments");
  DPT_ENTER_DEVICE(device);
  const CUmemAllocationProp prop = device_prop(device);
  CUmemGenericAllocationHandle h = 0;
  DPT_CU_OK(d.MemCreate(&h, bytes, &prop, 0));
  void* p = nullptr;
  if (map_rw(h, bytes, 0, device, &p)) {
    d.MemRelease(h);
    return -1;
  }
  const cudaError_t e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) return fail("vmm_alloc: cudaMemset failed: %s", cudaGetErrorString(e));
  int out_fd = -1;
  DPT_CU_OK(d.MemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *ptr = p;
  *handle = static_cast<unsigned long long>(h);
  *fd = out_fd;
  return 0;
}

// Map a peer's allocation (its exported fd, received over a unix socket) into this device's address space.
extern "C" int dpt_vmm_import(int device, int fd, size_t bytes, void** ptr, unsigned long long* handle) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!ptr || !handle || fd < 0) return fail("vmm_import: bad arguments");
  DPT_ENTER_DEVICE(device);
  CUmemGenericAllocationHandle h = 0;
  DPT_CU_OK(d.MemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  void* p = nullptr;
  if (map_rw(h, bytes, 0, device, &p)) {
    d.MemRelease(h);
    return -1;
  }
  *ptr = p;
  *handle = static_cast<unsigned long long>(h);
  return 0;
}

// Same process, another device: map an allocation made by dpt_vmm_alloc on a peer device (no fd round trip).
extern "C" int dpt_vmm_map_local(int device, unsigned long long handle, size_t bytes, void** ptr) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!ptr) return fail("vmm_map_local: bad arguments");
  DPT_ENTER_DEVICE(device);
  return map_rw(static_cast<CUmemGenericAllocationHandle>(handle), bytes, 0, device, ptr);
}

// Same process, another device: let `device` read and write an existing mapping too (one address for every device of
// the process, like cudaMalloc + peer access), instead of a second mapping of the same allocation.
extern "C" int dpt_vmm_grant(void* ptr, size_t bytes, int device) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!ptr || bytes == 0) return fail("vmm_grant: bad arguments");
  DPT_ENTER_DEVICE(device);
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DPT_CU_OK(d.MemSetAccess(reinterpret_cast<CUdeviceptr>(ptr), bytes, &acc, 1));
  return 0;
}

extern "C" int dpt_vmm_unmap(void* ptr, size_t bytes) {
  const DriverApi& d = drv();
  if (!d.ok || !ptr) return 0;
  DPT_CU_OK(d.MemUnmap(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  DPT_CU_OK(d.MemAddressFree(reinterpret_cast<CUdeviceptr>(ptr), bytes));
  return 0;
}

extern "C" int dpt_vmm_release(unsigned long long handle) {
  const DriverApi& d = drv();
  if (!d.ok || !handle) return 0;
  DPT_CU_OK(d.MemRelease(static_cast<CUmemGenericAllocationHandle>(handle)));
  return 0;
}

extern "C" int dpt_close_fd(int fd) {
  if (fd >= 0) close(fd);
  return 0;
}

// ---- multicast objects ------------------------------------------------------------------------------------
extern "C" int dpt_mc_create(int world, size_t bytes, unsigned long long* mc, int* fd) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!mc || !fd || world < 2) return fail("mc_create: bad arguments");
  const CUmulticastObjectProp prop = mc_prop(world, bytes);
  CUmemGenericAllocationHandle h = 0;
  DPT_CU_OK(d.MulticastCreate(&h, &prop));
  int out_fd = -1;
  DPT_CU_OK(d.MemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *mc = static_cast<unsigned long long>(h);
  *fd = out_fd;
  return 0;
}

extern "C" int dpt_mc_import(int fd, unsigned long long* mc) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!mc || fd < 0) return fail("mc_import: bad arguments");
  CUmemGenericAllocationHandle h = 0;
  DPT_CU_OK(d.MemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc = static_cast<unsigned long long>(h);
  return 0;
}

extern "C" int dpt_mc_add_device(unsigned long long mc, int device) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  DPT_ENTER_DEVICE(device);
  CUdevice dev;
  DPT_CU_OK(d.DeviceGet(&dev, device));
  DPT_CU_OK(d.MulticastAddDevice(static_cast<CUmemGenericAllocationHandle>(mc), dev));
  return 0;
}

// Bind this device's allocation at offset 0 of the multicast object (every device must have been added first).
extern "C" int dpt_mc_bind(unsigned long long mc, int device, unsigned long long mem_handle, size_t bytes) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  DPT_ENTER_DEVICE(device);
  DPT_CU_OK(d.MulticastBindMem(static_cast<CUmemGenericAllocationHandle>(mc), 0,
                               static_cast<CUmemGenericAllocationHandle>(mem_handle), 0, bytes, 0));
  return 0;
}

// Map the multicast object: stores to the returned address land in every bound copy, multimem.ld_reduce sums them.
extern "C" int dpt_mc_map(unsigned long long mc, int device, size_t bytes, void** ptr) {
  const DriverApi& d = drv();
  if (!d.ok) return fail("vmm: CUDA driver entry points unavailable");
  if (!ptr) return fail("mc_map: bad arguments");
  DPT_ENTER_DEVICE(device);
  const CUmulticastObjectProp prop = mc_prop(2, bytes);
  size_t gran = 0;
  DPT_CU_OK(d.MulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  return map_rw(static_cast<CUmemGenericAllocationHandle>(mc), bytes, gran, device, ptr);
}
