// Symmetric memory on the CUDA virtual-memory-management API, with an NVSwitch multicast mapping.
//
// The NVLS path of the gradient engine needs physical memory that can be
// bound to a multicast object, which cudaMalloc memory cannot.  So the arena is created with
// cuMemCreate (shareable as a POSIX file descriptor), mapped locally, imported and mapped by every
// peer, and additionally bound to one multicast object whose mapping gives a pointer on which
// `multimem.ld_reduce` returns the switch-reduced sum over all ranks and `multimem.st` writes to all
// ranks at once.  File descriptors travel between the per-GPU processes over a unix socket
// (parallel/vmm.py); this file only contains the driver calls.
//
// Everything degrades cleanly: ddpb_vmm_supported() reports whether the device/driver exposes
// multicast, and every failure returns an error string instead of aborting, so the caller can fall
// back to the cudaMalloc + CUDA-IPC arena (runtime/symm_mem.cpp) and the unicast kernels.
//
// All driver entry points are resolved at run time (no link against libcuda).
#include <cuda.h>
#include <cuda_runtime.h>
#include <string.h>
#include <unistd.h>

#include <mutex>

#include "../../include/ddp_b200.h"
#include "../common.cuh"

namespace {

struct Driver {
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  bool ok = false;
};

Driver& drv() {
  static Driver d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool all = true;
    auto get = [&](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !*fn) {
        (void)cudaGetLastError();
        all = false;
      }
    };
    get("cuGetErrorString", (void**)&d.GetErrorString);
    get("cuDeviceGetAttribute", (void**)&d.DeviceGetAttribute);
    get("cuMemGetAllocationGranularity", (void**)&d.MemGetAllocationGranularity);
    get("cuMemCreate", (void**)&d.MemCreate);
    get("cuMemRelease", (void**)&d.MemRelease);
    get("cuMemAddressReserve", (void**)&d.MemAddressReserve);
    get("cuMemAddressFree", (void**)&d.MemAddressFree);
    get("cuMemMap", (void**)&d.MemMap);
    get("cuMemUnmap", (void**)&d.MemUnmap);
    get("cuMemSetAccess", (void**)&d.MemSetAccess);
    get("cuMemExportToShareableHandle", (void**)&d.MemExportToShareableHandle);
    get("cuMemImportFromShareableHandle", (void**)&d.MemImportFromShareableHandle);
    get("cuMulticastCreate", (void**)&d.MulticastCreate);
    get("cuMulticastAddDevice", (void**)&d.MulticastAddDevice);
    get("cuMulticastBindMem", (void**)&d.MulticastBindMem);
    get("cuMulticastGetGranularity", (void**)&d.MulticastGetGranularity);
    d.ok = all;
  });
  return d;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().GetErrorString && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown driver error";
}

#define CU_CHECK(expr)                                                                       \
  do {                                                                                       \
    CUresult _r = (expr);                                                                    \
    if (_r != CUDA_SUCCESS) {                                                                \
      ::ddpb::set_last_error("%s:%d: %s -> %s (%d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r); \
      return DDPB_ERR_CUDA;                                                                  \
    }                                                                                        \
  } while (0)

int current_device(int* dev) {
  DDPB_CUDA_CHECK(cudaGetDevice(dev));
  DDPB_CUDA_CHECK(cudaFree(0));          // make sure the primary context exists
  return DDPB_OK;
}

CUmemAllocationProp mem_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_rw(CUmemGenericAllocationHandle h, size_t bytes, size_t align, int dev, void** out) {
  CUdeviceptr va = 0;
  CU_CHECK(drv().MemAddressReserve(&va, bytes, align, 0, 0));
  CU_CHECK(drv().MemMap(va, bytes, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_CHECK(drv().MemSetAccess(va, bytes, &acc, 1));
  *out = (void*)va;
  return DDPB_OK;
}

}  // namespace

// 1 when the current device and driver expose NVSwitch multicast objects.
extern "C" int ddpb_vmm_supported(int* multicast_supported) {
  DDPB_REQUIRE(multicast_supported, "vmm_supported: null argument");
  *multicast_supported = 0;
  if (!drv().ok) return DDPB_OK;
  int dev = 0, v = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS) *multicast_supported = v;
  return DDPB_OK;
}

// Size every rank must use: `bytes` rounded up to the allocation and multicast granularities.
extern "C" int ddpb_vmm_padded_size(size_t bytes, int num_devices, size_t* padded) {
  DDPB_REQUIRE(drv().ok && padded, "vmm: driver entry points unavailable");
  int dev = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  CUmemAllocationProp p = mem_prop(dev);
  size_t g = 0, mg = 0;
  CU_CHECK(drv().MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = num_devices; mp.size = bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (drv().MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) mg = g;
  if (mg > g) g = mg;
  *padded = (bytes + g - 1) / g * g;
  return DDPB_OK;
}

// ALLOCATES `bytes` (already padded) of zeroed device memory with cuMemCreate on the current device,
// maps it, and exports a POSIX fd for the peers.  `handle` is the driver allocation handle (needed
// for the multicast bind).
extern "C" int ddpb_vmm_alloc(size_t bytes, void** ptr, int* fd, unsigned long long* handle) {
  DDPB_REQUIRE(drv().ok && ptr && fd && handle, "vmm_alloc: driver entry points unavailable");
  int dev = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  CUmemAllocationProp p = mem_prop(dev);
  CUmemGenericAllocationHandle h = 0;
  CU_CHECK(drv().MemCreate(&h, bytes, &p, 0));
  rc = map_rw(h, bytes, 0, dev, ptr);
  if (rc) return rc;
  int out_fd = -1;
  CU_CHECK(drv().MemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  DDPB_CUDA_CHECK(cudaMemset(*ptr, 0, bytes));
  *fd = out_fd;
  *handle = (unsigned long long)h;
  return DDPB_OK;
}

// Map a peer's allocation (received as a POSIX fd) into the current device's address space.
extern "C" int ddpb_vmm_import(int fd, size_t bytes, void** ptr) {
  DDPB_REQUIRE(drv().ok && ptr, "vmm_import: driver entry points unavailable");
  int dev = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  CUmemGenericAllocationHandle h = 0;
  CU_CHECK(drv().MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return map_rw(h, bytes, 0, dev, ptr);
}

// Rank 0: create the multicast object for `num_devices` GPUs and export it as a POSIX fd.
extern "C" int ddpb_mc_create(size_t bytes, int num_devices, int* fd, unsigned long long* mc_handle) {
  DDPB_REQUIRE(drv().ok && fd && mc_handle, "mc_create: driver entry points unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = num_devices; mp.size = bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mc = 0;
  CU_CHECK(drv().MulticastCreate(&mc, &mp));
  int out_fd = -1;
  CU_CHECK(drv().MemExportToShareableHandle(&out_fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = out_fd;
  *mc_handle = (unsigned long long)mc;
  return DDPB_OK;
}

extern "C" int ddpb_mc_import(int fd, unsigned long long* mc_handle) {
  DDPB_REQUIRE(drv().ok && mc_handle, "mc_import: driver entry points unavailable");
  CUmemGenericAllocationHandle mc = 0;
  CU_CHECK(drv().MemImportFromShareableHandle(&mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *mc_handle = (unsigned long long)mc;
  return DDPB_OK;
}

// Every rank, step 1 (all ranks must have done this before anyone binds): join the multicast group.
extern "C" int ddpb_mc_add_device(unsigned long long mc_handle) {
  int dev = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  CU_CHECK(drv().MulticastAddDevice((CUmemGenericAllocationHandle)mc_handle, dev));
  return DDPB_OK;
}

// Every rank, step 2: bind the local allocation at offset 0 and map the multicast object.
extern "C" int ddpb_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes,
                                    void** mc_ptr) {
  DDPB_REQUIRE(mc_ptr, "mc_bind_and_map: null argument");
  int dev = 0;
  int rc = current_device(&dev);
  if (rc) return rc;
  CU_CHECK(drv().MulticastBindMem((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle,
                                  0, bytes, 0));
  return map_rw((CUmemGenericAllocationHandle)mc_handle, bytes, 0, dev, mc_ptr);
}

extern "C" int ddpb_vmm_unmap(void* ptr, size_t bytes) {
  if (!ptr) return DDPB_OK;
  CU_CHECK(drv().MemUnmap((CUdeviceptr)ptr, bytes));
  CU_CHECK(drv().MemAddressFree((CUdeviceptr)ptr, bytes));
  return DDPB_OK;
}

extern "C" int ddpb_vmm_release(unsigned long long handle) {
  if (!handle) return DDPB_OK;
  CU_CHECK(drv().MemRelease((CUmemGenericAllocationHandle)handle));
  return DDPB_OK;
}
