// CUDA driver entry points for virtual memory management + multicast objects (NVLS), and the per-engine multicast state of the
// multi-PROCESS data-parallel path: every rank (process) binds its own gradient buffer into ONE multicast object, so that a
// multimem.ld_reduce on the multicast address returns the sum over all ranks' gradients, reduced inside the NVSwitch.
//
// Set-up across processes (driven from Python, parallel/dist.py, with host barriers in between):
//   rank 0: cuMulticastCreate(handleTypes = POSIX fd) -> cuMemExportToShareableHandle -> the fd travels to the other ranks over a
//           Unix-domain socket (SCM_RIGHTS);  other ranks: cuMemImportFromShareableHandle
//   all   : cuMulticastAddDevice(own GPU)                 -- every device must be added before any memory is bound
//   all   : cuMemCreate(own GPU) -> cuMulticastBindMem    -- same offset on every rank
//   all   : map the own memory (unicast, read/write: the wgrad kernels accumulate into it) and the multicast object
// Replaces the NCCL/MPI NVLS transport Horovod would pick up on such a box (reference rpv.py:63-65).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <string>

namespace cib {

struct NvlsDrv {
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  bool ok = false;
};
static NvlsDrv g_nvls;

static int nvls_load_driver() {
  if (g_nvls.ok) return 0;
  auto get = [](const char* name, void** out) -> int {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q) != cudaSuccess || !*out || q != cudaDriverEntryPointSuccess)
      return fail(std::string("driver entry point missing: ") + name);
    return 0;
  };
  CIB_TRY(get("cuMulticastCreate", (void**)&g_nvls.MulticastCreate));
  CIB_TRY(get("cuMulticastAddDevice", (void**)&g_nvls.MulticastAddDevice));
  CIB_TRY(get("cuMulticastBindMem", (void**)&g_nvls.MulticastBindMem));
  CIB_TRY(get("cuMulticastGetGranularity", (void**)&g_nvls.MulticastGetGranularity));
  CIB_TRY(get("cuMulticastUnbind", (void**)&g_nvls.MulticastUnbind));
  CIB_TRY(get("cuMemCreate", (void**)&g_nvls.MemCreate));
  CIB_TRY(get("cuMemRelease", (void**)&g_nvls.MemRelease));
  CIB_TRY(get("cuMemAddressReserve", (void**)&g_nvls.MemAddressReserve));
  CIB_TRY(get("cuMemAddressFree", (void**)&g_nvls.MemAddressFree));
  CIB_TRY(get("cuMemMap", (void**)&g_nvls.MemMap));
  CIB_TRY(get("cuMemUnmap", (void**)&g_nvls.MemUnmap));
  CIB_TRY(get("cuMemSetAccess", (void**)&g_nvls.MemSetAccess));
  CIB_TRY(get("cuMemGetAllocationGranularity", (void**)&g_nvls.MemGetAllocationGranularity));
  CIB_TRY(get("cuDeviceGetAttribute", (void**)&g_nvls.DeviceGetAttribute));
  CIB_TRY(get("cuDeviceGet", (void**)&g_nvls.DeviceGet));
  CIB_TRY(get("cuMemExportToShareableHandle", (void**)&g_nvls.MemExportToShareableHandle));
  CIB_TRY(get("cuMemImportFromShareableHandle", (void**)&g_nvls.MemImportFromShareableHandle));
  g_nvls.ok = true;
  return 0;
}

#define CIB_CU(call)                                                                                              \
  do {                                                                                                            \
    CUresult r_ = (call);                                                                                         \
    if (r_ != CUDA_SUCCESS) return fail(std::string(#call " -> CUresult ") + std::to_string((int)r_));            \
  } while (0)


struct EngineNvls {
  int world = 0;
  size_t bytes = 0, gran = 0;                      // size of the bound range (gradient buffer rounded up to the multicast granularity)
  CUmemGenericAllocationHandle mc = 0, mem = 0;
  CUdeviceptr uc = 0, mcva = 0;                    // own memory (unicast) / the multicast mapping
  int device = 0;
  bool added = false, bound = false, ready = false;
};

static int engine_nvls_size(EngineNvls* s, int device, int world, size_t payload_bytes) {
  CIB_TRY(nvls_load_driver());
  CUdevice d; CIB_CU(g_nvls.DeviceGet(&d, device));
  int sup = 0;
  CIB_CU(g_nvls.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d));
  if (!sup) return fail("nvls: device " + std::to_string(device) + " reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 0");
  s->device = device; s->world = world;
  CUmulticastObjectProp mp; memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)world; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR; mp.size = payload_bytes;
  size_t gran = 0;
  CUresult r0 = g_nvls.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (r0 != CUDA_SUCCESS || gran == 0) return fail("nvls: cuMulticastGetGranularity -> CUresult " + std::to_string((int)r0));
  CUmemAllocationProp ap; memset(&ap, 0, sizeof ap);
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t g2 = 0;
  CIB_CU(g_nvls.MemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (g2 > gran) gran = g2;
  s->gran = gran;
  s->bytes = (payload_bytes + gran - 1) / gran * gran;
  return 0;
}
// rank 0: create the multicast object and export it as a POSIX file descriptor
static int engine_nvls_create(EngineNvls* s, int* fd_out) {
  CUmulticastObjectProp mp; memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)s->world; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR; mp.size = s->bytes;
  CUresult rc = g_nvls.MulticastCreate(&s->mc, &mp);
  if (rc != CUDA_SUCCESS) { s->mc = 0; return fail("nvls: cuMulticastCreate -> CUresult " + std::to_string((int)rc) + " (multicast not enabled on this platform)"); }
  int fd = -1;
  CIB_CU(g_nvls.MemExportToShareableHandle(&fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd_out = fd;
  return 0;
}
static int engine_nvls_import(EngineNvls* s, int fd) {
  CIB_CU(g_nvls.MemImportFromShareableHandle(&s->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return 0;
}
static int engine_nvls_add_device(EngineNvls* s) {
  CUdevice d; CIB_CU(g_nvls.DeviceGet(&d, s->device));
  CIB_CU(g_nvls.MulticastAddDevice(s->mc, d));
  s->added = true;
  return 0;
}
// after EVERY rank has added its device: own memory, bind, map both views
static int engine_nvls_bind_map(EngineNvls* s) {
  CUmemAllocationProp ap; memset(&ap, 0, sizeof ap);
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = s->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;     // must match the multicast object's handle type, or the bind is refused
  CIB_CU(g_nvls.MemCreate(&s->mem, s->bytes, &ap, 0));
  CIB_CU(g_nvls.MulticastBindMem(s->mc, 0, s->mem, 0, s->bytes, 0));
  s->bound = true;
  CUmemAccessDesc acc; memset(&acc, 0, sizeof acc);
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc.location.id = s->device; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CIB_CU(g_nvls.MemAddressReserve(&s->uc, s->bytes, s->gran, 0, 0));
  CIB_CU(g_nvls.MemMap(s->uc, s->bytes, 0, s->mem, 0));
  CIB_CU(g_nvls.MemSetAccess(s->uc, s->bytes, &acc, 1));
  CIB_CU(g_nvls.MemAddressReserve(&s->mcva, s->bytes, s->gran, 0, 0));
  CIB_CU(g_nvls.MemMap(s->mcva, s->bytes, 0, s->mc, 0));
  CIB_CU(g_nvls.MemSetAccess(s->mcva, s->bytes, &acc, 1));
  if (cudaMemset((void*)s->uc, 0, s->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return fail("nvls: clearing the gradient buffer failed");
  s->ready = true;
  return 0;
}
static void engine_nvls_destroy(EngineNvls* s) {
  if (!s) return;
  if (g_nvls.ok) {
    if (s->mcva) { g_nvls.MemUnmap(s->mcva, s->bytes); g_nvls.MemAddressFree(s->mcva, s->bytes); }
    if (s->bound) { CUdevice d; if (g_nvls.DeviceGet(&d, s->device) == CUDA_SUCCESS) g_nvls.MulticastUnbind(s->mc, d, 0, s->bytes); }
    if (s->uc) { g_nvls.MemUnmap(s->uc, s->bytes); g_nvls.MemAddressFree(s->uc, s->bytes); }
    if (s->mem) g_nvls.MemRelease(s->mem);
    if (s->mc) g_nvls.MemRelease(s->mc);
  }
  delete s;
}

}  // namespace cib
