// Symmetric memory through the CUDA virtual-memory-management API + NVLink SHARP (NVLS) multicast objects.
//
// cudaMalloc + cudaIpc (allreduce.cu) is enough for peer loads/stores, but an NVLS multicast address -- one
// virtual address whose loads are reduced *inside the NVSwitch* (multimem.ld_reduce) and whose stores are
// replicated to every GPU (multimem.st) -- can only be bound to physical memory created with cuMemCreate.
// So when the host exposes multicast, arenas are allocated here instead:
//   every rank : cuMemCreate (POSIX-fd shareable) -> map locally -> export fd -> peers import + map (unicast P2P)
//   rank 0     : cuMulticastCreate(W devices) -> export fd -> peers import
//   every rank : cuMulticastAddDevice -> (barrier) -> cuMulticastBindMem(own physical handle) -> map the
//                multicast handle -> `mc_ptr`
// File descriptors travel between the per-GPU processes over AF_UNIX sockets (SCM_RIGHTS) in parallel/peer.py.
// Driver entry points are resolved with cudaGetDriverEntryPoint so the library does not link libcuda.
#include <cuda.h>
#include "common.cuh"

namespace {

#define DRV(name) PFN_##name
typedef CUresult (*PFN_cuMemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*PFN_cuMemRelease)(CUmemGenericAllocationHandle);
typedef CUresult (*PFN_cuMemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*PFN_cuMemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*PFN_cuMemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*PFN_cuMemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*PFN_cuMemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*PFN_cuMemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*PFN_cuMulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*PFN_cuMulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*PFN_cuMulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
typedef CUresult (*PFN_cuMulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
typedef CUresult (*PFN_cuDeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*PFN_cuDeviceGet)(CUdevice*, int);

struct Drv {
  PFN_cuMemCreate MemCreate; PFN_cuMemRelease MemRelease; PFN_cuMemExportToShareableHandle Export;
  PFN_cuMemImportFromShareableHandle Import; PFN_cuMemAddressReserve Reserve; PFN_cuMemMap Map;
  PFN_cuMemSetAccess SetAccess; PFN_cuMemGetAllocationGranularity Gran; PFN_cuMulticastCreate McCreate;
  PFN_cuMulticastAddDevice McAdd; PFN_cuMulticastBindMem McBind; PFN_cuMulticastGetGranularity McGran;
  PFN_cuDeviceGetAttribute DevAttr; PFN_cuDeviceGet DevGet;
  bool ok = false;
} g;

template <typename F> bool load(const char* name, F* f) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || !p) return false;
  *f = (F)p;
  return true;
}

int load_all() {
  if (g.ok) return 0;
  bool ok = load("cuMemCreate", &g.MemCreate) && load("cuMemRelease", &g.MemRelease) &&
            load("cuMemExportToShareableHandle", &g.Export) && load("cuMemImportFromShareableHandle", &g.Import) &&
            load("cuMemAddressReserve", &g.Reserve) && load("cuMemMap", &g.Map) && load("cuMemSetAccess", &g.SetAccess) &&
            load("cuMemGetAllocationGranularity", &g.Gran) && load("cuMulticastCreate", &g.McCreate) &&
            load("cuMulticastAddDevice", &g.McAdd) && load("cuMulticastBindMem", &g.McBind) &&
            load("cuMulticastGetGranularity", &g.McGran) && load("cuDeviceGetAttribute", &g.DevAttr) &&
            load("cuDeviceGet", &g.DevGet);
  g.ok = ok;
  return ok ? 0 : -20;
}

int current_device(CUdevice* dev, int* ordinal) {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) return -21;
  cudaFree(0);  // make sure the primary context exists
  if (g.DevGet(dev, d) != CUDA_SUCCESS) return -22;
  *ordinal = d;
  return 0;
}

CUmemAllocationProp alloc_prop(int ordinal) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(CUmemGenericAllocationHandle h, size_t bytes, int ordinal, void** ptr) {
  CUdeviceptr va = 0;
  if (g.Reserve(&va, bytes, 0, 0, 0) != CUDA_SUCCESS) return -30;
  if (g.Map(va, bytes, 0, h, 0) != CUDA_SUCCESS) return -31;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (g.SetAccess(va, bytes, &acc, 1) != CUDA_SUCCESS) return -32;
  *ptr = (void*)va;
  return 0;
}

}  // namespace

// 1 if this device can take part in NVLS multicast objects and fd-shareable VMM allocations.
MPB_API int mpb_vmm_multicast_supported(void) {
  if (load_all()) return 0;
  CUdevice dev; int ord;
  if (current_device(&dev, &ord)) return 0;
  int mc = 0, fd = 0, vmm = 0;
  g.DevAttr(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  g.DevAttr(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  g.DevAttr(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  return (mc && fd && vmm) ? 1 : 0;
}

// Allocation size that satisfies both the VMM and the multicast granularity for `num_devices` GPUs.
MPB_API long long mpb_vmm_round_size(long long bytes, int num_devices) {
  if (load_all()) return -1;
  CUdevice dev; int ord;
  if (current_device(&dev, &ord)) return -1;
  CUmemAllocationProp prop = alloc_prop(ord);
  size_t gran = 0, mcg = 0;
  if (g.Gran(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return -1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)num_devices; mp.size = (size_t)bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (g.McGran(&mcg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) mcg = gran;
  size_t a = gran > mcg ? gran : mcg;
  if (a == 0) a = 2u << 20;
  return (long long)(((size_t)bytes + a - 1) / a * a);
}

// Physical allocation on the current device, mapped read/write locally; `fd_out` can be sent to peer processes.
MPB_API int mpb_vmm_alloc(long long bytes, void** ptr, int* fd_out, unsigned long long* handle_out) {
  int rc = load_all();
  if (rc) return rc;
  CUdevice dev; int ord;
  if ((rc = current_device(&dev, &ord))) return rc;
  CUmemAllocationProp prop = alloc_prop(ord);
  CUmemGenericAllocationHandle h;
  if (g.MemCreate(&h, (size_t)bytes, &prop, 0) != CUDA_SUCCESS) return -40;
  if ((rc = map_rw(h, (size_t)bytes, ord, ptr))) return rc;
  int fd = -1;
  if (g.Export(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) return -41;
  *fd_out = fd;
  *handle_out = (unsigned long long)h;
  MPB_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
  return 0;
}

// Map a peer's allocation (received as a file descriptor) into this process for unicast loads/stores.
MPB_API int mpb_vmm_import(int fd, long long bytes, void** ptr) {
  int rc = load_all();
  if (rc) return rc;
  CUdevice dev; int ord;
  if ((rc = current_device(&dev, &ord))) return rc;
  CUmemGenericAllocationHandle h;
  if (g.Import(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) return -42;
  return map_rw(h, (size_t)bytes, ord, ptr);
}

MPB_API int mpb_mc_create(long long bytes, int num_devices, int* fd_out, unsigned long long* mc_out) {
  int rc = load_all();
  if (rc) return rc;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)num_devices; mp.size = (size_t)bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mc;
  if (g.McCreate(&mc, &mp) != CUDA_SUCCESS) return -50;
  int fd = -1;
  if (g.Export(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) return -51;
  *fd_out = fd;
  *mc_out = (unsigned long long)mc;
  return 0;
}

MPB_API int mpb_mc_import(int fd, unsigned long long* mc_out) {
  int rc = load_all();
  if (rc) return rc;
  CUmemGenericAllocationHandle mc;
  if (g.Import(&mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) return -52;
  *mc_out = (unsigned long long)mc;
  return 0;
}

MPB_API int mpb_mc_add_device(unsigned long long mc) {
  int rc = load_all();
  if (rc) return rc;
  CUdevice dev; int ord;
  if ((rc = current_device(&dev, &ord))) return rc;
  return g.McAdd((CUmemGenericAllocationHandle)mc, dev) == CUDA_SUCCESS ? 0 : -53;
}

// Call after EVERY rank has added its device.  Binds this rank's physical memory and maps the multicast address.
MPB_API int mpb_mc_bind_map(unsigned long long mc, unsigned long long mem_handle, long long bytes, void** mc_ptr) {
  int rc = load_all();
  if (rc) return rc;
  CUdevice dev; int ord;
  if ((rc = current_device(&dev, &ord))) return rc;
  if (g.McBind((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem_handle, 0, (size_t)bytes, 0) != CUDA_SUCCESS)
    return -54;
  return map_rw((CUmemGenericAllocationHandle)mc, (size_t)bytes, ord, mc_ptr);
}
