// Symmetric memory through the CUDA virtual-memory-management API, and NVLS multicast objects.
//
// cudaMalloc + CUDA IPC (collectives.cu) gives every rank a pointer to each peer's copy; it
// cannot give the *multicast* address that NVSwitch needs for in-switch reduction
// (multimem.ld_reduce) and broadcast stores (multimem.st): a multicast object only binds memory
// that came from cuMemCreate, and across processes only memory created as shareable.  This file
// is that second allocator:
//
//   per rank   cuMemCreate (POSIX-fd shareable) -> export fd -> peers import -> cuMemMap
//   rank 0     cuMulticastCreate -> export fd -> peers import
//   every rank cuMulticastAddDevice, barrier, cuMulticastBindMem, cuMemMap of the multicast handle
//
// The fds travel between processes over Unix sockets (parallel/peer.py); inside one process
// the handles are used directly.  All driver calls go through cudaGetDriverEntryPoint, so the
// library still does not link libcuda.
//
// The reference reaches NVLS, if at all, only inside NCCL (pytorch/distributed_data_parallel.py:61-62
// passes backend "nccl"); SURVEY.md section 5.8 lists the multicast path as part of the build.
#include "../common/host.h"
#include "../../include/dtdl_b200.h"

#include <unistd.h>

namespace dtdl {

template <typename Fn>
static Fn drv(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess) return nullptr;
  return reinterpret_cast<Fn>(p);
}

static std::string cu_err(CUresult r) {
  typedef CUresult (*NameFn)(CUresult, const char**);
  static NameFn f = drv<NameFn>("cuGetErrorName");
  const char* s = nullptr;
  if (f && f(r, &s) == CUDA_SUCCESS && s) return s;
  return "CUresult " + std::to_string((int)r);
}

#define DTDL_DRV(fn_type, name) \
  static fn_type name##_p = drv<fn_type>(#name); \
  DTDL_REQUIRE(name##_p != nullptr, "driver entry point " #name " not found")

#define DTDL_CHECK_CU(expr)                                                               \
  do {                                                                                    \
    CUresult _r = (expr);                                                                 \
    if (_r != CUDA_SUCCESS) {                                                             \
      ::dtdl::set_last_error(std::string(#expr) + " failed: " + cu_err(_r) +              \
                             " (" __FILE__ ":" + std::to_string(__LINE__) + ")");         \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

typedef CUresult (*MemCreateFn)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
typedef CUresult (*MemReleaseFn)(CUmemGenericAllocationHandle);
typedef CUresult (*MemGranFn)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
typedef CUresult (*MemExportFn)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
typedef CUresult (*MemImportFn)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
typedef CUresult (*MemReserveFn)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
typedef CUresult (*MemAddrFreeFn)(CUdeviceptr, size_t);
typedef CUresult (*MemMapFn)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
typedef CUresult (*MemUnmapFn)(CUdeviceptr, size_t);
typedef CUresult (*MemSetAccessFn)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
typedef CUresult (*McCreateFn)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
typedef CUresult (*McAddDevFn)(CUmemGenericAllocationHandle, CUdevice);
typedef CUresult (*McBindMemFn)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
typedef CUresult (*McGranFn)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);

static CUmemAllocationProp mem_prop(int dev, int shareable) {
  CUmemAllocationProp p = {};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = shareable ? CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR : CU_MEM_HANDLE_TYPE_NONE;
  return p;
}

static CUmulticastObjectProp mc_prop(int world, size_t bytes, int shareable) {
  CUmulticastObjectProp p = {};
  p.numDevices = (unsigned)world;
  p.size = bytes;
  p.handleTypes = shareable ? CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR : 0;
  p.flags = 0;
  return p;
}

// Makes sure `dev` has a primary context, without changing the caller's current device.
static int ensure_context(int dev) {
  int cur = 0;
  DTDL_CHECK_CUDA(cudaGetDevice(&cur));
  if (cur != dev) DTDL_CHECK_CUDA(cudaSetDevice(dev));
  cudaError_t e = cudaFree(nullptr);  // forces the primary context into existence
  if (cur != dev) cudaSetDevice(cur);
  DTDL_CHECK_CUDA(e);
  return 0;
}

}  // namespace dtdl

using namespace dtdl;

// Size granule that satisfies both the physical allocator and the multicast object of `world`
// devices; symmetric NVLS buffers are rounded up to a multiple of it.
extern "C" long long dtdl_vmm_granularity(int dev, int world) {
  MemGranFn mg = drv<MemGranFn>("cuMemGetAllocationGranularity");
  McGranFn cg = drv<McGranFn>("cuMulticastGetGranularity");
  if (!mg || !cg) return -1;
  size_t g1 = 0, g2 = 0;
  CUmemAllocationProp mp = mem_prop(dev, 1);
  if (mg(&g1, &mp, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return -1;
  CUmulticastObjectProp cp = mc_prop(world, 0, 1);
  if (cg(&g2, &cp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return -1;
  size_t g = g1 > g2 ? g1 : g2;
  if (g == 0 || (g % g1) != 0 || (g % g2) != 0) g = g1 * g2;  // not powers of two: be safe
  return (long long)g;
}

extern "C" int dtdl_vmm_create(int dev, long long bytes, int shareable, long long* handle) {
  DTDL_DRV(MemCreateFn, cuMemCreate);
  if (ensure_context(dev)) return 1;
  CUmemAllocationProp p = mem_prop(dev, shareable);
  CUmemGenericAllocationHandle h = 0;
  DTDL_CHECK_CU(cuMemCreate_p(&h, (size_t)bytes, &p, 0));
  *handle = (long long)h;
  return 0;
}

extern "C" int dtdl_vmm_export_fd(long long handle, int* fd) {
  DTDL_DRV(MemExportFn, cuMemExportToShareableHandle);
  int out = -1;
  DTDL_CHECK_CU(cuMemExportToShareableHandle_p(&out, (CUmemGenericAllocationHandle)handle,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = out;
  return 0;
}

extern "C" int dtdl_vmm_import_fd(int fd, long long* handle) {
  DTDL_DRV(MemImportFn, cuMemImportFromShareableHandle);
  CUmemGenericAllocationHandle h = 0;
  DTDL_CHECK_CU(cuMemImportFromShareableHandle_p(&h, (void*)(uintptr_t)fd,
                                                 CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = (long long)h;
  return 0;
}

// Maps a physical or multicast handle into a fresh virtual range, readable and writable from
// the `ndevs` listed devices.
extern "C" int dtdl_vmm_map(long long handle, long long bytes, long long align, const int* devs,
                            int ndevs, void** ptr) {
  DTDL_DRV(MemReserveFn, cuMemAddressReserve);
  DTDL_DRV(MemMapFn, cuMemMap);
  DTDL_DRV(MemSetAccessFn, cuMemSetAccess);
  DTDL_DRV(MemAddrFreeFn, cuMemAddressFree);
  DTDL_DRV(MemUnmapFn, cuMemUnmap);
  DTDL_REQUIRE(ndevs >= 1 && ndevs <= 16, "bad device list");
  CUdeviceptr va = 0;
  DTDL_CHECK_CU(cuMemAddressReserve_p(&va, (size_t)bytes, (size_t)align, 0, 0));
  CUresult r = cuMemMap_p(va, (size_t)bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
  if (r != CUDA_SUCCESS) {
    cuMemAddressFree_p(va, (size_t)bytes);
    DTDL_CHECK_CU(r);
  }
  CUmemAccessDesc acc[16];
  for (int i = 0; i < ndevs; ++i) {
    acc[i] = {};
    acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc[i].location.id = devs[i];
    acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  r = cuMemSetAccess_p(va, (size_t)bytes, acc, (size_t)ndevs);
  if (r != CUDA_SUCCESS) {
    cuMemUnmap_p(va, (size_t)bytes);
    cuMemAddressFree_p(va, (size_t)bytes);
    DTDL_CHECK_CU(r);
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int dtdl_vmm_unmap(void* ptr, long long bytes) {
  DTDL_DRV(MemUnmapFn, cuMemUnmap);
  DTDL_DRV(MemAddrFreeFn, cuMemAddressFree);
  DTDL_CHECK_CU(cuMemUnmap_p((CUdeviceptr)ptr, (size_t)bytes));
  DTDL_CHECK_CU(cuMemAddressFree_p((CUdeviceptr)ptr, (size_t)bytes));
  return 0;
}

extern "C" int dtdl_vmm_release(long long handle) {
  DTDL_DRV(MemReleaseFn, cuMemRelease);
  DTDL_CHECK_CU(cuMemRelease_p((CUmemGenericAllocationHandle)handle));
  return 0;
}

extern "C" int dtdl_vmm_close_fd(int fd) {
  DTDL_REQUIRE(fd >= 0, "bad file descriptor");
  close(fd);
  return 0;
}

extern "C" int dtdl_mc_create(int world, long long bytes, int shareable, long long* handle) {
  DTDL_DRV(McCreateFn, cuMulticastCreate);
  CUmulticastObjectProp p = mc_prop(world, (size_t)bytes, shareable);
  CUmemGenericAllocationHandle h = 0;
  DTDL_CHECK_CU(cuMulticastCreate_p(&h, &p));
  *handle = (long long)h;
  return 0;
}

extern "C" int dtdl_mc_add_device(long long mc, int dev) {
  DTDL_DRV(McAddDevFn, cuMulticastAddDevice);
  typedef CUresult (*GetDevFn)(CUdevice*, int);
  DTDL_DRV(GetDevFn, cuDeviceGet);
  if (ensure_context(dev)) return 1;
  CUdevice d;
  DTDL_CHECK_CU(cuDeviceGet_p(&d, dev));
  DTDL_CHECK_CU(cuMulticastAddDevice_p((CUmemGenericAllocationHandle)mc, d));
  return 0;
}

// Blocks until every device of the team has been added (driver semantics).
extern "C" int dtdl_mc_bind(long long mc, long long mem, long long bytes) {
  DTDL_DRV(McBindMemFn, cuMulticastBindMem);
  DTDL_CHECK_CU(cuMulticastBindMem_p((CUmemGenericAllocationHandle)mc, 0,
                                     (CUmemGenericAllocationHandle)mem, 0, (size_t)bytes, 0));
  return 0;
}
