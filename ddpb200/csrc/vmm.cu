// Own bootstrap of the symmetric gradient buffer and its NVLS multicast alias (SURVEY.md par. 5.8:
// "every rank allocates its arena with the CUDA VMM API, exports it, maps all peers' arenas and binds the
// arena to an NVLS multicast object").  The reference's counterpart is the Gloo rendezvous of
// /root/reference/part2/part2b/main.py:52-58; torch's symmetric memory is the default provider here, this
// file removes the dependency on it (backend "vmm" of parallel/comm.py): without it a torch lacking
// symmetric memory would silently lose the multimem kernels, because plain CUDA IPC cannot carry multicast.
//
// The driver entry points are resolved at run time (the library does not link libcuda).  Handles travel
// between the ranks' processes as POSIX file descriptors (Unix-domain sockets, done in Python).
#include "common.cuh"

#include <unistd.h>

namespace ddpb {

template <typename Fn>
static Fn drv(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<Fn>(p);
}

#define DDPB_DRV(var, name, ...)                                                   \
  static auto var = drv<CUresult (*)(__VA_ARGS__)>(name);                          \
  DDPB_REQUIRE(var != nullptr, "vmm: driver entry point %s is not available", name)

#define DDPB_CHECK_CU(expr)                                                        \
  do {                                                                             \
    CUresult _r = (expr);                                                          \
    if (_r != CUDA_SUCCESS) {                                                      \
      ::ddpb::set_error("%s failed with CUresult %d (%s:%d)", #expr, static_cast<int>(_r), __FILE__, __LINE__); \
      return 1;                                                                    \
    }                                                                              \
  } while (0)

static CUmemAllocationProp local_prop(int dev) {
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace ddpb

using namespace ddpb;

// Size granularity a buffer must be padded to so that it can be created, mapped AND bound to a multicast
// object over `ndev` devices (the larger of the allocation and the multicast granularity).
extern "C" int ddpb_vmm_granularity(int ndev, long long* out) {
  int dev = 0;
  DDPB_CHECK_CUDA(cudaGetDevice(&dev));
  DDPB_CHECK_CUDA(cudaFree(nullptr));   // make sure the primary context exists
  DDPB_DRV(get_gran, "cuMemGetAllocationGranularity", size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  const CUmemAllocationProp prop = local_prop(dev);
  size_t g = 0;
  DDPB_CHECK_CU(get_gran(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  size_t mg = 0;
  auto mc_gran = drv<CUresult (*)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags)>(
      "cuMulticastGetGranularity");
  if (mc_gran != nullptr && ndev > 1) {
    CUmulticastObjectProp mp = {};
    mp.numDevices = static_cast<unsigned int>(ndev);
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (mc_gran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) mg = 0;
  }
  *out = static_cast<long long>(g > mg ? g : mg);
  return 0;
}

// Physical allocation of `bytes` (a multiple of the granularity) on the current device; returns the generic
// handle and a file descriptor other processes can import it from (close it after it has been sent).
extern "C" int ddpb_vmm_create(long long bytes, unsigned long long* handle, int* fd) {
  int dev = 0;
  DDPB_CHECK_CUDA(cudaGetDevice(&dev));
  DDPB_DRV(mem_create, "cuMemCreate", CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  DDPB_DRV(mem_export, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
           unsigned long long);
  const CUmemAllocationProp prop = local_prop(dev);
  CUmemGenericAllocationHandle h;
  DDPB_CHECK_CU(mem_create(&h, static_cast<size_t>(bytes), &prop, 0));
  int out_fd = -1;
  DDPB_CHECK_CU(mem_export(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = static_cast<unsigned long long>(h);
  *fd = out_fd;
  return 0;
}

// Imports an allocation (or a multicast object) another process exported as a file descriptor.
extern "C" int ddpb_vmm_import(int fd, unsigned long long* handle) {
  DDPB_DRV(mem_import, "cuMemImportFromShareableHandle", CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUmemGenericAllocationHandle h;
  DDPB_CHECK_CU(mem_import(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle = static_cast<unsigned long long>(h);
  return 0;
}

// Maps a handle (local, imported, or multicast) read-write into the current device's address space.
extern "C" int ddpb_vmm_map(unsigned long long handle, long long bytes, void** ptr) {
  int dev = 0;
  DDPB_CHECK_CUDA(cudaGetDevice(&dev));
  DDPB_DRV(reserve, "cuMemAddressReserve", CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  DDPB_DRV(map, "cuMemMap", CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  DDPB_DRV(set_access, "cuMemSetAccess", CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUdeviceptr va = 0;
  DDPB_CHECK_CU(reserve(&va, static_cast<size_t>(bytes), 0, 0, 0));
  DDPB_CHECK_CU(map(va, static_cast<size_t>(bytes), 0, static_cast<CUmemGenericAllocationHandle>(handle), 0));
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DDPB_CHECK_CU(set_access(va, static_cast<size_t>(bytes), &acc, 1));
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int ddpb_vmm_unmap(void* ptr, long long bytes) {
  DDPB_DRV(unmap, "cuMemUnmap", CUdeviceptr, size_t);
  DDPB_DRV(addr_free, "cuMemAddressFree", CUdeviceptr, size_t);
  DDPB_CHECK_CU(unmap(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(bytes)));
  DDPB_CHECK_CU(addr_free(reinterpret_cast<CUdeviceptr>(ptr), static_cast<size_t>(bytes)));
  return 0;
}

extern "C" int ddpb_vmm_release(unsigned long long handle) {
  DDPB_DRV(release, "cuMemRelease", CUmemGenericAllocationHandle);
  DDPB_CHECK_CU(release(static_cast<CUmemGenericAllocationHandle>(handle)));
  return 0;
}

extern "C" int ddpb_vmm_close_fd(int fd) { return close(fd) == 0 ? 0 : 1; }

// NVLS multicast object over `ndev` devices (created by one rank, imported by the others through its fd).
extern "C" int ddpb_mc_create(int ndev, long long bytes, unsigned long long* handle, int* fd) {
  DDPB_DRV(mc_create, "cuMulticastCreate", CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  DDPB_DRV(mem_export, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
           unsigned long long);
  CUmulticastObjectProp mp = {};
  mp.numDevices = static_cast<unsigned int>(ndev);
  mp.size = static_cast<size_t>(bytes);
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DDPB_CHECK_CU(mc_create(&h, &mp));
  int out_fd = -1;
  DDPB_CHECK_CU(mem_export(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle = static_cast<unsigned long long>(h);
  *fd = out_fd;
  return 0;
}

// Every rank: add the current device to the multicast object (all ranks must have done so before any binds).
extern "C" int ddpb_mc_add_device(unsigned long long mc_handle) {
  int dev = 0;
  DDPB_CHECK_CUDA(cudaGetDevice(&dev));
  DDPB_DRV(add_dev, "cuMulticastAddDevice", CUmemGenericAllocationHandle, CUdevice);
  DDPB_DRV(dev_get, "cuDeviceGet", CUdevice*, int);
  CUdevice cudev;
  DDPB_CHECK_CU(dev_get(&cudev, dev));
  DDPB_CHECK_CU(add_dev(static_cast<CUmemGenericAllocationHandle>(mc_handle), cudev));
  return 0;
}

// Every rank: bind its local allocation to the multicast object at offset 0.
extern "C" int ddpb_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes) {
  DDPB_DRV(bind, "cuMulticastBindMem", CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
           unsigned long long);
  DDPB_CHECK_CU(bind(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0,
                     static_cast<CUmemGenericAllocationHandle>(mem_handle), 0, static_cast<size_t>(bytes), 0));
  return 0;
}
