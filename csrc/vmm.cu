// Symmetric device memory for the communication kernels: CUDA virtual-memory-management allocations that peers map
// through POSIX file descriptors, and an NVSwitch multicast object bound over the same physical memory (NVLS), so
// that comm_adam.cu can use `multimem.ld_reduce` / `multimem.st` next to plain peer loads and stores.
//
// Reference behaviour replaced: the NCCL communicator that DistributedDataParallel builds at
// cifar10_mpi_mobilenet_224.py:142-145 (NCCL allocates and registers the same kind of memory internally).
//
// Protocol (one process per GPU, driven by parallel/comm.py):
//   every rank:  vmm_alloc(bytes)            -> local pointer + fd          (fd sent to every peer over AF_UNIX)
//   every rank:  vmm_import(fd of peer p)    -> pointer to p's arena as mapped here
//   rank 0:      mc_create(world, bytes)     -> multicast object + fd       (fd sent to every peer)
//   others:      mc_import(fd)
//   every rank:  mc_add_device(); <host barrier>; mc_bind(own allocation); <host barrier>; mc_map() -> multicast pointer
// The driver API is reached through cudaGetDriverEntryPoint, so the library keeps linking without libcuda.
// Only `multimem.*` instructions and plain ld/st ever touch these mappings: no TMA descriptor is built over a peer
// or multicast address anywhere in this library.
#include <cuda.h>
#include <unistd.h>

#include <map>
#include <mutex>

#include "common.cuh"

namespace b200dp {
namespace {

struct Drv {
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
  decltype(&cuMemAddressFree) MemAddressFree = nullptr;
  decltype(&cuMemMap) MemMap = nullptr;
  decltype(&cuMemUnmap) MemUnmap = nullptr;
  decltype(&cuMemSetAccess) MemSetAccess = nullptr;
  decltype(&cuMemGetAllocationGranularity) MemGetAllocationGranularity = nullptr;
  decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle = nullptr;
  decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle = nullptr;
  decltype(&cuMulticastCreate) MulticastCreate = nullptr;
  decltype(&cuMulticastAddDevice) MulticastAddDevice = nullptr;
  decltype(&cuMulticastBindMem) MulticastBindMem = nullptr;
  decltype(&cuMulticastUnbind) MulticastUnbind = nullptr;
  decltype(&cuMulticastGetGranularity) MulticastGetGranularity = nullptr;
  decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
  decltype(&cuGetErrorString) GetErrorString = nullptr;
  bool ok = false;
};

template <typename F>
bool fetch(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      fn == nullptr) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}

Drv& drv() {
  static Drv d;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFree(nullptr);   // make sure the primary context exists
    bool ok = true;
    ok &= fetch("cuMemCreate", &d.MemCreate);
    ok &= fetch("cuMemRelease", &d.MemRelease);
    ok &= fetch("cuMemAddressReserve", &d.MemAddressReserve);
    ok &= fetch("cuMemAddressFree", &d.MemAddressFree);
    ok &= fetch("cuMemMap", &d.MemMap);
    ok &= fetch("cuMemUnmap", &d.MemUnmap);
    ok &= fetch("cuMemSetAccess", &d.MemSetAccess);
    ok &= fetch("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
    ok &= fetch("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
    ok &= fetch("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
    ok &= fetch("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    ok &= fetch("cuGetErrorString", &d.GetErrorString);
    // multicast is optional (needs NVSwitch + a driver with NVLS enabled)
    fetch("cuMulticastCreate", &d.MulticastCreate);
    fetch("cuMulticastAddDevice", &d.MulticastAddDevice);
    fetch("cuMulticastBindMem", &d.MulticastBindMem);
    fetch("cuMulticastUnbind", &d.MulticastUnbind);
    fetch("cuMulticastGetGranularity", &d.MulticastGetGranularity);
    d.ok = ok;
  });
  return d;
}

int drv_fail(CUresult r, const char* what) {
  const char* msg = nullptr;
  if (drv().GetErrorString) drv().GetErrorString(r, &msg);
  set_error("%s failed: %s (%d)", what, msg ? msg : "?", (int)r);
  return B200DP_ECUDA;
}

#define DRV(call, what)                                  \
  do {                                                   \
    CUresult r__ = (call);                               \
    if (r__ != CUDA_SUCCESS) return drv_fail(r__, what); \
  } while (0)

struct Mapping {
  CUmemGenericAllocationHandle handle;
  size_t bytes;
  bool multicast;
};
std::mutex g_mu;
std::map<CUdeviceptr, Mapping> g_maps;            // by mapped address
std::map<unsigned long long, size_t> g_mc_objects;   // multicast handle -> bytes

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

CUmemAllocationProp local_prop(int dev) {
  CUmemAllocationProp p{};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_handle(CUmemGenericAllocationHandle h, size_t bytes, bool multicast, void** out) {
  Drv& d = drv();
  CUdeviceptr va = 0;
  DRV(d.MemAddressReserve(&va, bytes, 0, 0, 0), "cuMemAddressReserve");
  CUresult r = d.MemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(va, bytes);
    return drv_fail(r, "cuMemMap");
  }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = current_device();
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.MemUnmap(va, bytes);
    d.MemAddressFree(va, bytes);
    return drv_fail(r, "cuMemSetAccess");
  }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_maps[va] = Mapping{h, bytes, multicast};
  }
  *out = reinterpret_cast<void*>(va);
  return B200DP_OK;
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" {

// flags_out bit 0: VMM allocations with POSIX-fd export are available; bit 1: multicast (NVLS) is available.
// granularity_out: allocation sizes must be multiples of this (covers both the memory and the multicast granularity).
int b200dp_vmm_query(int* flags_out, size_t* granularity_out) {
  B200DP_REQUIRE(flags_out && granularity_out, "vmm_query: null argument");
  *flags_out = 0;
  *granularity_out = 2u << 20;
  Drv& d = drv();
  if (!d.ok) return B200DP_OK;
  const int dev = current_device();
  int vmm = 0, fd = 0, mc = 0;
  d.DeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
  d.DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
  if (d.MulticastCreate && d.MulticastAddDevice && d.MulticastBindMem && d.MulticastGetGranularity)
    d.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  size_t g = 0;
  CUmemAllocationProp p = local_prop(dev);
  if (vmm && fd && d.MemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) {
    *flags_out |= 1;
    if (g > *granularity_out) *granularity_out = g;
  }
  if ((*flags_out & 1) && mc) {
    CUmulticastObjectProp mp{};
    mp.numDevices = 2;
    mp.size = *granularity_out;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) {
      *flags_out |= 2;
      if (mg > *granularity_out) *granularity_out = mg;
    }
  }
  return B200DP_OK;
}

// Allocate `bytes` (a multiple of the granularity) on the current device, map it, zero it, and export a file
// descriptor that another process can import.  `handle_out` identifies the allocation for mc_bind.
int b200dp_vmm_alloc(size_t bytes, void** ptr_out, int* fd_out, unsigned long long* handle_out) {
  B200DP_REQUIRE(ptr_out && fd_out && handle_out && bytes > 0, "vmm_alloc: bad arguments");
  Drv& d = drv();
  B200DP_REQUIRE(d.ok, "vmm_alloc: driver entry points unavailable");
  const int dev = current_device();
  CUmemAllocationProp p = local_prop(dev);
  CUmemGenericAllocationHandle h;
  DRV(d.MemCreate(&h, bytes, &p, 0), "cuMemCreate");
  int rc = map_handle(h, bytes, false, ptr_out);
  if (rc != B200DP_OK) {
    d.MemRelease(h);
    return rc;
  }
  int fd = -1;
  DRV(d.MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
  B200DP_CUDA(cudaMemset(*ptr_out, 0, bytes));
  B200DP_CUDA(cudaDeviceSynchronize());
  *fd_out = fd;
  *handle_out = (unsigned long long)h;
  return B200DP_OK;
}

// Map a peer's allocation (file descriptor received from that process) into this process for the current device.
int b200dp_vmm_import(int fd, size_t bytes, void** ptr_out) {
  B200DP_REQUIRE(ptr_out && fd >= 0 && bytes > 0, "vmm_import: bad arguments");
  Drv& d = drv();
  B200DP_REQUIRE(d.ok, "vmm_import: driver entry points unavailable");
  CUmemGenericAllocationHandle h;
  DRV(d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
      "cuMemImportFromShareableHandle");
  int rc = map_handle(h, bytes, false, ptr_out);
  if (rc != B200DP_OK) d.MemRelease(h);
  return rc;
}

// Unmap a pointer returned by vmm_alloc / vmm_import / mc_map and drop its handle.
int b200dp_vmm_free(void* ptr) {
  if (!ptr) return B200DP_OK;
  Drv& d = drv();
  Mapping m;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_maps.find((CUdeviceptr)ptr);
    B200DP_REQUIRE(it != g_maps.end(), "vmm_free: unknown pointer %p", ptr);
    m = it->second;
    g_maps.erase(it);
  }
  cudaDeviceSynchronize();
  d.MemUnmap((CUdeviceptr)ptr, m.bytes);
  d.MemAddressFree((CUdeviceptr)ptr, m.bytes);
  if (!m.multicast) d.MemRelease(m.handle);     // multicast objects are released by mc_release
  return B200DP_OK;
}

// Rank 0: create the multicast object for `world` devices and export its file descriptor.
int b200dp_mc_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out) {
  B200DP_REQUIRE(mc_out && fd_out && world >= 2, "mc_create: bad arguments");
  Drv& d = drv();
  B200DP_REQUIRE(d.ok && d.MulticastCreate, "mc_create: multicast entry points unavailable");
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)world;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  DRV(d.MulticastCreate(&h, &mp), "cuMulticastCreate");
  int fd = -1;
  DRV(d.MemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
      "cuMemExportToShareableHandle(multicast)");
  *mc_out = (unsigned long long)h;
  *fd_out = fd;
  std::lock_guard<std::mutex> lk(g_mu);
  g_mc_objects[*mc_out] = bytes;
  return B200DP_OK;
}

int b200dp_mc_import(int fd, size_t bytes, unsigned long long* mc_out) {
  B200DP_REQUIRE(mc_out && fd >= 0, "mc_import: bad arguments");
  Drv& d = drv();
  B200DP_REQUIRE(d.ok && d.MulticastCreate, "mc_import: multicast entry points unavailable");
  CUmemGenericAllocationHandle h;
  DRV(d.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
      "cuMemImportFromShareableHandle(multicast)");
  *mc_out = (unsigned long long)h;
  std::lock_guard<std::mutex> lk(g_mu);
  g_mc_objects[*mc_out] = bytes;
  return B200DP_OK;
}

// Every rank, before anyone binds: add the current device to the multicast team.
int b200dp_mc_add_device(unsigned long long mc) {
  Drv& d = drv();
  DRV(d.MulticastAddDevice((CUmemGenericAllocationHandle)mc, (CUdevice)current_device()), "cuMulticastAddDevice");
  return B200DP_OK;
}

// After ALL ranks have added their device: bind this rank's allocation (handle from vmm_alloc) at offset 0.
int b200dp_mc_bind(unsigned long long mc, unsigned long long mem_handle, size_t bytes) {
  Drv& d = drv();
  DRV(d.MulticastBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem_handle, 0, bytes, 0),
      "cuMulticastBindMem");
  return B200DP_OK;
}

// Map the multicast object: stores through the returned pointer land in every rank's bound allocation,
// `multimem.ld_reduce` through it returns the element-wise reduction over all of them.
int b200dp_mc_map(unsigned long long mc, size_t bytes, void** ptr_out) {
  B200DP_REQUIRE(ptr_out, "mc_map: null argument");
  return map_handle((CUmemGenericAllocationHandle)mc, bytes, true, ptr_out);
}

int b200dp_mc_release(unsigned long long mc) {
  Drv& d = drv();
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_mc_objects.erase(mc);
  }
  if (d.MemRelease) d.MemRelease((CUmemGenericAllocationHandle)mc);
  return B200DP_OK;
}

int b200dp_close_fd(int fd) {
  if (fd >= 0) close(fd);
  return B200DP_OK;
}

}  // extern "C"
