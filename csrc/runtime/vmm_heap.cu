// Symmetric heap on the CUDA virtual-memory-management driver API, with an NVSwitch multicast alias.
//
// The reference moves every operand through Horovod / MPI collectives (multiplication/functions.py:95,143,206);
// here the fused kernels store straight into their peers' memory.  This file provides that memory:
//
//   * ddp_vmm_alloc   - cuMemCreate (pinned device memory, exportable as a POSIX file descriptor) + map + zero
//   * ddp_vmm_open    - map a peer's allocation from the descriptor it sent us (cuMemImportFromShareableHandle)
//   * ddp_mc_*        - one multicast object per heap (cuMulticastCreate on rank 0, every rank adds its device and
//                       binds its own allocation): a store to the multicast address (multimem.st in the pusher
//                       warps) is replicated by the switch into EVERY rank's copy, so a gathered byte leaves the
//                       GPU once instead of N-1 times
//
// The descriptors travel between the ranks' processes over unix-domain sockets (parallel/runtime.py).  Every entry
// point reports the failing driver call and its CUresult through ddp_last_error(), so a box without descriptor
// export or multicast support falls back to the cudaMalloc + CUDA-IPC heap of runtime.cu with the reason on record.
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "../../include/ddp_b200.h"
#include "../common/host_util.h"

namespace ddp {
namespace {

struct Driver {
  bool ok = false;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
};

template <typename F>
bool resolve(F& fn, const char* name, int version) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult st;
  if (cudaGetDriverEntryPointByVersion(name, &p, version, cudaEnableDefault, &st) != cudaSuccess ||
      st != cudaDriverEntryPointSuccess || p == nullptr)
    return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Driver& driver() {
  static Driver d;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFree(nullptr);  // a current context on this thread
    bool ok = true;
    ok &= resolve(d.MemCreate, "cuMemCreate", 10020);
    ok &= resolve(d.MemRelease, "cuMemRelease", 10020);
    ok &= resolve(d.MemAddressReserve, "cuMemAddressReserve", 10020);
    ok &= resolve(d.MemAddressFree, "cuMemAddressFree", 10020);
    ok &= resolve(d.MemMap, "cuMemMap", 10020);
    ok &= resolve(d.MemUnmap, "cuMemUnmap", 10020);
    ok &= resolve(d.MemSetAccess, "cuMemSetAccess", 10020);
    ok &= resolve(d.MemExport, "cuMemExportToShareableHandle", 10020);
    ok &= resolve(d.MemImport, "cuMemImportFromShareableHandle", 10020);
    ok &= resolve(d.MemGetGranularity, "cuMemGetAllocationGranularity", 10020);
    ok &= resolve(d.DeviceGet, "cuDeviceGet", 2000);
    ok &= resolve(d.DeviceGetAttribute, "cuDeviceGetAttribute", 2000);
    // multicast objects exist since CUDA 12.1; their absence only disables the multicast alias
    bool mc = true;
    mc &= resolve(d.McCreate, "cuMulticastCreate", 12010);
    mc &= resolve(d.McAddDevice, "cuMulticastAddDevice", 12010);
    mc &= resolve(d.McBindMem, "cuMulticastBindMem", 12010);
    mc &= resolve(d.McGetGranularity, "cuMulticastGetGranularity", 12010);
    if (!mc) d.McCreate = nullptr;
    d.ok = ok;
  });
  return d;
}

int cu_check(CUresult r, const char* what) {
  if (r == CUDA_SUCCESS) return 0;
  char buf[160];
  snprintf(buf, sizeof buf, "%s failed with CUresult %d", what, static_cast<int>(r));
  return set_error(-200 - static_cast<int>(r), buf);
}

struct Segment {
  CUmemGenericAllocationHandle handle;
  size_t size;
};
std::mutex g_mu;
std::map<CUdeviceptr, Segment> g_segments;

int current_device(CUdevice* dev, int* ordinal) {
  int o = 0;
  if (int e = cuda_check(cudaGetDevice(&o), "cudaGetDevice")) return e;
  if (ordinal) *ordinal = o;
  return cu_check(driver().DeviceGet(dev, o), "cuDeviceGet");
}

CUmemAllocationProp device_prop(int ordinal) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof prop);
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = ordinal;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// reserve a range, map `handle` into it and give the current device read/write access
int map_handle(CUmemGenericAllocationHandle handle, size_t size, size_t align, int ordinal, CUdeviceptr* va_out) {
  Driver& d = driver();
  CUdeviceptr va = 0;
  if (int e = cu_check(d.MemAddressReserve(&va, size, align, 0, 0), "cuMemAddressReserve")) return e;
  if (int e = cu_check(d.MemMap(va, size, 0, handle, 0), "cuMemMap")) {
    d.MemAddressFree(va, size);
    return e;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof acc);
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if (int e = cu_check(d.MemSetAccess(va, size, &acc, 1), "cuMemSetAccess")) {
    d.MemUnmap(va, size);
    d.MemAddressFree(va, size);
    return e;
  }
  *va_out = va;
  return 0;
}

}  // namespace
}  // namespace ddp

using namespace ddp;

// What this device and driver can do: descriptors for cuMemCreate allocations, multicast objects, and the
// allocation granularity that satisfies both (sizes passed to the calls below must be multiples of it).
extern "C" int ddp_vmm_query(int* fd_handles, int* multicast, size_t* granularity) {
  Driver& d = driver();
  if (fd_handles) *fd_handles = 0;
  if (multicast) *multicast = 0;
  if (granularity) *granularity = 2u << 20;
  if (!d.ok) return set_error(-10, "the CUDA driver does not export the virtual-memory-management entry points");
  CUdevice dev;
  int ord = 0;
  if (int e = current_device(&dev, &ord)) return e;
  int v = 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS &&
      fd_handles)
    *fd_handles = v;
  int mc = 0;
  if (d.McCreate != nullptr) d.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
  if (multicast) *multicast = mc;
  CUmemAllocationProp prop = device_prop(ord);
  size_t g = 0;
  if (int e = cu_check(d.MemGetGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
                       "cuMemGetAllocationGranularity"))
    return e;
  if (mc) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = 2;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.McGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  if (granularity) *granularity = g;
  return 0;
}

extern "C" int ddp_vmm_alloc(size_t bytes, void** ptr, int* fd_out) {
  Driver& d = driver();
  if (!d.ok) return set_error(-10, "the CUDA driver does not export the virtual-memory-management entry points");
  CUdevice dev;
  int ord = 0;
  if (int e = current_device(&dev, &ord)) return e;
  size_t gran = 0;
  if (int e = ddp_vmm_query(nullptr, nullptr, &gran)) return e;
  if (bytes == 0 || bytes % gran) return set_error(-4, "ddp_vmm_alloc: size must be a positive multiple of the granularity");
  CUmemAllocationProp prop = device_prop(ord);
  CUmemGenericAllocationHandle h;
  if (int e = cu_check(d.MemCreate(&h, bytes, &prop, 0), "cuMemCreate")) return e;
  CUdeviceptr va = 0;
  if (int e = map_handle(h, bytes, gran, ord, &va)) {
    d.MemRelease(h);
    return e;
  }
  if (int e = cuda_check(cudaMemset(reinterpret_cast<void*>(va), 0, bytes), "cudaMemset(symmetric heap)")) return e;
  if (fd_out) {
    int fd = -1;
    if (int e = cu_check(d.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
                         "cuMemExportToShareableHandle")) {
      d.MemUnmap(va, bytes);
      d.MemAddressFree(va, bytes);
      d.MemRelease(h);
      return e;
    }
    *fd_out = fd;
  }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_segments[va] = Segment{h, bytes};
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

// Maps the allocation behind a descriptor received from a peer process (the caller closes the descriptor).
extern "C" int ddp_vmm_open(int fd, size_t bytes, void** ptr) {
  Driver& d = driver();
  if (!d.ok) return set_error(-10, "the CUDA driver does not export the virtual-memory-management entry points");
  CUdevice dev;
  int ord = 0;
  if (int e = current_device(&dev, &ord)) return e;
  size_t gran = 0;
  if (int e = ddp_vmm_query(nullptr, nullptr, &gran)) return e;
  CUmemGenericAllocationHandle h;
  if (int e = cu_check(d.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                       "cuMemImportFromShareableHandle"))
    return e;
  CUdeviceptr va = 0;
  if (int e = map_handle(h, bytes, gran, ord, &va)) {
    d.MemRelease(h);
    return e;
  }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_segments[va] = Segment{h, bytes};
  }
  *ptr = reinterpret_cast<void*>(va);
  return 0;
}

// Unmaps a range obtained from ddp_vmm_alloc / ddp_vmm_open / ddp_mc_bind and drops its handle.
extern "C" int ddp_vmm_release(void* ptr) {
  Driver& d = driver();
  Segment seg;
  const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(ptr);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_segments.find(va);
    if (it == g_segments.end()) return set_error(-4, "ddp_vmm_release: unknown address");
    seg = it->second;
    g_segments.erase(it);
  }
  int rc = cu_check(d.MemUnmap(va, seg.size), "cuMemUnmap");
  if (int e = cu_check(d.MemAddressFree(va, seg.size), "cuMemAddressFree")) rc = rc ? rc : e;
  if (int e = cu_check(d.MemRelease(seg.handle), "cuMemRelease")) rc = rc ? rc : e;
  return rc;
}

// ---------------------------------------------------------------------------------------------- multicast
// Protocol (every step collective, driven by parallel/runtime.py):
//   rank 0: ddp_mc_create(size, world) -> (object, descriptor); the descriptor goes to every other rank
//   others: ddp_mc_import(descriptor)  -> object
//   all:    ddp_mc_add_device(object); barrier (binding needs every device added);
//           ddp_mc_bind(object, own heap address, size) -> multicast address; barrier
// The multicast address covers the same byte offsets as every rank's heap.
extern "C" int ddp_mc_create(size_t bytes, int n_devices, unsigned long long* object, int* fd_out) {
  Driver& d = driver();
  if (!d.ok || d.McCreate == nullptr) return set_error(-10, "the CUDA driver does not export cuMulticastCreate");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof mp);
  mp.numDevices = static_cast<unsigned int>(n_devices);
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  if (int e = cu_check(d.McCreate(&h, &mp), "cuMulticastCreate")) return e;
  int fd = -1;
  if (int e = cu_check(d.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
                       "cuMemExportToShareableHandle(multicast object)")) {
    d.MemRelease(h);
    return e;
  }
  *object = static_cast<unsigned long long>(h);
  *fd_out = fd;
  return 0;
}

extern "C" int ddp_mc_import(int fd, unsigned long long* object) {
  Driver& d = driver();
  if (!d.ok || d.McCreate == nullptr) return set_error(-10, "the CUDA driver does not export cuMulticastCreate");
  CUmemGenericAllocationHandle h;
  if (int e = cu_check(d.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                       "cuMemImportFromShareableHandle(multicast object)"))
    return e;
  *object = static_cast<unsigned long long>(h);
  return 0;
}

extern "C" int ddp_mc_add_device(unsigned long long object) {
  Driver& d = driver();
  CUdevice dev;
  if (int e = current_device(&dev, nullptr)) return e;
  return cu_check(d.McAddDevice(static_cast<CUmemGenericAllocationHandle>(object), dev), "cuMulticastAddDevice");
}

// Binds this rank's heap (an address returned by ddp_vmm_alloc) to the multicast object and maps the object:
// *mc_ptr + offset addresses byte `offset` of EVERY rank's heap at once (stores and reductions only).
extern "C" int ddp_mc_bind(unsigned long long object, void* heap_ptr, size_t bytes, void** mc_ptr) {
  Driver& d = driver();
  int ord = 0;
  CUdevice dev;
  if (int e = current_device(&dev, &ord)) return e;
  Segment seg;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_segments.find(reinterpret_cast<CUdeviceptr>(heap_ptr));
    if (it == g_segments.end()) return set_error(-4, "ddp_mc_bind: the heap was not allocated by ddp_vmm_alloc");
    seg = it->second;
  }
  if (bytes != seg.size) return set_error(-4, "ddp_mc_bind: size differs from the heap's");
  const CUmemGenericAllocationHandle mc = static_cast<CUmemGenericAllocationHandle>(object);
  if (int e = cu_check(d.McBindMem(mc, 0, seg.handle, 0, bytes, 0), "cuMulticastBindMem")) return e;
  size_t gran = 0;
  if (int e = ddp_vmm_query(nullptr, nullptr, &gran)) return e;
  CUdeviceptr va = 0;
  if (int e = map_handle(mc, bytes, gran, ord, &va)) return e;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_segments[va] = Segment{mc, bytes};
  }
  *mc_ptr = reinterpret_cast<void*>(va);
  return 0;
}

extern "C" int ddp_close_fd(int fd) { return close(fd) == 0 ? 0 : set_error(-4, "close(fd) failed"); }
