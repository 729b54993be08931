// Runtime pieces of libddp_b200.so: error reporting, device queries and the symmetric heap that the
// fused collective kernels address through peer pointers (CUDA IPC over NVLink / NVSwitch).
// Replaces the reference's Horovod/MPI substrate (reference: distributed_dot_product/utils/comm.py:1-30).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/ddp_b200.h"
#include "../common/host_util.h"

namespace ddp {
static thread_local std::string g_err;
int set_error(int code, const char* msg) {
  g_err = msg ? msg : "";
  return code;
}
int cuda_check(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  std::string m = std::string(what) + ": " + cudaGetErrorString(e);
  return set_error(-100 - static_cast<int>(e), m.c_str());
}
}  // namespace ddp

using namespace ddp;

extern "C" const char* ddp_last_error(void) { return g_err.c_str(); }
extern "C" int ddp_version(void) { return 100; }

extern "C" int ddp_num_sms(void) {
  static int sms[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!sms[dev]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

extern "C" int ddp_heap_alloc(size_t bytes, void** ptr, void* ipc_handle_out64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  if (int e = cuda_check(cudaMalloc(&p, bytes), "cudaMalloc(symmetric heap)")) return e;
  if (int e = cuda_check(cudaMemset(p, 0, bytes), "cudaMemset(symmetric heap)")) return e;
  if (ipc_handle_out64) {
    cudaIpcMemHandle_t h;
    if (int e = cuda_check(cudaIpcGetMemHandle(&h, p), "cudaIpcGetMemHandle")) {
      cudaFree(p);
      return e;
    }
    memcpy(ipc_handle_out64, &h, 64);
  }
  *ptr = p;
  return 0;
}

extern "C" int ddp_heap_open(const void* ipc_handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle64, 64);
  return cuda_check(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
}

extern "C" int ddp_heap_close(void* ptr) { return cuda_check(cudaIpcCloseMemHandle(ptr), "cudaIpcCloseMemHandle"); }
extern "C" int ddp_heap_free(void* ptr) { return cuda_check(cudaFree(ptr), "cudaFree"); }
extern "C" int ddp_memset_async(void* ptr, int value, size_t bytes, void* stream) {
  return cuda_check(cudaMemsetAsync(ptr, value, bytes, static_cast<cudaStream_t>(stream)), "cudaMemsetAsync");
}

extern "C" int ddp_read_u32(const void* dev_ptr, uint32_t* out) {
  return cuda_check(cudaMemcpy(out, dev_ptr, sizeof(uint32_t), cudaMemcpyDeviceToHost), "cudaMemcpy(status word)");
}

// Stream-ordered copy of a device word into PINNED host memory: the Python runtime polls the sticky status word
// this way every few collective calls without ever synchronising the device (parallel/runtime.py).
extern "C" int ddp_read_u32_async(const void* dev_ptr, uint32_t* pinned_host, void* stream) {
  return cuda_check(cudaMemcpyAsync(pinned_host, dev_ptr, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                    static_cast<cudaStream_t>(stream)),
                    "cudaMemcpyAsync(status word)");
}
