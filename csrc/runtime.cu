// Library runtime: error slot, launch counter, peer-shareable device arenas (CUDA IPC).
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "common.cuh"

namespace b200dp {

std::atomic<unsigned long long> g_launches{0};
static bool pdl_default() {
  const char* v = getenv("B200DP_PDL");
  return v != nullptr && atoi(v) != 0;     // measured: no gain inside a CUDA graph (7.86 vs 7.75 ms/step), so off by default
}
bool g_pdl = pdl_default();

namespace {
thread_local char t_err[1024] = "";
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

}  // namespace b200dp

using namespace b200dp;

extern "C" const char* b200dp_last_error(void) { return t_err; }
extern "C" int b200dp_version(void) { return 100; }
extern "C" void b200dp_set_pdl(int on) { g_pdl = on != 0; }
extern "C" int b200dp_get_pdl(void) { return g_pdl ? 1 : 0; }
extern "C" unsigned long long b200dp_launch_count(void) { return g_launches.load(); }
extern "C" void b200dp_reset_launch_count(void) { g_launches.store(0); }

extern "C" int b200dp_arena_alloc(size_t bytes, void** out_ptr) {
  B200DP_REQUIRE(out_ptr != nullptr && bytes > 0, "arena_alloc: bad arguments");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    set_error("arena_alloc: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    return B200DP_ENOMEM;
  }
  B200DP_CUDA(cudaMemset(p, 0, bytes));
  B200DP_CUDA(cudaDeviceSynchronize());
  *out_ptr = p;
  return B200DP_OK;
}

extern "C" int b200dp_arena_free(void* ptr) {
  if (ptr) B200DP_CUDA(cudaFree(ptr));
  return B200DP_OK;
}

extern "C" int b200dp_ipc_export(void* ptr, unsigned char handle_out[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  B200DP_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out, &h, 64);
  return B200DP_OK;
}

extern "C" int b200dp_ipc_open(const unsigned char handle[64], void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  B200DP_CUDA(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200DP_OK;
}

extern "C" int b200dp_ipc_close(void* ptr) {
  if (ptr) B200DP_CUDA(cudaIpcCloseMemHandle(ptr));
  return B200DP_OK;
}

extern "C" int b200dp_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return B200DP_OK;
  }
  B200DP_CUDA(e);
  return B200DP_OK;
}
