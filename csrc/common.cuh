// Shared host/device helpers for the b200dp native library (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../include/b200dp.h"

namespace b200dp {

// ---- error reporting ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;

inline int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("%s failed at %s:%d: %s", what, file, line, cudaGetErrorString(e));
  return B200DP_ECUDA;
}

#define B200DP_CUDA(call)                                                        \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) return ::b200dp::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

#define B200DP_CHECK_LAUNCH()                                                    \
  do {                                                                           \
    ::b200dp::g_launches.fetch_add(1, std::memory_order_relaxed);                \
    cudaError_t e__ = cudaGetLastError();                                        \
    if (e__ != cudaSuccess) return ::b200dp::cuda_fail(e__, "kernel launch", __FILE__, __LINE__); \
  } while (0)

#define B200DP_REQUIRE(cond, ...)                                                \
  do {                                                                           \
    if (!(cond)) {                                                               \
      ::b200dp::set_error(__VA_ARGS__);                                          \
      return B200DP_EINVAL;                                                      \
    }                                                                            \
  } while (0)

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------
// Every kernel of the training step is launched with the programmatic-stream-serialisation attribute: its CTAs may
// be scheduled (and run their shared-memory / TMEM set-up) while the previous kernel drains, and block in
// `griddepcontrol.wait` until that kernel has completed and flushed.  Opt-in (B200DP_PDL=1): inside the captured
// CUDA graph the launch gaps are already ~1 us, and the measured step time did not improve.
extern bool g_pdl;

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

// ---- device helpers ------------------------------------------------------------------------------
#ifdef __CUDACC__

// PDL: let the next kernel's CTAs start launching / wait for the previous kernel's results (no-ops without the
// launch attribute).  pdl_wait() must precede the first access to anything a predecessor wrote.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

__device__ __forceinline__ float relu6f(float x) { return fminf(fmaxf(x, 0.0f), 6.0f); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// streaming 16-byte global accesses (data touched once: keep it out of L1)
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

#endif  // __CUDACC__

}  // namespace b200dp
