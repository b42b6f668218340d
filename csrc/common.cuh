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

// SM budget: launches made while a budget is set size their grids for that many SMs instead of the whole device, so that
// two branches of the step graph can share the machine without queueing behind each other's persistent CTAs.
inline int& sm_budget() {
  static thread_local int b = 0;
  return b;
}
inline int num_sms() {
  if (sm_budget() > 0) return sm_budget();
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

// ---- packed fp32 pairs (Blackwell FFMA2 / FADD2): one issue slot for two channels -------------------------------
typedef unsigned long long f2;
__device__ __forceinline__ f2 f2_make(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2 f2_from_bf16x2(uint32_t v) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(v << 16), "r"(v & 0xffff0000u));
  return r;
}
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
  f2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// the same on two scalar FFMAs (experiment: FFMA2 may issue on one of the two FMA pipes only)
__device__ __forceinline__ f2 f2_fma_s(f2 a, f2 b, f2 c) {
  float alo, ahi, blo, bhi, clo, chi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(alo), "=f"(ahi) : "l"(a));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(blo), "=f"(bhi) : "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(clo), "=f"(chi) : "l"(c));
  f2 d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(fmaf(alo, blo, clo)), "f"(fmaf(ahi, bhi, chi)));
  return d;
}
#ifdef B200DP_DW_SCALAR_FMA   // measured: forward equal, backward 8-10 % slower than FFMA2
#define F2_STENCIL_FMA f2_fma_s
#else
#define F2_STENCIL_FMA f2_fma
#endif
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) {
  f2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint32_t f2_to_bf16x2(f2 a) {
  uint32_t r;
  asm("{\n\t.reg .b32 lo, hi;\n\tmov.b64 {lo, hi}, %1;\n\tcvt.rn.bf16x2.f32 %0, hi, lo;\n\t}" : "=r"(r) : "l"(a));
  return r;
}
__device__ __forceinline__ uint32_t f2_to_bf16x2_relu(f2 a) {        // max(x, 0) folded into the conversion
  uint32_t r;
  asm("{\n\t.reg .b32 lo, hi;\n\tmov.b64 {lo, hi}, %1;\n\tcvt.rn.relu.bf16x2.f32 %0, hi, lo;\n\t}" : "=r"(r) : "l"(a));
  return r;
}
// clamp to [0, 6] on the packed bf16 pair (both bounds are exact in bf16 and rounding is monotonic, so this equals
// rounding the clamped fp32 value).  max / min return the non-NaN operand: a NaN input becomes 0.
__device__ __forceinline__ uint32_t bf16x2_relu6(uint32_t v) {
  uint32_t r;
  asm("{\n\t.reg .b32 t;\n\tmax.bf16x2 t, %1, %2;\n\tmin.bf16x2 %0, t, %3;\n\t}" : "=r"(r) : "r"(v), "r"(0u), "r"(0x40C040C0u));
  return r;
}
__device__ __forceinline__ f2 f2_pair(uint32_t lo_bits, uint32_t hi_bits) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo_bits), "r"(hi_bits));
  return r;
}
__device__ __forceinline__ void f2_split(f2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// NaN -> 0, everything else unchanged: max / min return the non-NaN operand, and one of max(v,0), min(v,0) is zero
__device__ __forceinline__ uint32_t bf16x2_nan0(uint32_t v) {
  uint32_t r;
  asm("{\n\t.reg .b32 p, q;\n\tmax.bf16x2 p, %1, %2;\n\tmin.bf16x2 q, %1, %2;\n\tor.b32 %0, p, q;\n\t}" : "=r"(r) : "r"(v), "r"(0u));
  return r;
}

#endif  // __CUDACC__

}  // namespace b200dp
