// Experiment: can a tcgen05.mma shared-memory descriptor start at an arbitrary ROW of a tile?
// (needed to run a 3x3 depthwise conv as nine tap-shifted MMAs over one haloed tile)
//   layout 0: K-major SWIZZLE_128B rows of 128 B, start shifted by s rows, base_offset field = 0
//   layout 1: same, base_offset = (addr >> 7) & 7
//   layout 2: no swizzle, 8-channel planes of 16-byte rows (core matrix = 8 rows x 16 B), LBO = plane pitch, SBO = 128
//   layout 3: as 2 with LBO / SBO swapped
// B is the 64x64 identity, so D[r][n] must equal A[r + s][n].
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -Icsrc experiments/exp_desc_shift.cu -o /tmp/exp_desc_shift
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>

#include "sm100.cuh"

using namespace b200dp::sm100;

constexpr int ROWS = 272;
constexpr int PLANE = ROWS * 16 + 16;   // plane pitch (odd multiple of 16 B)

__device__ __forceinline__ float aval(int r, int k) { return (float)(((r * 64 + k) % 251) - 125); }

__global__ void __launch_bounds__(128, 1) exp_kernel(int layout, int shift, int* mismatches, float* sample) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                        // 272 * 128 = 34816 B (or 8 planes)
  uint8_t* sB = smem + 40 * 1024;            // 64 * 128 = 8192 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid / 32;
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(sA);
  __nv_bfloat16* Bm = reinterpret_cast<__nv_bfloat16*>(sB);
  for (int i = tid; i < ROWS * 64; i += 128) {
    const int r = i / 64, k = i % 64;
    size_t off;
    if (layout < 2) off = (size_t)r * 128 + (((k / 8) ^ (r % 8)) * 16) + (k % 8) * 2;
    else if (layout < 4) off = (size_t)(k / 8) * PLANE + (size_t)r * 16 + (k % 8) * 2;
    else if (layout == 4) { if (k >= 32) continue; off = (size_t)r * 64 + (((k / 8) ^ ((r >> 1) & 3)) * 16) + (k % 8) * 2; }
    else { if (k >= 16) continue; off = (size_t)r * 32 + (((k / 8) ^ ((r >> 2) & 1)) * 16) + (k % 8) * 2; }
    A[off / 2] = __float2bfloat16(aval(r, k));
  }
  for (int i = tid; i < 64 * 64; i += 128) {
    const int n = i / 64, k = i % 64;
    size_t off;
    if (layout < 2) off = (size_t)n * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2;
    else if (layout < 4) off = (size_t)(k / 8) * 1024 + (size_t)n * 16 + (k % 8) * 2;
    else if (layout == 4) { if (k >= 32) continue; off = (size_t)n * 64 + (((k / 8) ^ ((n >> 1) & 3)) * 16) + (k % 8) * 2; }
    else { if (k >= 16) continue; off = (size_t)n * 32 + (((k / 8) ^ ((n >> 2) & 1)) * 16) + (k % 8) * 2; }
    Bm[off / 2] = __float2bfloat16(n == k ? 1.f : 0.f);
  }
  if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 64);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
    const int ksteps = layout == 4 ? 2 : (layout == 5 ? 1 : 4);
    for (int j = 0; j < ksteps; ++j) {
      uint64_t da, db;
      if (layout >= 4) {
        const int w = layout == 4 ? 32 : 16;
        da = umma_desc_kmajor(smem_u32(sA) + shift * 2 * w + j * 32, w);
        db = umma_desc_kmajor(smem_u32(sB) + j * 32, w);
      } else if (layout < 2) {
        const uint32_t a_addr = smem_u32(sA) + shift * 128 + j * 32;
        da = umma_desc_kmajor_sw128(a_addr);
        if (layout == 1) da |= (uint64_t)((a_addr >> 7) & 7) << 49;
        db = umma_desc_kmajor_sw128(smem_u32(sB) + j * 32);
      } else {
        const uint32_t a_addr = smem_u32(sA) + (2 * j) * PLANE + shift * 16;
        const uint32_t b_addr = smem_u32(sB) + (2 * j) * 1024;
        const uint64_t lboA = PLANE >> 4, sboA = 128 >> 4, lboB = 1024 >> 4, sboB = 128 >> 4;
        da = (uint64_t)((a_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 46);
        db = (uint64_t)((b_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 46);
        if (layout == 2) { da |= lboA << 16 | sboA << 32; db |= lboB << 16 | sboB << 32; }
        else             { da |= sboA << 16 | lboA << 32; db |= sboB << 16 | lboB << 32; }
      }
      umma_bf16(tmem, da, db, idesc, j > 0 ? 1u : 0u);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  int bad = 0;
  const int r = warp * 32 + (tid % 32);
  for (int c = 0; c < 64; c += 16) {
    uint32_t v[16];
    tmem_ld_x16(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    for (int i = 0; i < 16; ++i) {
      const float got = __uint_as_float(v[i]);
      const int kmax = layout == 4 ? 32 : (layout == 5 ? 16 : 64);
      const float want = (c + i) < kmax ? aval(r + shift, c + i) : 0.f;
      if (got != want) ++bad;
      if (r == 5 && c + i < 8) sample[c + i] = got;
    }
  }
  atomicAdd(mismatches, bad);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

int main() {
  int* d_bad; float* d_s;
  cudaMalloc(&d_bad, 4); cudaMalloc(&d_s, 32);
  cudaFuncSetAttribute(exp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int shifts[] = {0, 1, 2, 3, 7, 8, 9, 18, 31, 130};
  for (int layout : {0, 2, 4, 5})
    for (int s : shifts) {
      cudaMemset(d_bad, 0, 4); cudaMemset(d_s, 0, 32);
      exp_kernel<<<1, 128, 64 * 1024>>>(layout, s, d_bad, d_s);
      cudaError_t e = cudaDeviceSynchronize();
      int bad = -1; float smp[8];
      cudaMemcpy(&bad, d_bad, 4, cudaMemcpyDeviceToHost); cudaMemcpy(smp, d_s, 32, cudaMemcpyDeviceToHost);
      printf("layout %d shift %3d: %s mismatches %5d of 8192   row5 got %g %g %g %g  want %g %g\n", layout, s,
             e == cudaSuccess ? "ok " : cudaGetErrorString(e), bad, smp[0], smp[1], smp[2], smp[3],
             (float)((((5 + s) * 64 + 0) % 251) - 125), (float)((((5 + s) * 64 + 1) % 251) - 125));
      if (e != cudaSuccess) return 1;
    }
  return 0;
}
