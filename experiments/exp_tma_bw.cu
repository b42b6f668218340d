// Experiment: TMA tiled-load throughput per SM as a function of box shape, tensor rank and bytes in flight.
// Persistent CTAs (one per SM) stream a [B,H,W,C] bf16 tensor through shared memory and do nothing else:
// a stage is re-armed as soon as its box has landed.
//   mode 0: rank-2 view [B*H*W][C], boxes {w, rows}
//   mode 1: rank-4 view, halo boxes {w, Wt+2, Ht+2, 1} at (x0-1, y0-1)
// usage: exp_tma_bw mode C w stages rows|Ht [Wt] [H]
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -Icsrc experiments/exp_tma_bw.cu -o build/exp_tma_bw -lcuda
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_bf16.h>

#include "sm100.cuh"

using namespace b200dp::sm100;

struct Args {
  int mode, w, stages, box_bytes, ntiles, tiles_x, tiles_y, Ht, Wt, rows;
};

__global__ void __launch_bounds__(128, 1) bw_kernel(const __grid_constant__ CUtensorMap tm, const Args a,
                                                    unsigned long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[16];
  const int stage_bytes = (a.box_bytes + 1023) & ~1023;
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) mbar_init(&full[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    int it = 0;
    auto issue = [&](int tile, int stage) {
      mbar_arrive_expect_tx(&full[stage], (uint32_t)a.box_bytes);
      if (a.mode == 0) {
        tma_load_2d(smem + stage * stage_bytes, &tm, &full[stage], 0, tile * a.rows);
      } else {
        const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
        tma_load_4d(smem + stage * stage_bytes, &tm, &full[stage], 0, tx * a.Wt - 1, ty * a.Ht - 1, n);
      }
    };
    // prime
    int tile = blockIdx.x, issued = 0;
    for (; issued < a.stages && tile < a.ntiles; ++issued, tile += gridDim.x) issue(tile, issued);
    int done = 0;
    while (done < issued) {
      const int stage = done % a.stages, par = (done / a.stages) & 1;
      mbar_wait(&full[stage], par);
      ++done;
      if (tile < a.ntiles) { issue(tile, stage); ++issued; tile += gridDim.x; }
    }
    (void)it;
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

int main(int argc, char** argv) {
  if (argc < 6) { printf("usage: mode C w stages rows|Ht [Wt] [H]\n"); return 1; }
  const int mode = atoi(argv[1]), C = atoi(argv[2]), w = atoi(argv[3]), stages = atoi(argv[4]);
  const int r = atoi(argv[5]);
  const int Wt = argc > 6 ? atoi(argv[6]) : 28;
  const int H = argc > 7 ? atoi(argv[7]) : 56;
  const int B = 128, W = H;
  const size_t elems = (size_t)B * H * W * C;
  __nv_bfloat16* x;
  cudaMalloc(&x, elems * 2);
  cudaMemset(x, 0, elems * 2);
  CUtensorMap tm;
  Args a{};
  a.mode = mode; a.w = w; a.stages = stages;
  const CUtensorMapSwizzle sw = w == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (w == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUresult rc;
  if (mode == 0) {
    a.rows = r;
    cuuint64_t gdim[2] = {(cuuint64_t)C, (cuuint64_t)B * H * W};
    cuuint64_t gstr[1] = {(cuuint64_t)C * 2};
    cuuint32_t box[2] = {(cuuint32_t)w, (cuuint32_t)r};
    cuuint32_t es[2] = {1, 1};
    rc = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, x, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    a.box_bytes = r * w * 2;
    a.ntiles = (int)((size_t)B * H * W / r);
  } else {
    a.Ht = r; a.Wt = Wt;
    cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {(cuuint32_t)w, (cuuint32_t)(Wt + 2), (cuuint32_t)(r + 2), 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    rc = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    a.box_bytes = (Wt + 2) * (r + 2) * w * 2;
    a.tiles_x = (W + Wt - 1) / Wt; a.tiles_y = (H + r - 1) / r;
    a.ntiles = a.tiles_x * a.tiles_y * B;
  }
  if (rc != CUDA_SUCCESS) { printf("encode failed %d\n", (int)rc); return 1; }
  const int stage_bytes = (a.box_bytes + 1023) & ~1023;
  const size_t smem = (size_t)stages * stage_bytes + 1024;
  if (smem > 227 * 1024) { printf("smem %zu too large\n", smem); return 1; }
  cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  unsigned long long* cyc;
  cudaMalloc(&cyc, 148 * 8);
  void* flush; cudaMalloc(&flush, 256 << 20);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaMemset(flush, rep, 256 << 20);
    cudaEventRecord(e0);
    bw_kernel<<<148, 128, smem>>>(tm, a, cyc);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const double loaded = (double)a.ntiles * a.box_bytes;
  printf("mode %d C %3d w %2d stages %d box %6d B (%s)  tiles %6d  %.1f us  loaded %.1f MB -> %.2f TB/s  = %.1f B/clk/SM @1.965GHz  (tensor %.1f MB)\n",
         mode, C, w, stages, a.box_bytes, mode == 0 ? "2d" : "4d halo", a.ntiles, best * 1e3, loaded / 1e6,
         loaded / (best * 1e-3) / 1e12, loaded / (best * 1e-3) / 148 / 1.965e9, elems * 2 / 1e6);
  return 0;
}
