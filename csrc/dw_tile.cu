// Depthwise 3x3 convolution, stride 1, shared-memory tiled and TMA-fed (K2, K4, K5).
//
// Same contract as dw_fwd_kernel in dwconv.cu:  y_raw = dwconv3x3( relu6(scale * x_raw + shift) ) + sum(y), sum(y^2).
// What is different is where the latency goes.  dwconv.cu keeps a sliding window AND a register prefetch of raw
// global words per thread (126 registers, 21 % of the warps resident, ~55 instructions per output: address
// arithmetic, per-row masks, unpacking every input three times).  Here
//
//   producer warp   one TMA box {w channels, Wt+2, Ht+2, NI images} of the RAW producer output per tile, halo
//                   zero-filled, NST boxes in flight per CTA and two CTAs per SM -- global latency is covered by
//                   bytes in flight in shared memory, not by registers;
//   8 compute warps pass 1: rewrite the tile in place, relu6(scale*x+shift) -> bf16, and 0 where the halo lies
//                   outside the image (padding is post-activation); each input element is transformed ONCE;
//                   pass 2: a thread owns 4 channels and walks one output row segment with a 3x3 window of fp32
//                   registers (3 LDS.64 + 12 unpacks + 36 FMA per 4 outputs), stores 8 bytes per output pixel
//                   (a half-warp writes one contiguous 128-byte pixel) and keeps the BatchNorm sums in registers.
//
// Channels are split into equal chunks of w <= 64 (144 -> 3 x 48, 960 -> 15 x 64); CTA b owns chunk b % nchunks for
// its whole life (weights, scale/shift and sums stay in registers) and neighbouring CTAs work on the chunks of
// the SAME tile at the same time, so the 128-byte lines of a pixel are fetched from HBM once (measured with
// scripts/exp_tma_bw.cu: streaming one 64-channel chunk of a 144-channel tensor costs as much DRAM time as
// streaming all of it).
//
// The tensor-core formulation of the same conv (dw_tc.cu) is kept as an experiment: ~100-200 cycles per
// tcgen05.mma make nine tap-shifted N=16 MMAs per 128 pixels slower than 36 FMAs per thread.
//
// Reference behaviour replaced: cuDNN/ATen depthwise conv + cuDNN batch-norm + hardtanh as separate passes
// (torchvision mobilenetv2.py ConvBNReLU; cifar10_mpi_mobilenet_224.py:177-179).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "sm100.cuh"

namespace b200dp {
namespace {

using namespace sm100;

constexpr int NST = 3;                        // TMA boxes in flight per CTA
constexpr int STAGE_BYTES = 36 * 1024;
constexpr int CONSUMERS = 256;                // 8 compute warps; thread 0 also issues the TMA loads
constexpr int TILE_THREADS = CONSUMERS;

struct DwTileParams {
  const float* scale;       // [C] BatchNorm scale / shift of the producer
  const float* shift;
  const float* wts;         // [9][C] fp32, tap-major
  __nv_bfloat16* y;         // [B, H, W, C]
  float* stats;             // [2][stats_stride] or null
  int stats_stride;
  int B, H, W, C;
  int w, nchunks;           // channels OWNED by a chunk (multiple of 16, <= 64), C / w
  int wbox;                 // channels LOADED per chunk (>= w; equal to w in practice, see the launcher)
  int Ht, Wt, NI, Hp, Wp;   // output tile, images per tile, haloed box (Wp may carry one padding column)
  int NSEG, SL;             // row segments per output row, segment length
  int nstrips, par;         // strips per tile = NI*Ht*NSEG; strips processed concurrently = CONSUMERS / (w/4)
  int tiles_x, tiles_y, ntiles;
  uint32_t magic_wp, magic_hp, magic_tx, magic_ty, magic_nseg, magic_ht;
};

__device__ __forceinline__ int fastdiv(int r, uint32_t m, int d) { return d == 1 ? r : (int)__umulhi((uint32_t)r, m); }

__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint2 lds8(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
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
__device__ __forceinline__ void red_shared_f2(uint32_t addr, f2 v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(lo) : "memory");
  asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr + 4), "f"(hi) : "memory");
}

__device__ __forceinline__ void decode_tile(const DwTileParams& P, int tile, int& n0, int& y0, int& x0) {
  const int t = fastdiv(tile, P.magic_tx, P.tiles_x);
  x0 = (tile - t * P.tiles_x) * P.Wt;
  const int tn = fastdiv(t, P.magic_ty, P.tiles_y);
  y0 = (t - tn * P.tiles_y) * P.Ht;
  n0 = tn * P.NI;
}

__global__ void __launch_bounds__(TILE_THREADS, 2)
dw_tile_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ DwTileParams P) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + NST * STAGE_BYTES);
  uint64_t* empty = full + NST;
  float* s_stats = reinterpret_cast<float*>(empty + NST);        // [2][64]
  float* s_coef = s_stats + 128;                                  // [2][64] scale, shift of this chunk

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = P.wbox, rowbytes = 2 * w;
  const int chunk = (int)blockIdx.x % P.nchunks;
  const int own0 = chunk * P.w;                                   // first channel this CTA is responsible for
  const int c0 = own0 + w <= P.C ? own0 : P.C - w;                // first channel of its box (never out of range)
  const int lo = own0 - c0, hi = lo + P.w;                        // owned channels, box-relative
  const int tile0 = (int)blockIdx.x / P.nchunks, tstep = (int)gridDim.x / P.nchunks;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], CONSUMERS / 32);
    }
    fence_barrier_init();
  }
  if (threadIdx.x < 128) s_stats[threadIdx.x] = 0.f;
  pdl_wait();
  if (threadIdx.x < w) {
    s_coef[threadIdx.x] = P.scale[c0 + threadIdx.x];
    s_coef[64 + threadIdx.x] = P.shift[c0 + threadIdx.x];
  }
  __syncthreads();

  // ---- TMA: thread 0 keeps NST boxes in flight (a stage is refilled as soon as all eight warps have left it)
  const uint32_t box_bytes = (uint32_t)(P.NI * P.Hp * P.Wp * rowbytes);
  auto issue = [&](int tile, int stage) {
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    mbar_arrive_expect_tx(&full[stage], box_bytes);
    tma_load_4d(base + stage * STAGE_BYTES, &tmX, &full[stage], c0, x0 - 1, y0 - 1, n0);
  };
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmX);
    for (int k = 0; k < NST; ++k)
      if (tile0 + k * tstep < P.ntiles) issue(tile0 + k * tstep, k);
  }

  const int t = threadIdx.x;
  const int Q = w >> 2;                               // channel quads per pixel
  const int j = t % Q, sidx = t / Q;                  // my quad, my strip (one strip per thread per tile)
  const int cq = c0 + 4 * j;
  // ---- my strip: geometry is the same for every tile, only the clipping at the tensor edge is not
  int st_img, st_yy, st_xs;
  {
    const int q1 = fastdiv(sidx, P.magic_nseg, P.NSEG);
    st_xs = (sidx - q1 * P.NSEG) * P.SL;
    st_img = fastdiv(q1, P.magic_ht, P.Ht);
    st_yy = q1 - st_img * P.Ht;
  }
  const bool active = sidx < P.nstrips && 4 * j >= lo && 4 * j < hi;
  const int st_xe = min(st_xs + P.SL, P.Wt);
  // haloed coordinates: output (yy, x) reads rows yy..yy+2, columns x..x+2
  const uint32_t st_off = ((st_img * P.Hp + st_yy) * P.Wp + st_xs) * rowbytes + (j << 3);
  const uint32_t rpitch = P.Wp * rowbytes;
  const size_t ystep = (size_t)P.C * 2;
  f2 wq[9][2], s0[2], s1[2];                          // packed pairs: channels (cq, cq+1) and (cq+2, cq+3)
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float4 wv = active ? *reinterpret_cast<const float4*>(P.wts + k * P.C + cq) : make_float4(0.f, 0.f, 0.f, 0.f);
    wq[k][0] = f2_make(wv.x, wv.y);
    wq[k][1] = f2_make(wv.z, wv.w);
  }
  s0[0] = s0[1] = s1[0] = s1[1] = 0ull;
  // ---- pass-1 role: 16-byte units; my unit column and its scale / shift
  const int U = w >> 3;                               // 16-byte units per pixel
  const int uj = t % U, ur0 = t / U, urstep = CONSUMERS / U;
  const bool xf_active = ur0 < urstep && 8 * uj >= lo && 8 * uj < hi;
  f2 sc2[4], sh2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    sc2[k] = f2_make(s_coef[8 * uj + 2 * k], s_coef[8 * uj + 2 * k + 1]);
    sh2[k] = f2_make(s_coef[64 + 8 * uj + 2 * k], s_coef[64 + 8 * uj + 2 * k + 1]);
  }
  const uint32_t xf_off = ur0 * rowbytes + (uj << 4), xf_step = urstep * rowbytes;
  const uint32_t xf_end = P.NI * P.Hp * P.Wp * rowbytes;

  int it = 0;
#pragma unroll 1
  for (int tile = tile0; tile < P.ntiles; tile += tstep, ++it) {
    const int stage = it % NST, par = (it / NST) & 1;
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    const uint32_t sbase = smem_u32(base + stage * STAGE_BYTES);
    mbar_wait(&full[stage], par);
    // ---- pass 1: BatchNorm + ReLU6 of the producer, in place.  The halo outside the image was filled with NaN by
    //      the copy engine; max(NaN, 0) = 0 (non-propagating max), which is exactly the post-activation zero padding,
    //      so no thread ever needs to know where in the image its pixel lies.
    if (xf_active) {
#pragma unroll 2
      for (uint32_t o = xf_off; o < xf_end; o += xf_step) {
        const uint4 raw = lds16(sbase + o);
        uint4 v;
        v.x = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw.x), sc2[0], sh2[0])));
        v.y = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw.y), sc2[1], sh2[1])));
        v.z = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw.z), sc2[2], sh2[2])));
        v.w = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw.w), sc2[3], sh2[3])));
        sts16(sbase + o, v);
      }
    }
    named_bar_sync(1, CONSUMERS);
    // ---- pass 2: 3x3 stencil over my strip
    const int oy = y0 + st_yy, on = n0 + st_img;
    const int xe = min(st_xe, P.W - x0);
    if (active && oy < P.H && on < P.B && st_xs < xe) {
      uint32_t a0 = sbase + st_off;
      uint32_t a1 = a0 + rpitch, a2 = a1 + rpitch;
      char* yp = reinterpret_cast<char*>(P.y + (((size_t)on * P.H + oy) * P.W + x0 + st_xs) * P.C + cq);
      f2 win[3][3][2];                              // [column slot][row][channel pair]
      auto load_col = [&](int slot) {
        const uint2 v0 = lds8(a0), v1 = lds8(a1), v2 = lds8(a2);
        a0 += rowbytes; a1 += rowbytes; a2 += rowbytes;
        win[slot][0][0] = f2_from_bf16x2(v0.x); win[slot][0][1] = f2_from_bf16x2(v0.y);
        win[slot][1][0] = f2_from_bf16x2(v1.x); win[slot][1][1] = f2_from_bf16x2(v1.y);
        win[slot][2][0] = f2_from_bf16x2(v2.x); win[slot][2][1] = f2_from_bf16x2(v2.y);
      };
      auto emit = [&](int sl, int sm, int sr) {     // column slots holding x-1, x, x+1
        uint2 o;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          f2 a = 0ull;
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            a = f2_fma(wq[r * 3 + 0][h], win[sl][r][h], a);
            a = f2_fma(wq[r * 3 + 1][h], win[sm][r][h], a);
            a = f2_fma(wq[r * 3 + 2][h], win[sr][r][h], a);
          }
          s0[h] = f2_add(s0[h], a);
          s1[h] = f2_fma(a, a, s1[h]);
          (h == 0 ? o.x : o.y) = f2_to_bf16x2(a);
        }
        *reinterpret_cast<uint2*>(yp) = o;
        yp += ystep;
      };
      load_col(0);
      load_col(1);
      int x = st_xs;
      // three outputs per trip so that the window rotates by renaming, not by moving registers
#pragma unroll 1
      for (; x + 3 <= xe; x += 3) {
        load_col(2); emit(0, 1, 2);
        load_col(0); emit(1, 2, 0);
        load_col(1); emit(2, 0, 1);
      }
      if (x < xe) { load_col(2); emit(0, 1, 2); ++x; }
      if (x < xe) { load_col(0); emit(1, 2, 0); ++x; }
    }
    fence_proxy_async_smem();          // my generic-proxy writes to the stage vs. the next TMA write into it
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[stage]);
    if (threadIdx.x == 0 && tile + NST * tstep < P.ntiles) {
      mbar_wait(&empty[stage], par);
      issue(tile + NST * tstep, stage);
    }
  }
  if (P.stats != nullptr && active) {
    const uint32_t sa = smem_u32(s_stats) + 16 * j;
    red_shared_f2(sa, s0[0]); red_shared_f2(sa + 8, s0[1]);
    red_shared_f2(sa + 256, s1[0]); red_shared_f2(sa + 264, s1[1]);
  }
  named_bar_sync(1, CONSUMERS);
  if (P.stats != nullptr && t < 2 * w) {
    const int h = t / w, c = t - h * w;
    if (c >= lo && c < hi) atomicAdd(P.stats + h * P.stats_stride + c0 + c, s_stats[h * 64 + c]);
  }
}

uint32_t magic(int d) { return (uint32_t)(((1ull << 32) + (uint64_t)d - 1) / (uint64_t)d); }

int encode_nhwc(CUtensorMap* out, const void* base, int B, int H, int W, int C, int box_c, int box_w, int box_h, int box_n) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    B200DP_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || fn == nullptr) {
      set_error("cuTensorMapEncodeTiled not available from the driver");
      return B200DP_ECUDA;
    }
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstride[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)box_n};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA);   // halo outside the tensor = NaN (see pass 1)
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(4d) failed (%d): dims %dx%dx%dx%d box %dx%dx%dx%d", (int)r, C, W, H, B, box_c, box_w,
              box_h, box_n);
    return B200DP_ECUDA;
  }
  return B200DP_OK;
}

// equal channel chunks of at most 64 (multiples of 16); 0 if C cannot be split that way
int chunk_width(int C) {
  if (C % 16 != 0) return 0;
  if (C <= 64) return C;
  for (int n = (C + 63) / 64; n <= C / 16; ++n)
    if (C % n == 0 && (C / n) % 16 == 0) return C / n;
  return 0;
}

// tile shape: maximise (thread utilisation) x (useful fraction of the haloed box) x (segment prologue) x (edge waste)
void pick_tile(DwTileParams& P) {
  const int Q = P.wbox / 4;
  P.par = CONSUMERS / Q;
  const int max_px = STAGE_BYTES / (2 * P.wbox);
  double best = -1.0;
  const int wcand[4] = {P.W < 28 ? P.W : 28, 21, 14, 7};
  for (int wi = 0; wi < 4; ++wi) {
    const int Wt = wcand[wi];
    if (Wt > P.W || (wi > 0 && Wt >= wcand[0])) continue;
    int Wp = Wt + 2;
    if (P.wbox < 64 && (Wp % 2) == 0) ++Wp;     // odd pitch: rows of < 128 bytes then land in different banks
    for (int Ht = 1; Ht <= P.H && Ht <= 32; ++Ht)
      for (int NI = 1; NI <= 4 && NI <= P.B; ++NI) {
        const int Hp = Ht + 2;
        if (NI * Hp * Wp > max_px) continue;
        for (int NSEG = 1; NSEG <= 4; ++NSEG) {
          const int SL = (Wt + NSEG - 1) / NSEG;
          if (SL < 5 && NSEG > 1) continue;
          const int strips = NI * Ht * NSEG;
          if (strips > P.par) continue;             // one strip per thread: its geometry is hoisted out of the tile loop
          const double util = (double)strips / P.par * ((double)Wt / (NSEG * SL));
          const double halo = (double)(Ht * Wt) / (Hp * Wp);
          const double prol = (double)SL / (SL + 2);
          const int tx = (P.W + Wt - 1) / Wt, ty = (P.H + Ht - 1) / Ht, tn = (P.B + NI - 1) / NI;
          const double edge = ((double)P.W / (tx * Wt)) * ((double)P.H / (ty * Ht)) * ((double)P.B / (tn * NI));
          const double size = NI * Hp * Wp * 2.0 * P.wbox >= 16384 ? 1.0 : 0.9;   // small boxes: more barrier trips
          const double score = util * (0.5 + 0.5 * halo) * (0.7 + 0.3 * prol) * edge * size;
          if (score > best) {
            best = score;
            P.Wt = Wt; P.Ht = Ht; P.NI = NI; P.Hp = Hp; P.Wp = Wp; P.NSEG = NSEG; P.SL = SL;
          }
        }
      }
  }
  P.nstrips = P.NI * P.Ht * P.NSEG;
  P.tiles_x = (P.W + P.Wt - 1) / P.Wt;
  P.tiles_y = (P.H + P.Ht - 1) / P.Ht;
  P.ntiles = P.tiles_x * P.tiles_y * ((P.B + P.NI - 1) / P.NI);
}

}  // namespace

bool dw_tile_supported(int B, int H, int W, int C, int stride) {
  return stride == 1 && chunk_width(C) > 0 && B >= 1 && H >= 1 && W >= 1 && H < 4096 && W < 4096;
}

int dw_tile_fwd_launch(const void* x, const float* scale, const float* shift, const float* wts, void* y, float* stats,
                       int stats_stride, int B, int H, int W, int C, cudaStream_t s) {
  B200DP_REQUIRE(dw_tile_supported(B, H, W, C, 1), "dw_tile_fwd: B=%d H=%d W=%d C=%d unsupported", B, H, W, C);
  DwTileParams P{};
  P.scale = scale; P.shift = shift; P.wts = wts; P.y = static_cast<__nv_bfloat16*>(y);
  P.stats = stats; P.stats_stride = stats_stride;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.w = chunk_width(C);
  P.nchunks = C / P.w;
  P.wbox = P.w;   // measured: loading 64-wide (conflict-free 128-byte) rows for a 48-wide chunk idles a quarter of the lanes
                  // and is slower than living with the bank conflicts of 96-byte rows (108 vs 90 us on 56x56x144)
  pick_tile(P);
  P.magic_wp = magic(P.Wp); P.magic_hp = magic(P.Hp); P.magic_tx = magic(P.tiles_x); P.magic_ty = magic(P.tiles_y);
  P.magic_nseg = magic(P.NSEG); P.magic_ht = magic(P.Ht);
  CUtensorMap tm;
  if (int rc = encode_nhwc(&tm, x, B, H, W, C, P.wbox, P.Wp, P.Hp, P.NI)) return rc;
  // CTA b owns chunk b % nchunks: the grid is a multiple of nchunks, at most two CTAs per SM
  int groups = 2 * num_sms() / P.nchunks;
  if (groups < 1) groups = 1;
  if (groups > P.ntiles) groups = P.ntiles;
  const int grid = groups * P.nchunks;
  const size_t smem = 128 + NST * STAGE_BYTES + 2 * NST * 8 + 256 * 4;
  B200DP_CUDA(cudaFuncSetAttribute(dw_tile_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024));
  B200DP_CUDA(launch_k(dw_tile_fwd_kernel, dim3((unsigned)grid), dim3(TILE_THREADS), smem, s, tm, P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_dw_fwd_tile(const void* x, const float* scale, const float* shift, const float* w, void* y,
                                  float* stats, int stats_stride, int B, int H, int W, int C, int stride, void* stream) {
  B200DP_REQUIRE(stride == 1, "dw_fwd_tile: stride 1 only");
  return dw_tile_fwd_launch(x, scale, shift, w, y, stats, stats_stride, B, H, W, C, static_cast<cudaStream_t>(stream));
}
