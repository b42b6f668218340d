// Depthwise 3x3 convolution, shared-memory tiled and TMA-fed (K2, K4, K5): forward for stride 1 and 2,
// backward for stride 1.
//
// Same contract as the kernels in dwconv.cu:  y_raw = dwconv3x3( relu6(scale * x_raw + shift) ) + sum(y), sum(y^2).
// What is different is where the latency goes.  dwconv.cu keeps a sliding window AND a register prefetch of raw
// global words per thread (126 registers, 21 % of the warps resident, ~55 instructions per output: address
// arithmetic, per-row masks, unpacking every input three times).  Here
//
//   thread 0        keeps NST TMA boxes {w channels, Wt+2, Ht+2, NI images} of the RAW producer output in flight per
//                   CTA (two CTAs per SM): global latency is covered by bytes in flight in shared memory, not by
//                   registers.  The halo outside the image is filled with NaN by the copy engine.
//   8 compute warps pass 1: rewrite the tile in place, min(max(bf16(scale*x+shift), 0), 6) on packed bf16 pairs; max /
//                   min return the non-NaN operand, so the NaN halo becomes 0 = the post-activation zero padding and
//                   no thread ever computes where in the image its pixel lies; each input is transformed ONCE;
//                   pass 2: a thread owns 4 channels and one output row segment (its geometry is a per-thread constant
//                   of the whole kernel), a 3x3 window of packed fp32 pairs: 3 LDS.64 + 12 unpacks + 18 FFMA2 per
//                   4-channel output, 8-byte stores (a half-warp writes one contiguous pixel), BatchNorm sums in
//                   registers for the whole kernel.
//
// Channels are split into equal chunks of w <= 64 (144 -> 3 x 48, 960 -> 15 x 64); CTA b owns chunk b % nchunks for
// its whole life (weights and sums stay in registers) and neighbouring CTAs work on the chunks of the SAME tile at the
// same time, so the 128-byte lines of a pixel are fetched from HBM once (measured with experiments/exp_tma_bw.cu:
// streaming one 64-channel chunk of a 144-channel tensor costs as much DRAM time as streaming all of it).
// Stride 2: four parity planes of the input per tile (TMA element strides {1, 2, 2, 1}); every tap of every output
// is then a plain (row, column) offset inside one plane; the chunk width is chosen per layer.
// The BatchNorm of the producer can be finalised in the prologue (bn_inline.cuh): this kernel is its first consumer.
//
// The tensor-core formulation of the same conv (experiments/dw_tc.cu, not part of the library) was measured: ~100-200 cycles per
// tcgen05.mma make nine tap-shifted N=16 MMAs per 128 pixels slower than 18 packed FMAs per thread.
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
  int S, Ho, Wo;            // stride (1 or 2; backward: 1 only) and output size; tiles are tiles of the OUTPUT
  int nplanes, plane_stride;  // stride 2: four parity planes of the input (row parity, column parity), bytes apart
  int w, nchunks;           // channels OWNED by a chunk (multiple of 16, <= 64), C / w
  int wbox;                 // channels LOADED per chunk (>= w; equal to w in practice, see the launcher)
  int Ht, Wt, NI, Hp, Wp;   // output tile, images per tile, haloed box (Wp may carry one padding column)
  int NSEG, SL;             // row segments per output row, segment length
  int nstrips, par;         // strips per tile = NI*Ht*NSEG; strips processed concurrently = CONSUMERS / (w/4)
  int tiles_x, tiles_y, ntiles;
  uint32_t magic_wp, magic_hp, magic_tx, magic_ty, magic_nseg, magic_ht;
  // backward only
  const float* dcoef;       // [3][C] ca, cb, cc of the BatchNorm backward of THIS conv's output
  float* dwg;               // [9][C] weight gradient (accumulated)
  int Wq;                   // pixel pitch of the un-haloed e_raw box
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

template <int S>
__global__ void __launch_bounds__(TILE_THREADS, 2)
dw_tile_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ DwTileParams P,
                   const __grid_constant__ BnFwdSrc bsrc) {
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
    float sc, sh;
    if (bsrc.sums != nullptr) {     // the producer's BatchNorm is finalised here; the first CTA of each chunk publishes it
      bn_fwd_channel(bsrc, c0 + threadIdx.x, tile0 == 0, sc, sh);
    } else {
      sc = P.scale[c0 + threadIdx.x];
      sh = P.shift[c0 + threadIdx.x];
    }
    s_coef[threadIdx.x] = sc;
    s_coef[64 + threadIdx.x] = sh;
  }
  __syncthreads();

  // ---- TMA: thread 0 keeps NST boxes in flight (a stage is refilled as soon as all eight warps have left it)
  const uint32_t box_bytes = (uint32_t)(P.NI * P.Hp * P.Wp * rowbytes);
  auto issue = [&](int tile, int stage) {
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    mbar_arrive_expect_tx(&full[stage], box_bytes * (S == 1 ? 1 : 4));
    if (S == 1) {
      tma_load_4d(base + stage * STAGE_BYTES, &tmX, &full[stage], c0, x0 - 1, y0 - 1, n0);
    } else {
      // plane (pr, pc), entry (i, j) = input pixel (2 (y0-1+i) + pr, 2 (x0-1+j) + pc): the map samples every other
      // row and column, so every tap of every output is a plain (row, column) offset inside ONE plane
#pragma unroll
      for (int pl = 0; pl < 4; ++pl)
        tma_load_4d(base + stage * STAGE_BYTES + pl * P.plane_stride, &tmX, &full[stage], c0, 2 * (x0 - 1) + (pl & 1),
                    2 * (y0 - 1) + (pl >> 1), n0);
    }
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
  const uint32_t coef_u32 = smem_u32(s_coef) + 32 * uj;     // my unit's scale / shift: re-read per tile (4 loads) so that
                                                            // they do not occupy 16 registers during the stencil
  const uint32_t xf_off = ur0 * rowbytes + (uj << 4), xf_step = urstep * rowbytes;
  // stride 2: the four planes are transformed as one range (the few alignment bytes between them hold nothing)
  const uint32_t xf_end = S == 1 ? P.NI * P.Hp * P.Wp * rowbytes : 3 * P.plane_stride + P.NI * P.Hp * P.Wp * rowbytes;

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
      f2 sc2[4], sh2[4];
      {
        const uint4 a0 = lds16(coef_u32), a1 = lds16(coef_u32 + 16), b0 = lds16(coef_u32 + 256), b1 = lds16(coef_u32 + 272);
        sc2[0] = f2_pair(a0.x, a0.y); sc2[1] = f2_pair(a0.z, a0.w); sc2[2] = f2_pair(a1.x, a1.y); sc2[3] = f2_pair(a1.z, a1.w);
        sh2[0] = f2_pair(b0.x, b0.y); sh2[1] = f2_pair(b0.z, b0.w); sh2[2] = f2_pair(b1.x, b1.y); sh2[3] = f2_pair(b1.z, b1.w);
      }
      // four independent 16-byte loads in flight, then the math, then the stores (the asm accesses are volatile:
      // load -> math -> store per unit would serialise on the shared-memory round trip)
      for (uint32_t o = xf_off; o < xf_end; o += 4 * xf_step) {
        uint4 raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (o + u * xf_step < xf_end) raw[u] = lds16(sbase + o + u * xf_step);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          raw[u].x = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw[u].x), sc2[0], sh2[0])));
          raw[u].y = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw[u].y), sc2[1], sh2[1])));
          raw[u].z = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw[u].z), sc2[2], sh2[2])));
          raw[u].w = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(raw[u].w), sc2[3], sh2[3])));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (o + u * xf_step < xf_end) sts16(sbase + o + u * xf_step, raw[u]);
      }
    }
    named_bar_sync(1, CONSUMERS);
    // ---- pass 2: 3x3 stencil over my strip
    const int oy = y0 + st_yy, on = n0 + st_img;
    const int xe = min(st_xe, P.Wo - x0);
    if (active && oy < P.Ho && on < P.B && st_xs < xe) {
      char* yp = reinterpret_cast<char*>(P.y + (((size_t)on * P.Ho + oy) * P.Wo + x0 + st_xs) * P.C + cq);
      f2 win[3][3][2];                              // [column slot][row][channel pair]
      auto emit = [&](int sl, int sm, int sr) {     // column slots holding taps kx = 0, 1, 2
        uint2 o;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          f2 a = 0ull;
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            a = F2_STENCIL_FMA(wq[r * 3 + 0][h], win[sl][r][h], a);
            a = F2_STENCIL_FMA(wq[r * 3 + 1][h], win[sm][r][h], a);
            a = F2_STENCIL_FMA(wq[r * 3 + 2][h], win[sr][r][h], a);
          }
          s0[h] = f2_add(s0[h], a);
          s1[h] = f2_fma(a, a, s1[h]);
          (h == 0 ? o.x : o.y) = f2_to_bf16x2(a);
        }
        *reinterpret_cast<uint2*>(yp) = o;
        yp += ystep;
      };
      if (S == 1) {
        uint32_t a0 = sbase + st_off;
        uint32_t a1 = a0 + rpitch, a2 = a1 + rpitch;
        auto load_col = [&](int slot) {
          const uint2 v0 = lds8(a0), v1 = lds8(a1), v2 = lds8(a2);
          a0 += rowbytes; a1 += rowbytes; a2 += rowbytes;
          win[slot][0][0] = f2_from_bf16x2(v0.x); win[slot][0][1] = f2_from_bf16x2(v0.y);
          win[slot][1][0] = f2_from_bf16x2(v1.x); win[slot][1][1] = f2_from_bf16x2(v1.y);
          win[slot][2][0] = f2_from_bf16x2(v2.x); win[slot][2][1] = f2_from_bf16x2(v2.y);
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
      } else {
        // stride 2.  Output (yy, x) reads, relative to the plane origin:
        //   ky = 0: odd-row planes, row yy      ky = 1: even-row planes, row yy+1      ky = 2: odd-row planes, row yy+1
        //   kx = 0: odd-col planes, col x       kx = 1: even-col planes, col x+1       kx = 2: odd-col planes, col x+1
        // so the kx = 2 column of one output is the kx = 0 column of the next: two new columns per output.
        const uint32_t pl_e = sbase + st_off;                          // (row, col) = (yy, xs) of plane 0 + my quad
        const uint32_t ps = P.plane_stride;
        // row r of the window: plane row-parity / row index as listed above
        uint32_t ae0 = pl_e + 2 * ps + rowbytes;                       // ky=0: plane (1,0), row yy,   col xs+1
        uint32_t ae1 = pl_e + rpitch + rowbytes;                       // ky=1: plane (0,0), row yy+1, col xs+1
        uint32_t ae2 = pl_e + 2 * ps + rpitch + rowbytes;              // ky=2: plane (1,0), row yy+1, col xs+1
        uint32_t ao0 = pl_e + 3 * ps;                                  // ky=0: plane (1,1), row yy,   col xs
        uint32_t ao1 = pl_e + ps + rpitch;                             // ky=1: plane (0,1), row yy+1, col xs
        uint32_t ao2 = pl_e + 3 * ps + rpitch;                         // ky=2: plane (1,1), row yy+1, col xs
        auto load_odd = [&](int slot) {
          const uint2 v0 = lds8(ao0), v1 = lds8(ao1), v2 = lds8(ao2);
          ao0 += rowbytes; ao1 += rowbytes; ao2 += rowbytes;
          win[slot][0][0] = f2_from_bf16x2(v0.x); win[slot][0][1] = f2_from_bf16x2(v0.y);
          win[slot][1][0] = f2_from_bf16x2(v1.x); win[slot][1][1] = f2_from_bf16x2(v1.y);
          win[slot][2][0] = f2_from_bf16x2(v2.x); win[slot][2][1] = f2_from_bf16x2(v2.y);
        };
        auto load_even = [&](int slot) {
          const uint2 v0 = lds8(ae0), v1 = lds8(ae1), v2 = lds8(ae2);
          ae0 += rowbytes; ae1 += rowbytes; ae2 += rowbytes;
          win[slot][0][0] = f2_from_bf16x2(v0.x); win[slot][0][1] = f2_from_bf16x2(v0.y);
          win[slot][1][0] = f2_from_bf16x2(v1.x); win[slot][1][1] = f2_from_bf16x2(v1.y);
          win[slot][2][0] = f2_from_bf16x2(v2.x); win[slot][2][1] = f2_from_bf16x2(v2.y);
        };
        load_odd(0);
        int x = st_xs;
#pragma unroll 1
        for (; x + 2 <= xe; x += 2) {
          load_even(1); load_odd(2); emit(0, 1, 2);
          load_even(1); load_odd(0); emit(2, 1, 0);
        }
        if (x < xe) { load_even(1); load_odd(2); emit(0, 1, 2); }
      }
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

// ---------------------------------------------------------------------------------------------------------------
// Backward, stride 1 (same contract as dw_bwd_kernel<1> in dwconv.cu).  Per tile the copy engine lands three boxes:
// the haloed dz and d_raw tiles of the conv OUTPUT side and the un-haloed e_raw tile of the input side.
//   pass 1 (16 warps)  dd = ca*dz + cb*d_raw + cc -> bf16, in place of dz; the NaN halo becomes 0 (max(v,0) | min(v,0))
//   pass 2, warps 0-7  data gradient: 3x3 window of dd against the flipped weights, ReLU6 mask from e_raw, store,
//                      sums (dz_e, dz_e * e_raw) for the producer's BatchNorm backward
//           warps 8-15 weight gradient: the same window times a = relu6(scale*e_raw + shift), 36 fp32 accumulators
//                      per thread that live for the whole kernel
// The two halves read the same shared-memory tile concurrently; neither has to fit both register sets.
constexpr int BW_THREADS = 512, BW_GROUP = 256, BW_NST = 2;
constexpr int HALO_BYTES = 36 * 1024, PLAIN_BYTES = 28 * 1024, BW_STAGE = 2 * HALO_BYTES + PLAIN_BYTES;

__global__ void __launch_bounds__(BW_THREADS, 1)
dw_tile_bwd_kernel(const __grid_constant__ CUtensorMap tmZ, const __grid_constant__ CUtensorMap tmG,
                   const __grid_constant__ CUtensorMap tmE, const __grid_constant__ DwTileParams P,
                   const __grid_constant__ BnBwdSrc bsrc) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + BW_NST * BW_STAGE);
  uint64_t* empty = full + BW_NST;
  float* s_stats = reinterpret_cast<float*>(empty + BW_NST);     // [2][64]
  float* s_ecoef = s_stats + 128;                                 // [2][64] scale, shift of the input-side BatchNorm
  float* s_dcoef = s_ecoef + 128;                                 // [3][64] ca, cb, cc
  float* s_dw = s_dcoef + 192;                                    // [9][64]

  const int lane = threadIdx.x & 31;
  const int w = P.w, rowbytes = 2 * w;
  const int chunk = (int)blockIdx.x % P.nchunks, c0 = chunk * w;
  const int tile0 = (int)blockIdx.x / P.nchunks, tstep = (int)gridDim.x / P.nchunks;

  if (threadIdx.x == 0) {
    for (int i = 0; i < BW_NST; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], BW_THREADS / 32);
    }
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 128 + 9 * 64; i += BW_THREADS) (i < 128 ? s_stats[i] : s_dw[i - 128]) = 0.f;
  pdl_wait();
  if (threadIdx.x < w) {
    s_ecoef[threadIdx.x] = P.scale[c0 + threadIdx.x];
    s_ecoef[64 + threadIdx.x] = P.shift[c0 + threadIdx.x];
    float ca, cb, cc;
    if (bsrc.sums != nullptr) {     // BatchNorm backward of this conv's output, finalised here
      bn_bwd_channel(bsrc, c0 + threadIdx.x, tile0 == 0, ca, cb, cc);
    } else {
      ca = P.dcoef[c0 + threadIdx.x];
      cb = P.dcoef[P.C + c0 + threadIdx.x];
      cc = P.dcoef[2 * P.C + c0 + threadIdx.x];
    }
    s_dcoef[threadIdx.x] = ca;
    s_dcoef[64 + threadIdx.x] = cb;
    s_dcoef[128 + threadIdx.x] = cc;
  }
  __syncthreads();

  const uint32_t halo_bytes = (uint32_t)(P.NI * P.Hp * P.Wp * rowbytes);
  const uint32_t plain_bytes = (uint32_t)(P.NI * P.Ht * P.Wq * rowbytes);
  auto issue = [&](int tile, int stage) {
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    uint8_t* st = base + stage * BW_STAGE;
    mbar_arrive_expect_tx(&full[stage], 2 * halo_bytes + plain_bytes);
    tma_load_4d(st, &tmZ, &full[stage], c0, x0 - 1, y0 - 1, n0);
    tma_load_4d(st + HALO_BYTES, &tmG, &full[stage], c0, x0 - 1, y0 - 1, n0);
    tma_load_4d(st + 2 * HALO_BYTES, &tmE, &full[stage], c0, x0, y0, n0);
  };
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmZ); prefetch_tmap(&tmG); prefetch_tmap(&tmE);
    for (int k = 0; k < BW_NST; ++k)
      if (tile0 + k * tstep < P.ntiles) issue(tile0 + k * tstep, k);
  }

  const bool grp_b = threadIdx.x >= BW_GROUP;         // warps 8-15: weight gradient
  const int t = threadIdx.x - (grp_b ? BW_GROUP : 0);
  const int Q = w >> 2;
  const int j = t % Q, sidx = t / Q;
  const int cq = c0 + 4 * j;
  int st_img, st_yy, st_xs;
  {
    const int q1 = fastdiv(sidx, P.magic_nseg, P.NSEG);
    st_xs = (sidx - q1 * P.NSEG) * P.SL;
    st_img = fastdiv(q1, P.magic_ht, P.Ht);
    st_yy = q1 - st_img * P.Ht;
  }
  const bool active = sidx < P.nstrips;
  const int st_xe = min(st_xs + P.SL, P.Wt);
  const uint32_t st_off = ((st_img * P.Hp + st_yy) * P.Wp + st_xs) * rowbytes + (j << 3);
  const uint32_t e_off = 2 * HALO_BYTES + ((st_img * P.Ht + st_yy) * P.Wq + st_xs) * rowbytes + (j << 3);
  const uint32_t rpitch = P.Wp * rowbytes;
  const size_t ystep = (size_t)P.C * 2;
  // group A: flipped weights; group B: weight-gradient accumulators.  Same registers, different meaning.
  f2 wa[9][2];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!grp_b && active) wv = *reinterpret_cast<const float4*>(P.wts + (8 - k) * P.C + cq);
    wa[k][0] = f2_make(wv.x, wv.y);
    wa[k][1] = f2_make(wv.z, wv.w);
  }
  f2 s0[2], s1[2];
  s0[0] = s0[1] = s1[0] = s1[1] = 0ull;
  const f2 esc0 = f2_make(s_ecoef[4 * j], s_ecoef[4 * j + 1]), esc1 = f2_make(s_ecoef[4 * j + 2], s_ecoef[4 * j + 3]);
  const f2 esh0 = f2_make(s_ecoef[64 + 4 * j], s_ecoef[64 + 4 * j + 1]);
  const f2 esh1 = f2_make(s_ecoef[64 + 4 * j + 2], s_ecoef[64 + 4 * j + 3]);
  // pass-1 role (all 512 threads): 16-byte units of the haloed tiles
  const int U = w >> 3;
  const int uj = threadIdx.x % U, ur0 = threadIdx.x / U, urstep = BW_THREADS / U;
  const bool xf_active = ur0 < urstep;
  const uint32_t xf_off = ur0 * rowbytes + (uj << 4), xf_step = urstep * rowbytes;
  const uint32_t dc_addr = smem_u32(s_dcoef) + 32 * uj;

  int it = 0;
#pragma unroll 1
  for (int tile = tile0; tile < P.ntiles; tile += tstep, ++it) {
    const int stage = it % BW_NST, par = (it / BW_NST) & 1;
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    const uint32_t sbase = smem_u32(base + stage * BW_STAGE);
    mbar_wait(&full[stage], par);
    // ---- pass 1: dd = ca*dz + cb*d_raw + cc, in place of dz; NaN (halo outside the image) -> 0
    if (xf_active) {
      const uint4 ka0 = lds16(dc_addr), ka1 = lds16(dc_addr + 16);
      const uint4 kb0 = lds16(dc_addr + 256), kb1 = lds16(dc_addr + 272);
      const uint4 kc0 = lds16(dc_addr + 512), kc1 = lds16(dc_addr + 528);
      const f2 ca[4] = {f2_pair(ka0.x, ka0.y), f2_pair(ka0.z, ka0.w), f2_pair(ka1.x, ka1.y), f2_pair(ka1.z, ka1.w)};
      const f2 cb[4] = {f2_pair(kb0.x, kb0.y), f2_pair(kb0.z, kb0.w), f2_pair(kb1.x, kb1.y), f2_pair(kb1.z, kb1.w)};
      const f2 cc[4] = {f2_pair(kc0.x, kc0.y), f2_pair(kc0.z, kc0.w), f2_pair(kc1.x, kc1.y), f2_pair(kc1.z, kc1.w)};
#pragma unroll 2
      for (uint32_t o = xf_off; o < halo_bytes; o += xf_step) {       // (batching two units here spills registers)
        const uint4 z = lds16(sbase + o), g = lds16(sbase + HALO_BYTES + o);
        uint4 v;
        v.x = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[0], f2_from_bf16x2(z.x), f2_fma(cb[0], f2_from_bf16x2(g.x), cc[0]))));
        v.y = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[1], f2_from_bf16x2(z.y), f2_fma(cb[1], f2_from_bf16x2(g.y), cc[1]))));
        v.z = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[2], f2_from_bf16x2(z.z), f2_fma(cb[2], f2_from_bf16x2(g.z), cc[2]))));
        v.w = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[3], f2_from_bf16x2(z.w), f2_fma(cb[3], f2_from_bf16x2(g.w), cc[3]))));
        sts16(sbase + o, v);
      }
    }
    named_bar_sync(1, BW_THREADS);
    // ---- pass 2
    const int oy = y0 + st_yy, on = n0 + st_img;
    const int xe = min(st_xe, P.W - x0);
    if (active && oy < P.H && on < P.B && st_xs < xe) {
      uint32_t a0 = sbase + st_off;
      uint32_t a1 = a0 + rpitch, a2 = a1 + rpitch;
      uint32_t ae = sbase + e_off;
      f2 win[3][3][2];
      auto load_col = [&](int slot) {
        const uint2 v0 = lds8(a0), v1 = lds8(a1), v2 = lds8(a2);
        a0 += rowbytes; a1 += rowbytes; a2 += rowbytes;
        win[slot][0][0] = f2_from_bf16x2(v0.x); win[slot][0][1] = f2_from_bf16x2(v0.y);
        win[slot][1][0] = f2_from_bf16x2(v1.x); win[slot][1][1] = f2_from_bf16x2(v1.y);
        win[slot][2][0] = f2_from_bf16x2(v2.x); win[slot][2][1] = f2_from_bf16x2(v2.y);
      };
      if (!grp_b) {
        char* yp = reinterpret_cast<char*>(P.y + (((size_t)on * P.H + oy) * P.W + x0 + st_xs) * P.C + cq);
        auto emit = [&](int sl, int sm, int sr) {   // window slots holding dd columns x-1, x, x+1
          const uint2 ev = lds8(ae);
          ae += rowbytes;
          const f2 e[2] = {f2_from_bf16x2(ev.x), f2_from_bf16x2(ev.y)};
          const f2 z[2] = {f2_fma(e[0], esc0, esh0), f2_fma(e[1], esc1, esh1)};
          uint2 o;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            f2 a = 0ull;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
              a = F2_STENCIL_FMA(wa[r * 3 + 0][h], win[sl][r][h], a);
              a = F2_STENCIL_FMA(wa[r * 3 + 1][h], win[sm][r][h], a);
              a = F2_STENCIL_FMA(wa[r * 3 + 2][h], win[sr][r][h], a);
            }
            float alo, ahi, zlo, zhi;
            f2_split(a, alo, ahi);
            f2_split(z[h], zlo, zhi);
            alo = (zlo > 0.f && zlo < 6.f) ? alo : 0.f;
            ahi = (zhi > 0.f && zhi < 6.f) ? ahi : 0.f;
            const f2 m = f2_make(alo, ahi);
            s0[h] = f2_add(s0[h], m);
            s1[h] = f2_fma(m, e[h], s1[h]);
            (h == 0 ? o.x : o.y) = f2_to_bf16x2(m);
          }
          *reinterpret_cast<uint2*>(yp) = o;
          yp += ystep;
        };
        load_col(0);
        load_col(1);
        int x = st_xs;
#pragma unroll 1
        for (; x + 3 <= xe; x += 3) {
          load_col(2); emit(0, 1, 2);
          load_col(0); emit(1, 2, 0);
          load_col(1); emit(2, 0, 1);
        }
        if (x < xe) { load_col(2); emit(0, 1, 2); ++x; }
        if (x < xe) { load_col(0); emit(1, 2, 0); ++x; }
      } else {
        auto accum = [&](int sl, int sm, int sr) {
          const uint2 ev = lds8(ae);
          ae += rowbytes;
          // a = the activation the forward pass fed to the conv: relu6 of the bf16-rounded BatchNorm output
          const f2 act[2] = {f2_from_bf16x2(bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(ev.x), esc0, esh0)))),
                             f2_from_bf16x2(bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(ev.y), esc1, esh1))))};
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int r = 0; r < 3; ++r) {
              // window (row r, column slot c) is dd[h+1-ky][w+1-kx] for ky = 2-r, kx = 2-c
              wa[8 - (r * 3 + 0)][h] = F2_STENCIL_FMA(act[h], win[sl][r][h], wa[8 - (r * 3 + 0)][h]);
              wa[8 - (r * 3 + 1)][h] = F2_STENCIL_FMA(act[h], win[sm][r][h], wa[8 - (r * 3 + 1)][h]);
              wa[8 - (r * 3 + 2)][h] = F2_STENCIL_FMA(act[h], win[sr][r][h], wa[8 - (r * 3 + 2)][h]);
            }
        };
        load_col(0);
        load_col(1);
        int x = st_xs;
#pragma unroll 1
        for (; x + 3 <= xe; x += 3) {
          load_col(2); accum(0, 1, 2);
          load_col(0); accum(1, 2, 0);
          load_col(1); accum(2, 0, 1);
        }
        if (x < xe) { load_col(2); accum(0, 1, 2); ++x; }
        if (x < xe) { load_col(0); accum(1, 2, 0); ++x; }
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[stage]);
    if (threadIdx.x == 0 && tile + BW_NST * tstep < P.ntiles) {
      mbar_wait(&empty[stage], par);
      issue(tile + BW_NST * tstep, stage);
    }
  }
  if (active) {
    if (!grp_b) {
      const uint32_t sa = smem_u32(s_stats) + 16 * j;
      red_shared_f2(sa, s0[0]); red_shared_f2(sa + 8, s0[1]);
      red_shared_f2(sa + 256, s1[0]); red_shared_f2(sa + 264, s1[1]);
    } else {
      const uint32_t sa = smem_u32(s_dw) + 16 * j;
#pragma unroll
      for (int k = 0; k < 9; ++k) { red_shared_f2(sa + 256 * k, wa[k][0]); red_shared_f2(sa + 256 * k + 8, wa[k][1]); }
    }
  }
  named_bar_sync(1, BW_THREADS);
  for (int i = threadIdx.x; i < 11 * w; i += BW_THREADS) {
    const int h = i / w, c = i - h * w;
    if (h < 2) atomicAdd(P.stats + h * P.stats_stride + c0 + c, s_stats[h * 64 + c]);
    else atomicAdd(P.dwg + (h - 2) * P.C + c0 + c, s_dw[(h - 2) * 64 + c]);
  }
}


// ---------------------------------------------------------------------------------------------------------------
// Backward, stride 2.  A tile is Ht x Wt positions of the conv OUTPUT; it owns the 2Ht x 2Wt input pixels under it.
// Boxes per tile: dz and d_raw {w, Wt+1, Ht+1, NI} at (x0, y0) -- the extra row / column is what odd input rows /
// columns of the tile's last line reach -- and e_raw {w, 2Wt, 2Ht, NI} at (2x0, 2y0).
// Input pixel (h, v) gets, with a = h/2, b = v/2:
//   h even: ky = 1 from dd row a                      h odd: ky = 0 from dd row a+1, ky = 2 from dd row a
//   v even: kx = 1 from dd col b                      v odd: kx = 0 from dd col b+1, kx = 2 from dd col b
// A thread owns 4 channels and ONE input row (so the row parity, i.e. which taps exist, is a per-thread constant of the
// whole kernel) and walks column PAIRS (2b, 2b+1) with a two-column window of dd.  "Row A" = dd row a with
// ky = 1 (even rows) or 2 (odd rows); "row B" = dd row a+1 with ky = 0, weights zero on even rows.
// Warps 0-7: data gradient + mask + sums; warps 8-15: weight gradient (12 accumulator pairs per thread).
constexpr int DD_BYTES = 18 * 1024, E2_BYTES = 64 * 1024;
static_assert(2 * DD_BYTES + E2_BYTES == BW_STAGE, "stride-2 backward uses the same stage size");

__global__ void __launch_bounds__(BW_THREADS, 1)
dw_tile_bwd2_kernel(const __grid_constant__ CUtensorMap tmZ, const __grid_constant__ CUtensorMap tmG,
                    const __grid_constant__ CUtensorMap tmE, const __grid_constant__ DwTileParams P,
                    const __grid_constant__ BnBwdSrc bsrc) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint64_t* full = reinterpret_cast<uint64_t*>(base + BW_NST * BW_STAGE);
  uint64_t* empty = full + BW_NST;
  float* s_stats = reinterpret_cast<float*>(empty + BW_NST);     // [2][64]
  float* s_ecoef = s_stats + 128;                                 // [2][64]
  float* s_dcoef = s_ecoef + 128;                                 // [3][64]
  float* s_dw = s_dcoef + 192;                                    // [9][64]

  const int lane = threadIdx.x & 31;
  const int w = P.w, rowbytes = 2 * w;
  const int chunk = (int)blockIdx.x % P.nchunks, c0 = chunk * w;
  const int tile0 = (int)blockIdx.x / P.nchunks, tstep = (int)gridDim.x / P.nchunks;

  if (threadIdx.x == 0) {
    for (int i = 0; i < BW_NST; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], BW_THREADS / 32);
    }
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 128 + 9 * 64; i += BW_THREADS) (i < 128 ? s_stats[i] : s_dw[i - 128]) = 0.f;
  pdl_wait();
  if (threadIdx.x < w) {
    s_ecoef[threadIdx.x] = P.scale[c0 + threadIdx.x];
    s_ecoef[64 + threadIdx.x] = P.shift[c0 + threadIdx.x];
    float ca, cb, cc;
    if (bsrc.sums != nullptr) {
      bn_bwd_channel(bsrc, c0 + threadIdx.x, tile0 == 0, ca, cb, cc);
    } else {
      ca = P.dcoef[c0 + threadIdx.x];
      cb = P.dcoef[P.C + c0 + threadIdx.x];
      cc = P.dcoef[2 * P.C + c0 + threadIdx.x];
    }
    s_dcoef[threadIdx.x] = ca;
    s_dcoef[64 + threadIdx.x] = cb;
    s_dcoef[128 + threadIdx.x] = cc;
  }
  __syncthreads();

  const uint32_t dd_bytes = (uint32_t)(P.NI * P.Hp * P.Wp * rowbytes);          // Hp = Ht + 1, Wp = Wt + 1 (+ pad)
  const uint32_t e_bytes = (uint32_t)(P.NI * 2 * P.Ht * P.Wq * rowbytes);        // Wq = 2 Wt (+ pad)
  auto issue = [&](int tile, int stage) {
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    uint8_t* st = base + stage * BW_STAGE;
    mbar_arrive_expect_tx(&full[stage], 2 * dd_bytes + e_bytes);
    tma_load_4d(st, &tmZ, &full[stage], c0, x0, y0, n0);
    tma_load_4d(st + DD_BYTES, &tmG, &full[stage], c0, x0, y0, n0);
    tma_load_4d(st + 2 * DD_BYTES, &tmE, &full[stage], c0, 2 * x0, 2 * y0, n0);
  };
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmZ); prefetch_tmap(&tmG); prefetch_tmap(&tmE);
    for (int k = 0; k < BW_NST; ++k)
      if (tile0 + k * tstep < P.ntiles) issue(tile0 + k * tstep, k);
  }

  const bool grp_b = threadIdx.x >= BW_GROUP;
  const int t = threadIdx.x - (grp_b ? BW_GROUP : 0);
  const int Q = w >> 2;
  const int j = t % Q, sidx = t / Q;
  const int cq = c0 + 4 * j;
  // my strip: image, INPUT row hh of the tile (0 .. 2Ht-1), first column pair
  int st_img, st_hh, st_bs;
  {
    const int q1 = fastdiv(sidx, P.magic_nseg, P.NSEG);
    st_bs = (sidx - q1 * P.NSEG) * P.SL;
    st_img = fastdiv(q1, P.magic_ht, 2 * P.Ht);                  // magic_ht is the magic of 2*Ht here
    st_hh = q1 - st_img * 2 * P.Ht;
  }
  const bool active = sidx < P.nstrips;
  const int st_be = min(st_bs + P.SL, P.Wt);
  const int ph = st_hh & 1, ra = st_hh >> 1;                      // row parity, dd row "A"
  const int kyA = ph ? 2 : 1;                                     // tap row served by dd row A; row B serves ky = 0
  const uint32_t dd_off = ((st_img * P.Hp + ra) * P.Wp + st_bs) * rowbytes + (j << 3);
  const uint32_t e_off = 2 * DD_BYTES + ((st_img * 2 * P.Ht + st_hh) * P.Wq + 2 * st_bs) * rowbytes + (j << 3);
  const uint32_t rpitch = P.Wp * rowbytes;
  const size_t ystep = (size_t)P.C * 2;
  // group A: weights of rows A / B by kx; group B: the matching weight-gradient accumulators
  f2 wA[3][2], wB[3][2];
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
    float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
    if (!grp_b && active) {
      va = *reinterpret_cast<const float4*>(P.wts + (kyA * 3 + kx) * P.C + cq);
      if (ph) vb = *reinterpret_cast<const float4*>(P.wts + kx * P.C + cq);
    }
    wA[kx][0] = f2_make(va.x, va.y); wA[kx][1] = f2_make(va.z, va.w);
    wB[kx][0] = f2_make(vb.x, vb.y); wB[kx][1] = f2_make(vb.z, vb.w);
  }
  f2 s0[2], s1[2];
  s0[0] = s0[1] = s1[0] = s1[1] = 0ull;
  const f2 esc[2] = {f2_make(s_ecoef[4 * j], s_ecoef[4 * j + 1]), f2_make(s_ecoef[4 * j + 2], s_ecoef[4 * j + 3])};
  const f2 esh[2] = {f2_make(s_ecoef[64 + 4 * j], s_ecoef[64 + 4 * j + 1]),
                     f2_make(s_ecoef[64 + 4 * j + 2], s_ecoef[64 + 4 * j + 3])};
  const int U = w >> 3;
  const int uj = threadIdx.x % U, ur0 = threadIdx.x / U, urstep = BW_THREADS / U;
  const bool xf_active = ur0 < urstep;
  const uint32_t xf_off = ur0 * rowbytes + (uj << 4), xf_step = urstep * rowbytes;
  const uint32_t dc_addr = smem_u32(s_dcoef) + 32 * uj;

  int it = 0;
#pragma unroll 1
  for (int tile = tile0; tile < P.ntiles; tile += tstep, ++it) {
    const int stage = it % BW_NST, par = (it / BW_NST) & 1;
    int n0, y0, x0;
    decode_tile(P, tile, n0, y0, x0);
    const uint32_t sbase = smem_u32(base + stage * BW_STAGE);
    mbar_wait(&full[stage], par);
    // ---- pass 1: dd = ca*dz + cb*d_raw + cc, in place of dz; NaN (outside the image) -> 0
    if (xf_active) {
      const uint4 ka0 = lds16(dc_addr), ka1 = lds16(dc_addr + 16);
      const uint4 kb0 = lds16(dc_addr + 256), kb1 = lds16(dc_addr + 272);
      const uint4 kc0 = lds16(dc_addr + 512), kc1 = lds16(dc_addr + 528);
      const f2 ca[4] = {f2_pair(ka0.x, ka0.y), f2_pair(ka0.z, ka0.w), f2_pair(ka1.x, ka1.y), f2_pair(ka1.z, ka1.w)};
      const f2 cb[4] = {f2_pair(kb0.x, kb0.y), f2_pair(kb0.z, kb0.w), f2_pair(kb1.x, kb1.y), f2_pair(kb1.z, kb1.w)};
      const f2 cc[4] = {f2_pair(kc0.x, kc0.y), f2_pair(kc0.z, kc0.w), f2_pair(kc1.x, kc1.y), f2_pair(kc1.z, kc1.w)};
#pragma unroll 2
      for (uint32_t o = xf_off; o < dd_bytes; o += xf_step) {
        const uint4 z = lds16(sbase + o), g = lds16(sbase + DD_BYTES + o);
        uint4 v;
        v.x = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[0], f2_from_bf16x2(z.x), f2_fma(cb[0], f2_from_bf16x2(g.x), cc[0]))));
        v.y = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[1], f2_from_bf16x2(z.y), f2_fma(cb[1], f2_from_bf16x2(g.y), cc[1]))));
        v.z = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[2], f2_from_bf16x2(z.z), f2_fma(cb[2], f2_from_bf16x2(g.z), cc[2]))));
        v.w = bf16x2_nan0(f2_to_bf16x2(f2_fma(ca[3], f2_from_bf16x2(z.w), f2_fma(cb[3], f2_from_bf16x2(g.w), cc[3]))));
        sts16(sbase + o, v);
      }
    }
    named_bar_sync(1, BW_THREADS);
    // ---- pass 2: my input row, column pairs [st_bs, be)
    const int h = 2 * y0 + st_hh, on = n0 + st_img;
    const int vcol0 = 2 * (x0 + st_bs);                             // first input column of my strip
    const int be = min(st_be, (P.W - 2 * x0 + 1) >> 1);
    if (active && h < P.H && on < P.B && st_bs < be) {
      uint32_t aA = sbase + dd_off, aB = aA + rpitch, ae = sbase + e_off;
      f2 cA[2], cB[2], nA[2], nB[2];
      auto load_dd = [&](f2 (&xa)[2], f2 (&xb)[2]) {
        const uint2 va = lds8(aA), vb = lds8(aB);
        aA += rowbytes; aB += rowbytes;
        xa[0] = f2_from_bf16x2(va.x); xa[1] = f2_from_bf16x2(va.y);
        xb[0] = f2_from_bf16x2(vb.x); xb[1] = f2_from_bf16x2(vb.y);
      };
      load_dd(cA, cB);
      int v = vcol0;
      if (!grp_b) {
        char* yp = reinterpret_cast<char*>(P.y + (((size_t)on * P.H + h) * P.W + vcol0) * P.C + cq);
        // mask with ReLU6' of the input activation, accumulate the sums, store one pixel
        auto finish = [&](const uint2 ev, const f2 (&acc)[2], char* dst) {
          uint2 o;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const f2 e = f2_from_bf16x2(hf == 0 ? ev.x : ev.y);
            const f2 z = f2_fma(e, esc[hf], esh[hf]);
            float alo, ahi, zlo, zhi;
            f2_split(acc[hf], alo, ahi);
            f2_split(z, zlo, zhi);
            alo = (zlo > 0.f && zlo < 6.f) ? alo : 0.f;
            ahi = (zhi > 0.f && zhi < 6.f) ? ahi : 0.f;
            const f2 m = f2_make(alo, ahi);
            s0[hf] = f2_add(s0[hf], m);
            s1[hf] = f2_fma(m, e, s1[hf]);
            (hf == 0 ? o.x : o.y) = f2_to_bf16x2(m);
          }
          *reinterpret_cast<uint2*>(dst) = o;
        };
#pragma unroll 1
        for (int b = st_bs; b < be; ++b, v += 2) {
          load_dd(nA, nB);
          const uint2 e0 = lds8(ae), e1 = lds8(ae + rowbytes);
          ae += 2 * rowbytes;
          f2 d0[2], d1[2];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            d0[hf] = f2_fma(wB[1][hf], cB[hf], f2_fma(wA[1][hf], cA[hf], 0ull));
            d1[hf] = f2_fma(wA[0][hf], nA[hf], f2_fma(wA[2][hf], cA[hf], 0ull));
            d1[hf] = f2_fma(wB[0][hf], nB[hf], f2_fma(wB[2][hf], cB[hf], d1[hf]));
          }
          finish(e0, d0, yp);
          if (v + 1 < P.W) finish(e1, d1, yp + ystep);
          yp += 2 * ystep;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) { cA[hf] = nA[hf]; cB[hf] = nB[hf]; }
        }
      } else {
        // wA / wB are the weight-gradient accumulators of rows A / B here
#pragma unroll 1
        for (int b = st_bs; b < be; ++b, v += 2) {
          load_dd(nA, nB);
          const uint2 e0 = lds8(ae), e1 = lds8(ae + rowbytes);   // a pixel outside the image is NaN -> activation 0
          ae += 2 * rowbytes;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const f2 a0 = f2_from_bf16x2(bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(hf == 0 ? e0.x : e0.y), esc[hf], esh[hf]))));
            const f2 a1 = f2_from_bf16x2(bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(hf == 0 ? e1.x : e1.y), esc[hf], esh[hf]))));
            wA[1][hf] = f2_fma(a0, cA[hf], wA[1][hf]);
            wA[0][hf] = f2_fma(a1, nA[hf], wA[0][hf]);
            wA[2][hf] = f2_fma(a1, cA[hf], wA[2][hf]);
            wB[1][hf] = f2_fma(a0, cB[hf], wB[1][hf]);
            wB[0][hf] = f2_fma(a1, nB[hf], wB[0][hf]);
            wB[2][hf] = f2_fma(a1, cB[hf], wB[2][hf]);
            cA[hf] = nA[hf]; cB[hf] = nB[hf];
          }
        }
      }
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[stage]);
    if (threadIdx.x == 0 && tile + BW_NST * tstep < P.ntiles) {
      mbar_wait(&empty[stage], par);
      issue(tile + BW_NST * tstep, stage);
    }
  }
  if (active) {
    if (!grp_b) {
      const uint32_t sa = smem_u32(s_stats) + 16 * j;
      red_shared_f2(sa, s0[0]); red_shared_f2(sa + 8, s0[1]);
      red_shared_f2(sa + 256, s1[0]); red_shared_f2(sa + 264, s1[1]);
    } else {
      const uint32_t sa = smem_u32(s_dw) + 16 * j;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        red_shared_f2(sa + 256 * (kyA * 3 + kx), wA[kx][0]); red_shared_f2(sa + 256 * (kyA * 3 + kx) + 8, wA[kx][1]);
        if (ph) { red_shared_f2(sa + 256 * kx, wB[kx][0]); red_shared_f2(sa + 256 * kx + 8, wB[kx][1]); }
      }
    }
  }
  named_bar_sync(1, BW_THREADS);
  for (int i = threadIdx.x; i < 11 * w; i += BW_THREADS) {
    const int hh = i / w, c = i - hh * w;
    if (hh < 2) atomicAdd(P.stats + hh * P.stats_stride + c0 + c, s_stats[hh * 64 + c]);
    else atomicAdd(P.dwg + (hh - 2) * P.C + c0 + c, s_dw[(hh - 2) * 64 + c]);
  }
}

uint32_t magic(int d) { return (uint32_t)(((1ull << 32) + (uint64_t)d - 1) / (uint64_t)d); }

int encode_nhwc(CUtensorMap* out, const void* base, int B, int H, int W, int C, int box_c, int box_w, int box_h, int box_n,
                int estride = 1) {
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
  // element stride e in x and y: the box spans e*box_w x e*box_h pixels and every e-th one is loaded
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)(box_w * estride), (cuuint32_t)(box_h * estride), (cuuint32_t)box_n};
  cuuint32_t estr[4] = {1, (cuuint32_t)estride, (cuuint32_t)estride, 1};
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

// tile shape: maximise (thread utilisation) x (useful fraction of the haloed box) x (segment prologue) x (edge waste).
// Tiles are tiles of the OUTPUT (P.Ho x P.Wo); stride 2 loads four parity planes of (Ht+1) x (Wt+1) input pixels.
// Returns the score of the choice it leaves in P (so that a caller can compare chunk widths).
double pick_tile(DwTileParams& P, int halo_limit = STAGE_BYTES, int plain_limit = 0) {
  const int Q = P.wbox / 4;
  P.par = CONSUMERS / Q;
  const int rowbytes = 2 * P.wbox;
  const int hal = P.S == 1 ? 2 : 1, planes = P.S == 1 ? 1 : 4;
  double best = -1.0;
  const int wcand[4] = {P.Wo < 28 ? P.Wo : 28, 21, 14, 7};
  for (int wi = 0; wi < 4; ++wi) {
    const int Wt = wcand[wi];
    if (Wt > P.Wo || (wi > 0 && Wt >= wcand[0])) continue;
    int Wp = Wt + hal;
    if (P.wbox < 64 && (Wp % 2) == 0) ++Wp;     // odd pitch: rows of < 128 bytes then land in different banks
    for (int Ht = 1; Ht <= P.Ho && Ht <= 32; ++Ht)
      for (int NI = 1; NI <= 4 && NI <= P.B; ++NI) {
        const int Hp = Ht + hal;
        // planes start 128-byte aligned (TMA) AND on a row boundary (pass 1 walks all planes as one array of rows)
        const int palign = (128 % rowbytes == 0 || rowbytes % 128 == 0) ? (rowbytes > 128 ? rowbytes : 128) : 3 * 128;
        const int plane_bytes = (NI * Hp * Wp * rowbytes + palign - 1) / palign * palign;
        if (planes * plane_bytes > halo_limit) continue;
        if (plain_limit > 0 && NI * Ht * (Wp - 2) * rowbytes > plain_limit) continue;
        for (int NSEG = 1; NSEG <= 4; ++NSEG) {
          const int SL = (Wt + NSEG - 1) / NSEG;
          if (SL < (P.S == 1 ? 5 : 4) && NSEG > 1) continue;
          const int strips = NI * Ht * NSEG;
          if (strips > P.par) continue;             // one strip per thread: its geometry is hoisted out of the tile loop
          const double util = (double)strips / P.par * ((double)Wt / (NSEG * SL));
          const double halo = (double)(Ht * Wt) / (Hp * Wp);
          const double prol = P.S == 1 ? (double)SL / (SL + 2) : (double)SL / (SL + 0.5);
          const int tx = (P.Wo + Wt - 1) / Wt, ty = (P.Ho + Ht - 1) / Ht, tn = (P.B + NI - 1) / NI;
          const double edge = ((double)P.Wo / (tx * Wt)) * ((double)P.Ho / (ty * Ht)) * ((double)P.B / (tn * NI));
          const double size = planes * plane_bytes >= 16384 ? 1.0 : 0.9;          // small boxes: more barrier trips
          const double score = util * (0.5 + 0.5 * halo) * (0.7 + 0.3 * prol) * edge * size;
          if (score > best) {
            best = score;
            P.Wt = Wt; P.Ht = Ht; P.NI = NI; P.Hp = Hp; P.Wp = Wp; P.NSEG = NSEG; P.SL = SL;
            P.plane_stride = plane_bytes;
          }
        }
      }
  }
  P.nplanes = planes;
  P.Wq = P.Wp - 2;           // un-haloed box (backward): same pitch parity as the haloed one
  P.nstrips = P.NI * P.Ht * P.NSEG;
  P.tiles_x = (P.Wo + P.Wt - 1) / P.Wt;
  P.tiles_y = (P.Ho + P.Ht - 1) / P.Ht;
  P.ntiles = P.tiles_x * P.tiles_y * ((P.B + P.NI - 1) / P.NI);
  return best;
}

}  // namespace

bool dw_tile_supported(int B, int H, int W, int C, int stride) {      // forward AND backward
  return stride == 1 && chunk_width(C) > 0 && B >= 1 && H >= 1 && W >= 1 && H < 4096 && W < 4096;
}
bool dw_tile_fwd_supported(int B, int H, int W, int C, int stride) {  // the forward kernel also does stride 2
  return (stride == 1 || stride == 2) && chunk_width(C) > 0 && B >= 1 && H >= 1 && W >= 1 && H < 4096 && W < 4096;
}

int dw_tile_fwd_launch(const void* x, const float* scale, const float* shift, const float* wts, void* y, float* stats,
                       int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s, const BnFwdSrc* src) {
  B200DP_REQUIRE(dw_tile_fwd_supported(B, H, W, C, stride), "dw_tile_fwd: B=%d H=%d W=%d C=%d stride=%d unsupported", B, H,
                 W, C, stride);
  DwTileParams P{};
  P.scale = scale; P.shift = shift; P.wts = wts; P.y = static_cast<__nv_bfloat16*>(y);
  P.stats = stats; P.stats_stride = stats_stride;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.S = stride; P.Ho = (H - 1) / stride + 1; P.Wo = (W - 1) / stride + 1;
  if (stride == 1) {
    P.w = chunk_width(C);
    P.wbox = P.w;   // measured: loading 64-wide (conflict-free 128-byte) rows for a 48-wide chunk idles a quarter of the
                    // lanes and is slower than living with the bank conflicts of 96-byte rows (108 vs 90 us on 56x56x144)
    pick_tile(P);
  } else {
    // stride 2 moves four input pixels per output: tiles hold few outputs, so a narrower chunk (more strips per
    // tile for the same bytes) can keep more threads busy.  Try every equal split; narrow TMA rows cost a little.
    DwTileParams best = P;
    double best_score = -1.0;
    for (int w = 64; w >= 16; w -= 16) {
      if (C % w != 0) continue;
      DwTileParams T = P;
      T.w = T.wbox = w;
      const double row_eff = w >= 64 ? 1.0 : (w >= 48 ? 0.97 : (w >= 32 ? 0.92 : 0.75));
      const double sc = pick_tile(T) * row_eff;
      if (sc > best_score) { best_score = sc; best = T; }
    }
    B200DP_REQUIRE(best_score > 0.0, "dw_tile_fwd: no tile fits (C=%d)", C);
    P = best;
  }
  P.nchunks = C / P.w;
  P.magic_wp = magic(P.Wp); P.magic_hp = magic(P.Hp); P.magic_tx = magic(P.tiles_x); P.magic_ty = magic(P.tiles_y);
  P.magic_nseg = magic(P.NSEG); P.magic_ht = magic(P.Ht);
  CUtensorMap tm;
  if (int rc = encode_nhwc(&tm, x, B, H, W, C, P.wbox, P.Wp, P.Hp, P.NI, stride)) return rc;
  // CTA b owns chunk b % nchunks: the grid is a multiple of nchunks, at most two CTAs per SM
  int groups = 2 * num_sms() / P.nchunks;
  if (groups < 1) groups = 1;
  if (groups > P.ntiles) groups = P.ntiles;
  const int grid = groups * P.nchunks;
  const size_t smem = 128 + NST * STAGE_BYTES + 2 * NST * 8 + 256 * 4;
  if (stride == 1) {
    B200DP_CUDA(cudaFuncSetAttribute(dw_tile_fwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024));
    B200DP_CUDA(launch_k(dw_tile_fwd_kernel<1>, dim3((unsigned)grid), dim3(TILE_THREADS), smem, s, tm, P,
                         src ? *src : BnFwdSrc{}));
  } else {
    B200DP_CUDA(cudaFuncSetAttribute(dw_tile_fwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024));
    B200DP_CUDA(launch_k(dw_tile_fwd_kernel<2>, dim3((unsigned)grid), dim3(TILE_THREADS), smem, s, tm, P,
                         src ? *src : BnFwdSrc{}));
  }
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int dw_tile_bwd_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                       const float* dcoef, const float* wts, void* dze, float* stats, int stats_stride, float* dw, int B,
                       int H, int W, int C, cudaStream_t s, const BnBwdSrc* src) {
  B200DP_REQUIRE(dw_tile_supported(B, H, W, C, 1), "dw_tile_bwd: B=%d H=%d W=%d C=%d unsupported", B, H, W, C);
  B200DP_REQUIRE(stats != nullptr && dw != nullptr, "dw_tile_bwd: stats / dw missing");
  DwTileParams P{};
  P.scale = escale; P.shift = eshift; P.wts = wts; P.y = static_cast<__nv_bfloat16*>(dze);
  P.stats = stats; P.stats_stride = stats_stride; P.dcoef = dcoef; P.dwg = dw;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.S = 1; P.Ho = H; P.Wo = W;
  P.w = chunk_width(C);
  P.nchunks = C / P.w;
  P.wbox = P.w;
  pick_tile(P, HALO_BYTES, PLAIN_BYTES);
  P.magic_wp = magic(P.Wp); P.magic_hp = magic(P.Hp); P.magic_tx = magic(P.tiles_x); P.magic_ty = magic(P.tiles_y);
  P.magic_nseg = magic(P.NSEG); P.magic_ht = magic(P.Ht);
  CUtensorMap tmZ, tmG, tmE;
  if (int rc = encode_nhwc(&tmZ, dzd, B, H, W, C, P.w, P.Wp, P.Hp, P.NI)) return rc;
  if (int rc = encode_nhwc(&tmG, d_raw, B, H, W, C, P.w, P.Wp, P.Hp, P.NI)) return rc;
  if (int rc = encode_nhwc(&tmE, e_raw, B, H, W, C, P.w, P.Wq, P.Ht, P.NI)) return rc;
  int groups = num_sms() / P.nchunks;
  if (groups < 1) groups = 1;
  if (groups > P.ntiles) groups = P.ntiles;
  const int grid = groups * P.nchunks;
  const size_t smem = 128 + BW_NST * BW_STAGE + 2 * BW_NST * 8 + (128 + 128 + 192 + 576) * 4;
  B200DP_CUDA(cudaFuncSetAttribute(dw_tile_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  B200DP_CUDA(launch_k(dw_tile_bwd_kernel, dim3((unsigned)grid), dim3(BW_THREADS), smem, s, tmZ, tmG, tmE, P,
                       src ? *src : BnBwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

// stride-2 backward: tile of the conv output, strips = input rows x column-pair segments
static double pick_tile_bwd2(DwTileParams& P) {
  const int Q = P.w / 4, rowbytes = 2 * P.w;
  P.par = BW_GROUP / Q;
  double best = -1.0;
  const int wcand[4] = {P.Wo < 28 ? P.Wo : 28, 21, 14, 7};
  for (int wi = 0; wi < 4; ++wi) {
    const int Wt = wcand[wi];
    if (Wt > P.Wo || (wi > 0 && Wt >= wcand[0])) continue;
    int Wp = Wt + 1, Wq = 2 * Wt;
    if (P.w < 64) { if (Wp % 2 == 0) ++Wp; ++Wq; }          // odd pitches for rows narrower than 128 bytes
    for (int Ht = 1; Ht <= P.Ho && Ht <= 16; ++Ht)
      for (int NI = 1; NI <= 4 && NI <= P.B; ++NI) {
        const int Hp = Ht + 1;
        if (NI * Hp * Wp * rowbytes > DD_BYTES || NI * 2 * Ht * Wq * rowbytes > E2_BYTES) continue;
        for (int NSEG = 1; NSEG <= 4; ++NSEG) {
          const int SL = (Wt + NSEG - 1) / NSEG;
          if (SL < 3 && NSEG > 1) continue;
          const int strips = NI * 2 * Ht * NSEG;
          if (strips > P.par) continue;
          const double util = (double)strips / P.par * ((double)Wt / (NSEG * SL));
          const double halo = (double)(Ht * Wt) / (Hp * Wp);
          const double prol = (double)SL / (SL + 0.5);
          const int tx = (P.Wo + Wt - 1) / Wt, ty = (P.Ho + Ht - 1) / Ht, tn = (P.B + NI - 1) / NI;
          const double edge = ((double)P.Wo / (tx * Wt)) * ((double)P.Ho / (ty * Ht)) * ((double)P.B / (tn * NI));
          const double size = NI * 2 * Ht * Wq * rowbytes >= 16384 ? 1.0 : 0.9;
          const double score = util * (0.7 + 0.3 * halo) * (0.8 + 0.2 * prol) * edge * size;
          if (score > best) {
            best = score;
            P.Wt = Wt; P.Ht = Ht; P.NI = NI; P.Hp = Hp; P.Wp = Wp; P.Wq = Wq; P.NSEG = NSEG; P.SL = SL;
          }
        }
      }
  }
  P.nstrips = P.NI * 2 * P.Ht * P.NSEG;
  P.tiles_x = (P.Wo + P.Wt - 1) / P.Wt;
  P.tiles_y = (P.Ho + P.Ht - 1) / P.Ht;
  P.ntiles = P.tiles_x * P.tiles_y * ((P.B + P.NI - 1) / P.NI);
  return best;
}

bool dw_tile_bwd_supported(int B, int H, int W, int C, int stride) {
  return (stride == 1 || stride == 2) && chunk_width(C) > 0 && B >= 1 && H >= 1 && W >= 1 && H < 4096 && W < 4096;
}

int dw_tile_bwd2_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                        const float* dcoef, const float* wts, void* dze, float* stats, int stats_stride, float* dw, int B,
                        int H, int W, int C, cudaStream_t s, const BnBwdSrc* src) {
  B200DP_REQUIRE(dw_tile_bwd_supported(B, H, W, C, 2), "dw_tile_bwd2: B=%d H=%d W=%d C=%d unsupported", B, H, W, C);
  B200DP_REQUIRE(stats != nullptr && dw != nullptr, "dw_tile_bwd2: stats / dw missing");
  DwTileParams P{};
  P.scale = escale; P.shift = eshift; P.wts = wts; P.y = static_cast<__nv_bfloat16*>(dze);
  P.stats = stats; P.stats_stride = stats_stride; P.dcoef = dcoef; P.dwg = dw;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.S = 2; P.Ho = (H - 1) / 2 + 1; P.Wo = (W - 1) / 2 + 1;
  DwTileParams best = P;
  double best_score = -1.0;
  for (int w = 64; w >= 16; w -= 16) {
    if (C % w != 0) continue;
    DwTileParams T = P;
    T.w = T.wbox = w;
    const double row_eff = w >= 64 ? 1.0 : (w >= 48 ? 0.97 : (w >= 32 ? 0.92 : 0.75));
    const double sc = pick_tile_bwd2(T) * row_eff;
    if (sc > best_score) { best_score = sc; best = T; }
  }
  B200DP_REQUIRE(best_score > 0.0, "dw_tile_bwd2: no tile fits (C=%d)", C);
  P = best;
  P.nchunks = C / P.w;
  P.magic_wp = magic(P.Wp); P.magic_hp = magic(P.Hp); P.magic_tx = magic(P.tiles_x); P.magic_ty = magic(P.tiles_y);
  P.magic_nseg = magic(P.NSEG); P.magic_ht = magic(2 * P.Ht);
  CUtensorMap tmZ, tmG, tmE;
  if (int rc = encode_nhwc(&tmZ, dzd, B, P.Ho, P.Wo, C, P.w, P.Wp, P.Hp, P.NI)) return rc;
  if (int rc = encode_nhwc(&tmG, d_raw, B, P.Ho, P.Wo, C, P.w, P.Wp, P.Hp, P.NI)) return rc;
  if (int rc = encode_nhwc(&tmE, e_raw, B, H, W, C, P.w, P.Wq, 2 * P.Ht, P.NI)) return rc;
  int groups = num_sms() / P.nchunks;
  if (groups < 1) groups = 1;
  if (groups > P.ntiles) groups = P.ntiles;
  const int grid = groups * P.nchunks;
  const size_t smem = 128 + BW_NST * BW_STAGE + 2 * BW_NST * 8 + (128 + 128 + 192 + 576) * 4;
  B200DP_CUDA(cudaFuncSetAttribute(dw_tile_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  B200DP_CUDA(launch_k(dw_tile_bwd2_kernel, dim3((unsigned)grid), dim3(BW_THREADS), smem, s, tmZ, tmG, tmE, P,
                       src ? *src : BnBwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_dw_fwd_tile(const void* x, const float* scale, const float* shift, const float* w, void* y,
                                  float* stats, int stats_stride, int B, int H, int W, int C, int stride, void* stream) {
  return dw_tile_fwd_launch(x, scale, shift, w, y, stats, stats_stride, B, H, W, C, stride,
                            static_cast<cudaStream_t>(stream), nullptr);
}

extern "C" int b200dp_dw_bwd_tile(const void* e_raw, const float* escale, const float* eshift, const void* dzd,
                                  const void* d_raw, const float* dcoef, const float* w, void* dze, float* stats,
                                  int stats_stride, float* dw, int B, int H, int W, int C, int stride, void* stream) {
  if (stride == 2)
    return dw_tile_bwd2_launch(e_raw, escale, eshift, dzd, d_raw, dcoef, w, dze, stats, stats_stride, dw, B, H, W, C,
                               static_cast<cudaStream_t>(stream), nullptr);
  return dw_tile_bwd_launch(e_raw, escale, eshift, dzd, d_raw, dcoef, w, dze, stats, stats_stride, dw, B, H, W, C,
                            static_cast<cudaStream_t>(stream), nullptr);
}
