// Depthwise 3x3 convolution on the 5th-generation tensor cores (K2, K4, K5) -- EXPERIMENT, not on the
// training path (opt-in through b200dp_dw_fwd_tc / ops.dw_fwd(impl="tc")): correct, tested, and slower than the
// CUDA-core tile kernel of dw_tile.cu, because every tcgen05.mma costs 100-200 cycles however small it is (measured
// with the B200DP_DWTC_DBG switches below) and this formulation needs 36 tiny ones per 128 pixels x 64 channels.
//
// A depthwise conv has no reduction over channels, so it is not a GEMM -- but it IS nine GEMMs with diagonal
// weight matrices:   y[p, c] = sum_t  a[p + shift_t, c] * w_t[c]  =  sum_t ( A_t * diag(w_t) )[p, c].
// A_t is the SAME shared-memory tile for every tap, only started `shift_t` pixel rows later: with pixels as rows
// (one 2w-byte row per pixel for a chunk of w channels, TMA-swizzled) a tcgen05 shared-memory descriptor may start
// at any row, because the hardware swizzle is a function of the absolute shared-memory address
// (experiments/exp_desc_shift.cu verifies this for the 128/64/32-byte swizzles and the unswizzled layout).  So:
//
//   TMA   one 4-D box {w channels, Wt+2, Ht+2, NI images} of the RAW producer output, halo zero-filled
//   8 warps  rewrite it in place: relu6(scale*x + shift), and 0 where the halo is outside the image (padding is
//            post-activation)                                         -- BatchNorm + ReLU6 of the producer, fused
//   1 thread  per 128 flattened output positions q (pitch Wp = Wt+2): 9 taps x (w/16) MMAs, M=128, N=16, K=16,
//            A = tile + (q0 + ky*Wp + kx) rows, B = 16x16 diag(w_t) (bf16, built once per CTA), D in TMEM
//   8 warps  TMEM -> bf16 -> swizzled staging -> coalesced 16-byte global stores of the valid positions
//            + per-channel sum / sum of squares for THIS conv's BatchNorm            -- statistics, fused
//
// Positions whose x or y falls into the halo columns / rows are computed and dropped (7-25 % of the MMA work);
// the tensor pipe has two orders of magnitude of headroom here, what counts is that the CUDA cores execute ~10
// instructions per element (transform + epilogue) instead of ~55 in the register-sliding-window kernel
// (dwconv.cu), and that no registers are held across global-memory latency.
// Stride 2 uses four parity planes of the input (TMA boxes with element stride 2 in x and y), so every tap is
// again a plain row shift inside one plane.
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

constexpr int NSTAGE = 2;
constexpr int STAGE_BYTES = 64 * 1024;            // up to 512 pixel rows of 128 bytes (all planes of one tile)
constexpr int STAGING_BYTES = 8 * 4096;           // 8 epilogue warps x 32 rows x 128 bytes
constexpr int BMAT_BYTES = 9 * 4 * 512;           // 9 taps x 4 channel slices x (16 x 16 bf16)
constexpr int TC_THREADS = 640;                   // 4 control warps + 8 epilogue warps + 8 transform warps
constexpr int XF_WARP0 = 12, XF_WARPS = 8, XF_THREADS = XF_WARPS * 32;
constexpr int MAX_CHUNKS = 16;
constexpr int ACC_COLS = 64, TMEM_COLS = 128;

struct DwTcChunk { int c0, w, cta0, nctas; };

struct DwTcParams {
  const float* scale;       // [C] BatchNorm scale / shift of the producer
  const float* shift;
  const float* wts;         // [9][C] fp32, tap-major
  __nv_bfloat16* y;         // [B, Ho, Wo, C]
  float* stats;             // [2][stats_stride] or null
  int stats_stride;
  int B, H, W, C, Ho, Wo, S;
  int Ht, Wt, NI, Hp, Wp, nplanes, plane_rows, nmb;
  int RB, NB;               // one TMA box = RB plane rows x NB images (RB | Hp; NB = NI when RB == Hp, else 1)
  int tiles_x, tiles_y, ntiles;
  uint32_t magic_wp, magic_hp, magic_tx, magic_ty;   // ceil(2^32 / d) for the small divisors used per row / tile
  int nchunks;
  int dbg;                  // timing experiments (B200DP_DWTC_DBG): 1 no MMAs, 2 no transform math, 4 no epilogue sweep
  DwTcChunk chunk[MAX_CHUNKS];
};

__device__ __forceinline__ void sts16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr)
               : "memory");
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

// swizzle phase of pixel row r for rows of 2w bytes (w = 64 / 32 / 16): what TMA and the UMMA descriptors apply
__device__ __forceinline__ int swz(int r, int w) { return w == 64 ? (r & 7) : (w == 32 ? ((r >> 1) & 3) : ((r >> 2) & 1)); }

// r / d for r, d < 65536 with m = ceil(2^32 / d) (which does not fit 32 bits for d = 1)
__device__ __forceinline__ int fastdiv(int r, uint32_t m, int d) { return d == 1 ? r : (int)__umulhi((uint32_t)r, m); }

// flattened row of a plane -> (image, y, x) inside the plane
__device__ __forceinline__ void decode_row(const DwTcParams& P, int r, int& img, int& yy, int& xx) {
  const int t = fastdiv(r, P.magic_wp, P.Wp);
  xx = r - t * P.Wp;
  img = fastdiv(t, P.magic_hp, P.Hp);
  yy = t - img * P.Hp;
}
__device__ __forceinline__ void decode_tile(const DwTcParams& P, int tile, int& n0, int& y0, int& x0) {
  const int t = fastdiv(tile, P.magic_tx, P.tiles_x);
  x0 = (tile - t * P.tiles_x) * P.Wt;
  const int tn = fastdiv(t, P.magic_ty, P.tiles_y);
  y0 = (t - tn * P.tiles_y) * P.Ht;
  n0 = tn * P.NI;
}

__global__ void __launch_bounds__(TC_THREADS, 1)
dw_tc_fwd_kernel(const __grid_constant__ CUtensorMap tm64, const __grid_constant__ CUtensorMap tm32,
                 const __grid_constant__ CUtensorMap tm16, const __grid_constant__ DwTcParams P) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage0 = base;
  uint8_t* staging = base + NSTAGE * STAGE_BYTES;
  uint8_t* bmat = staging + STAGING_BYTES;
  float* s_stats = reinterpret_cast<float*>(bmat + BMAT_BYTES);          // [2][64]
  int* s_rowinfo = reinterpret_cast<int*>(s_stats + 128);                 // [8][32]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_rowinfo + 256);
  uint64_t* raw_full = bars;                  // TMA landed
  uint64_t* xf_done = bars + NSTAGE;          // transformed in place
  uint64_t* stage_empty = bars + 2 * NSTAGE;  // all MMAs reading the stage have completed
  uint64_t* tmem_full = bars + 3 * NSTAGE;    // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // ---- which channel chunk does this CTA own?
  int ci = 0;
  for (int i = 1; i < P.nchunks; ++i)
    if ((int)blockIdx.x >= P.chunk[i].cta0) ci = i;
  const int c0 = P.chunk[ci].c0, w = P.chunk[ci].w;
  const int cta_in = (int)blockIdx.x - P.chunk[ci].cta0, nctas = P.chunk[ci].nctas;
  const int lsh = w == 64 ? 3 : (w == 32 ? 2 : 1);      // log2(16-byte chunks per row)
  const int cpr = 1 << lsh;
  const int rowbytes = 2 * w;
  const int plane_stride = (P.plane_rows * rowbytes + 1023) & ~1023;
  const int nslice = w >> 4;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NSTAGE; ++i) {
      mbar_init(&raw_full[i], 1);
      mbar_init(&xf_done[i], XF_WARPS);
      mbar_init(&stage_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
  pdl_wait();
  // ---- B operands: 16x16 diagonal matrices, K-major rows of 32 bytes with the 32-byte swizzle
  for (int i = threadIdx.x; i < BMAT_BYTES / 16; i += blockDim.x) reinterpret_cast<uint4*>(bmat)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < 128; i += blockDim.x) s_stats[i] = 0.f;
  __syncthreads();
  for (int i = threadIdx.x; i < 9 * nslice * 16; i += blockDim.x) {
    const int n = i & 15, s = (i >> 4) % nslice, t = (i >> 4) / nslice;
    const float wv = P.wts[t * P.C + c0 + 16 * s + n];
    const int off = (t * 4 + s) * 512 + n * 32 + ((((n >> 3) ^ ((n >> 2) & 1))) << 4) + (n & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(bmat + off) = __float2bfloat16(wv);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      const CUtensorMap* tm = w == 64 ? &tm64 : (w == 32 ? &tm32 : &tm16);
      prefetch_tmap(tm);
      int it = 0;
      for (int tile = cta_in; tile < P.ntiles; tile += nctas, ++it) {
        const int stage = it % NSTAGE, par = (it / NSTAGE) & 1;
        mbar_wait(&stage_empty[stage], par ^ 1);
        int n0, y0, x0;
        decode_tile(P, tile, n0, y0, x0);
        mbar_arrive_expect_tx(&raw_full[stage], (uint32_t)(P.nplanes * P.plane_rows * rowbytes));
        uint8_t* dst = stage0 + stage * STAGE_BYTES;
        // several smaller boxes per plane: the copy engine overlaps boxes, not the rows of one box
        const int box_bytes = P.RB * P.Wp * rowbytes;
        for (int p = 0; p < P.nplanes; ++p) {
          const int bx = P.S == 1 ? x0 - 1 : 2 * (x0 - 1) + (p & 1);
          const int by = P.S == 1 ? y0 - 1 : 2 * (y0 - 1) + (p >> 1);
          uint8_t* pd = dst + p * plane_stride;
          for (int img = 0; img < P.NI; img += P.NB)
            for (int yb = 0; yb < P.Hp; yb += P.RB, pd += box_bytes * P.NB)
              tma_load_4d(pd, tm, &raw_full[stage], c0, bx, by + yb * P.S, n0 + img);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, 16, 0, 0);
      const uint32_t bbase = smem_u32(bmat);
      int it = 0, m = 0;
      for (int tile = cta_in; tile < P.ntiles; tile += nctas, ++it) {
        const int stage = it % NSTAGE, par = (it / NSTAGE) & 1;
        mbar_wait(&xf_done[stage], par);
        tc_fence_after();
        const uint32_t sbase = smem_u32(stage0 + stage * STAGE_BYTES);
        for (int mb = 0; mb < P.nmb; ++mb, ++m) {
          const int slot = m & 1;
          mbar_wait(&tmem_empty[slot], ((m >> 1) & 1) ^ 1);
          tc_fence_after();
          if (P.dbg & 32) {   // experiment: same-accumulator MMAs back to back
            for (int s = 0; s < nslice; ++s)
              for (int t = 0; t < 9; ++t) {
                const int ky = t / 3, kx = t - 3 * ky;
                int plane = 0, off = ky * P.Wp + kx;
                if (P.S == 2) { plane = 2 * (ky != 1) + (kx != 1); off = (ky > 0) * P.Wp + (kx > 0); }
                const uint32_t arow = sbase + plane * plane_stride + (mb * 128 + off) * rowbytes;
                umma_bf16(tmem_base + slot * ACC_COLS + s * 16, umma_desc_kmajor(arow + s * 32, w),
                          umma_desc_kmajor(bbase + (t * 4 + s) * 512, 16), idesc, t > 0 ? 1u : 0u);
              }
          } else
          for (int t = 0; t < ((P.dbg & 64) ? 3 : 9); ++t) {
            const int ky = t / 3, kx = t - 3 * ky;
            int plane = 0, off = ky * P.Wp + kx;
            if (P.S == 2) { plane = 2 * (ky != 1) + (kx != 1); off = (ky > 0) * P.Wp + (kx > 0); }
            const uint32_t arow = sbase + plane * plane_stride + (mb * 128 + off) * rowbytes;
            if (P.dbg & 1) continue;
            for (int s = 0; s < nslice; ++s)
              umma_bf16(tmem_base + slot * ACC_COLS + s * 16, umma_desc_kmajor(arow + s * 32, w),
                        umma_desc_kmajor(bbase + (t * 4 + s) * 512, 16), idesc, t > 0 ? 1u : 0u);
          }
          umma_commit(&tmem_full[slot]);
        }
        umma_commit(&stage_empty[stage]);
      }
    }
  } else if (warp >= 4 && warp < 12) {
    // ================= epilogue: TMEM -> bf16 -> staging -> coalesced stores + statistics =================
    const int e = warp - 4, g = e >> 2, qd = warp & 3;
    const uint32_t stg = smem_u32(staging + e * 4096);
    int* rinfo = s_rowinfo + e * 32;
    const int j = lane & (cpr - 1), rsub = lane >> lsh, rpp = 32 >> lsh;
    float s0[8], s1[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s0[i] = s1[i] = 0.f;
    __nv_bfloat16* ybase = P.y + c0 + 8 * j;
    const int c16 = P.C >> 3;                       // 16-byte units per pixel
    int m = 0;
    for (int tile = cta_in; tile < P.ntiles; tile += nctas) {
      int n0, y0, x0;
      decode_tile(P, tile, n0, y0, x0);
      for (int mb = 0; mb < P.nmb; ++mb, ++m) {
        if ((m & 1) != g) continue;
        // where does my row go?  (computed while the MMAs run)
        {
          const int q = mb * 128 + qd * 32 + lane;
          int img, yy, xx;
          decode_row(P, q, img, yy, xx);
          const bool ok = img < P.NI && yy < P.Ht && xx < P.Wt && (y0 + yy) < P.Ho && (x0 + xx) < P.Wo && (n0 + img) < P.B;
          rinfo[lane] = ok ? (((n0 + img) * P.Ho + (y0 + yy)) * P.Wo + (x0 + xx)) * c16 : -1;
        }
        mbar_wait(&tmem_full[g], (m >> 1) & 1);
        tc_fence_after();
        const int sw = swz(lane, w);
        for (int s = 0; s < ((P.dbg & 16) ? 0 : nslice); ++s) {
          uint32_t v[16];
          tmem_ld_x16(tmem_base + ((uint32_t)(qd * 32) << 16) + g * ACC_COLS + s * 16, v);
          tmem_ld_wait();
          uint4 lo, hi;
          lo.x = pack_bf16x2(__uint_as_float(v[0]), __uint_as_float(v[1]));
          lo.y = pack_bf16x2(__uint_as_float(v[2]), __uint_as_float(v[3]));
          lo.z = pack_bf16x2(__uint_as_float(v[4]), __uint_as_float(v[5]));
          lo.w = pack_bf16x2(__uint_as_float(v[6]), __uint_as_float(v[7]));
          hi.x = pack_bf16x2(__uint_as_float(v[8]), __uint_as_float(v[9]));
          hi.y = pack_bf16x2(__uint_as_float(v[10]), __uint_as_float(v[11]));
          hi.z = pack_bf16x2(__uint_as_float(v[12]), __uint_as_float(v[13]));
          hi.w = pack_bf16x2(__uint_as_float(v[14]), __uint_as_float(v[15]));
          sts16(stg + lane * rowbytes + (((2 * s) ^ sw) << 4), lo);
          sts16(stg + lane * rowbytes + (((2 * s + 1) ^ sw) << 4), hi);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[g]);
        for (int pass = 0; pass < ((P.dbg & 4) ? 0 : cpr); ++pass) {
          const int rr = pass * rpp + rsub;
          const int info = rinfo[rr];
          if (info >= 0) {
            const uint4 v = lds16(stg + rr * rowbytes + ((j ^ swz(rr, w)) << 4));
            stg_stream(ybase + (size_t)info * 8, v);
            float f[8];
            unpack8(v, f);
#pragma unroll
            for (int i = 0; i < 8; ++i) { s0[i] += f[i]; s1[i] = fmaf(f[i], f[i], s1[i]); }
          }
        }
        __syncwarp();
      }
    }
    if (P.stats != nullptr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        atomicAdd(&s_stats[8 * j + i], s0[i]);
        atomicAdd(&s_stats[64 + 8 * j + i], s1[i]);
      }
    }
  } else if (warp >= XF_WARP0) {
    // ================= transform: BatchNorm + ReLU6 of the producer, in place; zero padding by position ==========
    const int t = threadIdx.x - XF_WARP0 * 32;
    const int j = t & (cpr - 1), r0 = t >> lsh, rstep = XF_THREADS >> lsh;
    float sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { sc[i] = P.scale[c0 + 8 * j + i]; sh[i] = P.shift[c0 + 8 * j + i]; }
    int it = 0;
    for (int tile = cta_in; tile < P.ntiles; tile += nctas, ++it) {
      const int stage = it % NSTAGE, par = (it / NSTAGE) & 1;
      int n0, y0, x0;
      decode_tile(P, tile, n0, y0, x0);
      mbar_wait(&raw_full[stage], par);
      for (int p = 0; p < P.nplanes; ++p) {
        const uint32_t pbase = smem_u32(stage0 + stage * STAGE_BYTES + p * plane_stride);
        for (int r = r0; r < ((P.dbg & 2) ? 0 : P.plane_rows); r += rstep) {
          int img, yy, xx;
          decode_row(P, r, img, yy, xx);
          int iy = y0 - 1 + yy, ix = x0 - 1 + xx;
          if (P.S == 2) { iy = 2 * iy + (p >> 1); ix = 2 * ix + (p & 1); }
          const bool ok = (unsigned)iy < (unsigned)P.H && (unsigned)ix < (unsigned)P.W && (n0 + img) < P.B;
          const uint32_t addr = pbase + r * rowbytes + ((j ^ swz(r, w)) << 4);
          uint4 v = make_uint4(0, 0, 0, 0);
          if (ok) {
            float f[8];
            unpack8(lds16(addr), f);
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = relu6f(fmaf(f[i], sc[i], sh[i]));
            v = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                           pack_bf16x2(f[6], f[7]));
          }
          sts16(addr, v);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&xf_done[stage]);
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (P.stats != nullptr && threadIdx.x < 2 * w) {
    const int h = threadIdx.x / w, c = threadIdx.x - h * w;
    atomicAdd(P.stats + h * P.stats_stride + c0 + c, s_stats[h * 64 + c]);
  }
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

uint32_t magic(int d) { return (uint32_t)(((1ull << 32) + (uint64_t)d - 1) / (uint64_t)d); }

int make_tmap_nhwc(CUtensorMap* out, const void* base, int B, int H, int W, int C, int box_c, int box_w, int box_h,
                   int box_n, int estride) {
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
  cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)(box_w * estride), (cuuint32_t)(box_h * estride), (cuuint32_t)box_n};
  cuuint32_t estr[4] = {1, (cuuint32_t)estride, (cuuint32_t)estride, 1};
  const CUtensorMapSwizzle sw = box_c == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                                            : (box_c == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(4d) failed (%d): dims %dx%dx%dx%d box %dx%dx%dx%d stride %d", (int)r, C, W, H, B,
              box_c, box_w, box_h, box_n, estride);
    return B200DP_ECUDA;
  }
  return B200DP_OK;
}

// channel chunks of 64 (+ one of 32 and/or 16) and the CTAs that work on each, proportional to estimated cost
int plan_chunks(DwTcParams& P, int C, int ntiles, int max_ctas) {
  int n = 0, c = 0;
  float cost[MAX_CHUNKS], total = 0.f;
  auto add = [&](int w) {
    P.chunk[n].c0 = c; P.chunk[n].w = w;
    cost[n] = w == 64 ? 5.2f : 3.0f;           // MMA-bound (w=64) vs TMA-row-bound (narrow rows), per tile
    total += cost[n];
    c += w; ++n;
  };
  while (C - c >= 64) { if (n >= MAX_CHUNKS) return -1; add(64); }
  if (C - c >= 32) { if (n >= MAX_CHUNKS) return -1; add(32); }
  if (C - c >= 16) { if (n >= MAX_CHUNKS) return -1; add(16); }
  if (c != C) return -1;
  P.nchunks = n;
  int ctas = max_ctas < n ? n : max_ctas;
  int used = 0;
  for (int i = 0; i < n; ++i) {
    int k = (int)(cost[i] / total * ctas);
    if (k < 1) k = 1;
    if (k > ntiles) k = ntiles;
    P.chunk[i].nctas = k;
    used += k;
  }
  // hand out what rounding left over, most expensive chunks first
  for (int guard = 0; used < ctas && guard < 4 * n; ++guard) {
    const int i = guard % n;
    if (P.chunk[i].nctas < ntiles) { ++P.chunk[i].nctas; ++used; }
  }
  int at = 0;
  for (int i = 0; i < n; ++i) { P.chunk[i].cta0 = at; at += P.chunk[i].nctas; }
  return at;
}

}  // namespace

bool dw_tc_supported(int B, int H, int W, int C, int stride) {
  const int nch = C / 64 + ((C % 64) >= 32) + ((C % 32) >= 16);
  return C % 16 == 0 && nch <= MAX_CHUNKS && (stride == 1 || stride == 2) && H >= 1 && W >= 1 && B >= 1;
}

int dw_tc_fwd_launch(const void* x, const float* scale, const float* shift, const float* wts, void* y, float* stats,
                     int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s) {
  B200DP_REQUIRE(dw_tc_supported(B, H, W, C, stride), "dw_tc_fwd: B=%d H=%d W=%d C=%d stride=%d unsupported", B, H, W, C,
                 stride);
  DwTcParams P{};
  P.scale = scale; P.shift = shift; P.wts = wts; P.y = static_cast<__nv_bfloat16*>(y);
  P.stats = stats; P.stats_stride = stats_stride;
  P.B = B; P.H = H; P.W = W; P.C = C; P.S = stride;
  P.Ho = (H - 1) / stride + 1; P.Wo = (W - 1) / stride + 1;
  if (stride == 1) {
    P.Wt = P.Wo < 28 ? P.Wo : 28;
    P.Ht = P.Ho < 14 ? P.Ho : 14;
    P.Hp = P.Ht + 2; P.Wp = P.Wt + 2; P.nplanes = 1;
    P.NI = 512 / (P.Hp * P.Wp);
  } else {
    P.Wt = P.Wo < 14 ? P.Wo : 14;
    P.Ht = P.Ho < 7 ? P.Ho : 7;
    P.Hp = P.Ht + 1; P.Wp = P.Wt + 1; P.nplanes = 4;
    P.NI = 128 / (P.Hp * P.Wp);
  }
  if (P.NI < 1) P.NI = 1;
  if (P.NI > B) P.NI = B;
  P.plane_rows = P.NI * P.Hp * P.Wp;
  {
    const char* rb = getenv("B200DP_DWTC_RB");
    P.RB = rb ? atoi(rb) : P.Hp;
    if (P.RB < 1 || P.RB > P.Hp || P.Hp % P.RB != 0) P.RB = P.Hp;
    P.NB = P.RB == P.Hp ? P.NI : 1;
  }
  B200DP_REQUIRE(P.nplanes * ((P.plane_rows * 128 + 1023) & ~1023) <= STAGE_BYTES, "dw_tc_fwd: tile does not fit a stage");
  const int qlast = (P.NI - 1) * P.Hp * P.Wp + (P.Ht - 1) * P.Wp + (P.Wt - 1);
  P.nmb = qlast / 128 + 1;
  P.tiles_x = (P.Wo + P.Wt - 1) / P.Wt;
  P.tiles_y = (P.Ho + P.Ht - 1) / P.Ht;
  P.ntiles = P.tiles_x * P.tiles_y * ((B + P.NI - 1) / P.NI);
  P.magic_wp = magic(P.Wp); P.magic_hp = magic(P.Hp); P.magic_tx = magic(P.tiles_x); P.magic_ty = magic(P.tiles_y);
  {
    const char* d = getenv("B200DP_DWTC_DBG");
    P.dbg = d ? atoi(d) : 0;
  }
  const int grid = plan_chunks(P, C, P.ntiles, num_sms());
  B200DP_REQUIRE(grid > 0, "dw_tc_fwd: cannot split C=%d into channel chunks", C);
  CUtensorMap tm[3];
  const int widths[3] = {64, 32, 16};
  bool have[3] = {false, false, false};
  for (int i = 0; i < P.nchunks; ++i) have[P.chunk[i].w == 64 ? 0 : (P.chunk[i].w == 32 ? 1 : 2)] = true;
  int first = -1;
  for (int k = 0; k < 3; ++k)
    if (have[k]) {
      if (int rc = make_tmap_nhwc(&tm[k], x, B, H, W, C, widths[k], P.Wp, P.RB, P.NB, stride)) return rc;
      if (first < 0) first = k;
    }
  for (int k = 0; k < 3; ++k)
    if (!have[k]) tm[k] = tm[first];
  const size_t smem = 1024 + NSTAGE * STAGE_BYTES + STAGING_BYTES + BMAT_BYTES + 128 * 4 + 256 * 4 + 16 * 8 + 16;
  B200DP_CUDA(cudaFuncSetAttribute(dw_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  B200DP_CUDA(launch_k(dw_tc_fwd_kernel, dim3((unsigned)grid), dim3(TC_THREADS), smem, s, tm[0], tm[1], tm[2], P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_dw_fwd_tc(const void* x, const float* scale, const float* shift, const float* w, void* y,
                                float* stats, int stats_stride, int B, int H, int W, int C, int stride, void* stream) {
  return dw_tc_fwd_launch(x, scale, shift, w, y, stats, stats_stride, B, H, W, C, stride,
                          static_cast<cudaStream_t>(stream));
}
