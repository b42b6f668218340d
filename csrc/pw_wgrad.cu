// Pointwise-convolution weight gradient on tcgen05 (SURVEY W5 "wgrad", K1).
//
//   dW[n][k] += sum_m G[m][n] * A[m][k]          (fp32, accumulated into the flat gradient arena)
//
// Both operands are [rows m][channels] NHWC tensors, i.e. the reduction index m is the *slow* index in memory:
// they are MN-major UMMA operands.  A TMA box {64 channels, MK rows} lands in shared memory as MK lines of 128
// bytes with the 128-byte swizzle, which IS the canonical MN-major tile -- no transpose pass exists anywhere.
//
// Every 1x1 conv in MobileNetV2 has one wide side ("hidden": 32..1280 channels) and one narrow side (16..320):
//   MODE 0 (expand / last conv):  hidden = G side = ca*dz + cb*y_raw + cc;  narrow = A side (block input).  The
//                                 affine map is per hidden channel, so it FACTORS out of the reduction over m:
//                                   dW[n][k] = ca[n] (dz^T A)[n][k] + cb[n] (y^T A)[n][k] + cc[n] sum_m A[m][k]
//                                 -> two MMAs on the raw bf16 tensors, no per-element transform at all; the
//                                 combination happens once per output element in the epilogue (fp32).
//   MODE 1 (project conv):        hidden = A side, formed as relu6(scale*d_raw + shift);  narrow = G side (dp)
// The hidden side is the UMMA M operand (128 TMEM lanes), the narrow side the N operand:
//   D[hidden chunk (128 lanes)][narrow chunk (<= 256 columns)] += H[m x 128]^T * Q[m x nq]
// Global->shared traffic is cp.async (LDGSTS, 16 bytes per thread) issued by 4 loader warps straight into the
// swizzled operand layout and tracked by mbarriers (cp.async.mbarrier.arrive.noinc): measured here, TMA costs
// ~4.5-8 cycles per box ROW whatever its width, which caps narrow NHWC rows (32-256 B) far below the HBM roofline,
// while LDGSTS has no per-row cost and holds no registers across the latency.  The raw hidden tile(s) are then
// rewritten in place by 4 transform warps (shared -> registers -> BatchNorm math -> shared, fence.proxy.async);
// the narrow side is consumed as loaded.  A tail of 16 / 32 channels uses lines of 32 / 64 bytes with the matching
// swizzle and, for the N operand, its own MMA into the next TMEM columns; missing lanes of a tail M chunk are zero
// blocks in shared memory.
// Work items = (hidden chunk) x (narrow chunk) x (split of the m range); each item's fp32 tile is added to dW with
// red.global.add (vectorised v4 when the narrow index is the contiguous one).
// Reference behaviour replaced: cuDNN convolution-backward-filter for the 34 1x1 convs.
#include <cstdlib>

#include "bn_inline.cuh"
#include "common.cuh"
#include "pw_gemm.h"
#include "sm100.cuh"

namespace b200dp {
namespace {

using namespace sm100;

constexpr int MAX_STAGES = 6;
constexpr int XF_THREADS = 128;        // transform warps 8..11
constexpr int LD_THREADS = 128;        // loader warps 12..15
constexpr int TMEM_COLS = 512;         // 2 accumulator stages x 256 columns

struct WgradParams {
  int M;                       // rows
  int hidden, narrow;          // channel counts of the wide / narrow side (after row packing)
  int hidden_true, narrow_true, pack;   // before packing: r consecutive rows are one row of r*C channels, only the
                               // r diagonal (hidden slot == narrow slot) blocks of the product are weight gradient
  int ldh, ldq, ldw;           // row pitches (elements): hidden tensors, narrow tensor, dW
  int mk;                      // m rows per pipeline stage (64 / 128 / 256)
  int nq, qblks, qchunks;      // narrow chunk width (multiple of 16), its 64-wide blocks, number of chunks
  int hchunks, msplit, chunks_per_split, stages;
  const float* coef;           // MODE 0: [3][hidden] ca, cb, cc     MODE 1: [2][hidden] scale, shift
  float* dw;
  const __nv_bfloat16* t0;     // hidden raw tensor 0 (MODE 0: dz, MODE 1: d_raw)
  const __nv_bfloat16* t1;     // hidden raw tensor 1 (MODE 0: y_raw)
  const __nv_bfloat16* qsrc;   // narrow tensor
  const float* asum;           // MODE 0: [narrow] column sums of the narrow tensor over all M rows
  int fuse_g;                  // MODE 0: 1 = form G = ca*dz + cb*y + cc in shared memory (transform warps) and issue ONE
                               // MMA per 16 rows instead of the factored two (these layers are MMA-issue bound)
  unsigned long long* trace;   // debugging: CTA 0 writes %globaltimer at its pipeline milestones (null = off)
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
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
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// 16-byte asynchronous global->shared copy (zero-fills when src_bytes == 0) and its mbarrier completion hook
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Roles (512 threads): warp 0 = MMA issuer + TMEM allocator | warps 4-7 = epilogue | warps 8-11 = transform (MODE 1)
// or publish-only (MODE 0) | warps 12-15 = cp.async loaders | warps 1-3 idle.
template <int MODE>
__global__ void __launch_bounds__(512, 1)
pw_wgrad_kernel(const __grid_constant__ WgradParams P, const __grid_constant__ BnBwdSrc bsrc) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int MK = P.mk;
  const int h_bytes = MK * 256;                                     // hidden tile: 2 blocks of 64 channels, 128-byte lines
  const int q_bytes = P.qblks * MK * 128;
  const int stage_bytes = h_bytes * (MODE == 0 ? 2 : 1) + q_bytes;
  const int stages = P.stages;
  float* s_coef = reinterpret_cast<float*>(base + (size_t)stages * stage_bytes);
  const int ncoef = (MODE == 0 ? 3 : 2) * P.hidden;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_coef + ((ncoef + 3) & ~3));
  uint64_t* full = bars;
  uint64_t* empty = bars + MAX_STAGES;
  uint64_t* raw_full = bars + 2 * MAX_STAGES;
  uint64_t* tmem_full = bars + 3 * MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto stamp = [&](int slot) {
    if (P.trace != nullptr && blockIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      P.trace[slot] = t;
    }
  };
  if (threadIdx.x == 0) stamp(0);
  const int items = P.hchunks * P.qchunks * P.msplit;
  const int total_chunks = (P.M + MK - 1) / MK;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], XF_THREADS / 32);      // one arrival per transform warp
      mbar_init(&empty[s], 1);
      mbar_init(&raw_full[s], LD_THREADS);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
  // operand tiles are zero-padded past the channels a chunk covers: start from all-zero stages
  for (int o = threadIdx.x * 16; o < stages * stage_bytes; o += blockDim.x * 16) sts16(smem_u32(base) + o, make_uint4(0u, 0u, 0u, 0u));
  // on-chip set-up above overlaps the previous kernel's tail (PDL); global data from here on
  pdl_trigger();
  pdl_wait();
  if (MODE == 0 && bsrc.sums != nullptr) {
    // BatchNorm backward of the hidden side finalised here (first consumer); CTA 0 publishes ca, cb, cc, dgamma, dbeta
    for (int c = threadIdx.x; c < P.hidden; c += blockDim.x) {
      float ca, cb, cc;
      bn_bwd_channel(bsrc, c % P.hidden_true, blockIdx.x == 0 && c < P.hidden_true, ca, cb, cc);
      s_coef[c] = ca;
      s_coef[P.hidden + c] = cb;
      s_coef[2 * P.hidden + c] = cc;
    }
  } else {
    for (int i = threadIdx.x; i < ncoef; i += blockDim.x) {
      const int j = i / P.hidden, c = i - j * P.hidden;
      s_coef[i] = P.coef[j * P.hidden_true + c % P.hidden_true];
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) stamp(1);

  // item -> hidden chunk (h0, hw <= 128), narrow chunk q0, m-chunk range.  Hidden chunk fastest: CTAs running side
  // by side read the same narrow rows, which stay in L2.
  auto decode = [&](int item, int& h0, int& hw, int& q0, int& c_lo, int& c_hi) {
    const int hc = item % P.hchunks;
    const int qc = (item / P.hchunks) % P.qchunks;
    const int ms = item / (P.hchunks * P.qchunks);
    h0 = hc * 128;
    hw = min(128, P.hidden - h0);
    q0 = qc * P.nq;
    c_lo = ms * P.chunks_per_split;
    c_hi = min(c_lo + P.chunks_per_split, total_chunks);
  };

  if (warp == 0) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, P.nq, 1, 1);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
        int h0, hw, q0, c_lo, c_hi;
        decode(item, h0, hw, q0, c_lo, c_hi);
        const int as = it & 1;
        mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int c = c_lo; c < c_hi; ++c) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          if (it == 0 && c == c_lo) stamp(3);
          const uint32_t st = smem_u32(base + (size_t)s * stage_bytes);
          const uint32_t sq = st + h_bytes * (MODE == 0 ? 2 : 1);
          const uint64_t dh = umma_desc_mnmajor_sw128(st, (uint32_t)MK * 128);
          const uint64_t dh2 = umma_desc_mnmajor_sw128(st + h_bytes, (uint32_t)MK * 128);   // MODE 0: y_raw tile
          const uint64_t dq = umma_desc_mnmajor_sw128(sq, (uint32_t)MK * 128);
          // 16 m-rows per MMA = 2048 B of 128-byte lines.  All MMAs of one accumulator back to back: alternating
          // between two accumulators costs about twice as much per instruction (measured in experiments/dw_tc.cu).
          for (int k = 0; k < (MK >> 4); ++k)
            umma_bf16(d_tmem, dh + (uint64_t)(128 * k), dq + (uint64_t)(128 * k), idesc, (c > c_lo) | (k > 0));
          if (MODE == 0 && !P.fuse_g)
            for (int k = 0; k < (MK >> 4); ++k)
              umma_bf16(d_tmem + 128, dh2 + (uint64_t)(128 * k), dq + (uint64_t)(128 * k), idesc, (c > c_lo) | (k > 0));
          umma_commit(&empty[s]);
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        umma_commit(&tmem_full[as]);
        if (it == 0) stamp(4);
      }
      stamp(8);
    }
  } else if (warp >= 4 && warp < 8) {
    // =========================== epilogue: TMEM -> red.global.add ===========================
    const int q = warp & 3;
    int it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      int h0, hw, q0, c_lo, c_hi;
      decode(item, h0, hw, q0, c_lo, c_hi);
      const int as = it & 1;
      mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      if (it == 0 && warp == 4 && lane == 0) stamp(5);
      const int hl = q * 32 + lane;                    // lane within the hidden chunk
      const bool ok = hl < hw && c_hi > c_lo;
      const int h = h0 + hl;
      const int hslot = h / P.hidden_true, ht = h - hslot * P.hidden_true;   // pixel slot / true channel of my lane
      const float inv_pack = 1.f / (float)P.pack;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * 256;
      float ka = 0.f, kb = 0.f, kc = 0.f;
      if (MODE == 0 && ok) {
        ka = s_coef[h]; kb = s_coef[P.hidden + h];
        kc = c_lo == 0 ? s_coef[2 * P.hidden + h] : 0.f;     // the cc * sum_m A term is added by the first m-split only
      }
      for (int c16 = 0; c16 < (P.nq >> 4); ++c16) {
        uint32_t r[16];
        tmem_ld_x16(taddr + c16 * 16, r);
        tmem_ld_wait();
        uint32_t r2[16];
        if (MODE == 0 && !P.fuse_g) {
          tmem_ld_x16(taddr + 128 + c16 * 16, r2);
          tmem_ld_wait();
        }
        const int qq = q0 + c16 * 16;
        const int qslot = qq / P.narrow_true, qt = qq - qslot * P.narrow_true;   // 16 columns never straddle a slot
        if (ok && qslot == hslot) {
          if (MODE == 0) {           // dW[n = hidden][k = narrow]: this lane's row, 16 consecutive k
            float* dst = P.dw + (size_t)ht * P.ldw + qt;
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)     // each of the `pack` diagonal blocks adds its share of cc * sum_m A
              v[j] = P.fuse_g ? __uint_as_float(r[j])
                              : fmaf(ka, __uint_as_float(r[j]),
                                     fmaf(kb, __uint_as_float(r2[j]), kc * inv_pack * __ldg(P.asum + qt + j)));
#pragma unroll
            for (int j = 0; j < 16; j += 4) red_add_v4(dst + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {                   // dW[n = narrow][k = hidden]: consecutive lanes -> consecutive k
#pragma unroll
            for (int j = 0; j < 16; ++j) atomicAdd(P.dw + (size_t)(qt + j) * P.ldw + ht, __uint_as_float(r[j]));
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
      if (it == 0 && warp == 4 && lane == 0) stamp(6);
    }
    if (warp == 4 && lane == 0) stamp(7);
  } else if (warp >= 8 && warp < 12) {
    // =========================== transform warps ===========================
    // MODE 1: rewrite the raw hidden tile in place with relu6(scale*x + shift).  MODE 0: nothing to compute (the
    // BatchNorm-backward affine is applied in the epilogue); the warps only publish the tile to the async proxy.
    // Pieces past the chunk's channels are zero (never written by the loaders) unless an earlier, wider item used
    // the stage: then they are re-zeroed once.
    const int tt = threadIdx.x - 256;            // 0..127
    const int pp = tt & 15, rb = tt >> 4;        // 16-byte piece pp (0..15 over the two 64-channel blocks) of rows rb + 8 i
    const int blk = pp >> 3, p = pp & 7;         // rows rb + 8 i all have (r & 7) == rb: the swizzle term is constant
    int s = 0;
    uint32_t ph = 0;
    int last_hw[MAX_STAGES];
#pragma unroll
    for (int i = 0; i < MAX_STAGES; ++i) last_hw[i] = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      int h0, hw, q0, c_lo, c_hi;
      decode(item, h0, hw, q0, c_lo, c_hi);
      const bool mine = pp * 8 < hw;
      f2 sc[4], sh[4];                           // packed fp32 pairs: scale, shift of my 8 channels
      if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          sc[j] = mine ? f2_make(s_coef[h0 + pp * 8 + 2 * j], s_coef[h0 + pp * 8 + 2 * j + 1]) : 0ull;
          sh[j] = mine ? f2_make(s_coef[P.hidden + h0 + pp * 8 + 2 * j], s_coef[P.hidden + h0 + pp * 8 + 2 * j + 1]) : 0ull;
        }
      }
      for (int c = c_lo; c < c_hi; ++c) {
        const int row0 = c * MK;
        mbar_wait(&raw_full[s], ph);
        const uint32_t st = smem_u32(base + (size_t)s * stage_bytes);
        const bool rezero = last_hw[s] > hw;
        if (MODE == 1 && mine) {
          // four rows per trip: loads, then math, then stores (the volatile shared accesses keep their order)
          for (int r = rb; r < MK; r += 32) {
            uint32_t off[4];
            uint4 xv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int ru = r + 8 * u;
              off[u] = blk * MK * 128 + ru * 128 + ((p ^ rb) << 4);
              if (ru < MK) xv[u] = lds16(st + off[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const bool ok = (row0 + r + 8 * u) < P.M;      // rows past M were zero-filled and must stay zero
              // packed fp32 FMA, one rounding, clamp on the packed bf16 pair
              xv[u].x = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(xv[u].x), sc[0], sh[0])));
              xv[u].y = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(xv[u].y), sc[1], sh[1])));
              xv[u].z = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(xv[u].z), sc[2], sh[2])));
              xv[u].w = bf16x2_relu6(f2_to_bf16x2(f2_fma(f2_from_bf16x2(xv[u].w), sc[3], sh[3])));
              if (!ok) xv[u] = make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (r + 8 * u < MK) sts16(st + off[u], xv[u]);
          }
        }
        if (MODE == 0 && P.fuse_g && mine) {
          // G = ca*dz + cb*y + cc into the dz tile (packed fp32, one rounding to bf16); rows past M stay zero
          f2 ka[4], kb[4], kc[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int hc = h0 + pp * 8 + 2 * j;
            ka[j] = f2_make(s_coef[hc], s_coef[hc + 1]);
            kb[j] = f2_make(s_coef[P.hidden + hc], s_coef[P.hidden + hc + 1]);
            kc[j] = f2_make(s_coef[2 * P.hidden + hc], s_coef[2 * P.hidden + hc + 1]);
          }
          for (int r = rb; r < MK; r += 16) {
            uint32_t off[2];
            uint4 zv[2], yv[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              off[u] = blk * MK * 128 + (r + 8 * u) * 128 + ((p ^ rb) << 4);
              zv[u] = lds16(st + off[u]);
              yv[u] = lds16(st + h_bytes + off[u]);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const bool ok = (row0 + r + 8 * u) < P.M;
              zv[u].x = f2_to_bf16x2(f2_fma(ka[0], f2_from_bf16x2(zv[u].x), f2_fma(kb[0], f2_from_bf16x2(yv[u].x), kc[0])));
              zv[u].y = f2_to_bf16x2(f2_fma(ka[1], f2_from_bf16x2(zv[u].y), f2_fma(kb[1], f2_from_bf16x2(yv[u].y), kc[1])));
              zv[u].z = f2_to_bf16x2(f2_fma(ka[2], f2_from_bf16x2(zv[u].z), f2_fma(kb[2], f2_from_bf16x2(yv[u].z), kc[2])));
              zv[u].w = f2_to_bf16x2(f2_fma(ka[3], f2_from_bf16x2(zv[u].w), f2_fma(kb[3], f2_from_bf16x2(yv[u].w), kc[3])));
              if (!ok) zv[u] = make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) sts16(st + off[u], zv[u]);
          }
        }
        if (rezero && !mine) {
          for (int r = rb; r < MK; r += 8) {
            const uint32_t off = blk * MK * 128 + r * 128 + ((p ^ rb) << 4);
            sts16(st + off, make_uint4(0u, 0u, 0u, 0u));
            if (MODE == 0) sts16(st + h_bytes + off, make_uint4(0u, 0u, 0u, 0u));
          }
        }
        last_hw[s] = hw;
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp >= 12) {
    // =========================== loader warps: cp.async straight into the swizzled operand layout ===========================
    const int lt = threadIdx.x - 384;            // 0..127
    const int pp = lt & 15, rb = lt >> 4;        // hidden tile: piece pp of rows rb + 8 i   ((r & 7) == rb)
    const int blk = pp >> 3, p = pp & 7;
    const int pq = lt & 31, rq = lt >> 5;        // narrow tile: piece pq of rows rq + 4 i
    const bool qmine = pq * 8 < P.nq;
    int s = 0;
    uint32_t ph = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      int h0, hw, q0, c_lo, c_hi;
      decode(item, h0, hw, q0, c_lo, c_hi);
      const bool mine = pp * 8 < hw;
      const uint32_t hoff = blk * MK * 128 + rb * 128 + ((p ^ rb) << 4);
      for (int c = c_lo; c < c_hi; ++c) {
        const int row0 = c * MK;
        mbar_wait(&empty[s], ph ^ 1);
        const uint32_t st = smem_u32(base + (size_t)s * stage_bytes);
        const uint32_t sq = st + h_bytes * (MODE == 0 ? 2 : 1);
        if (mine) {
          const __nv_bfloat16* g0 = P.t0 + (size_t)(row0 + rb) * P.ldh + h0 + pp * 8;
          const __nv_bfloat16* g1 = MODE == 0 ? P.t1 + (size_t)(row0 + rb) * P.ldh + h0 + pp * 8 : nullptr;
          uint32_t d = st + hoff;
          for (int r = rb; r < MK; r += 8) {
            const bool ok = (row0 + r) < P.M;
            cp_async16(d, ok ? g0 : P.t0, ok ? 16u : 0u);
            if (MODE == 0) cp_async16(d + h_bytes, ok ? g1 : P.t1, ok ? 16u : 0u);
            d += 8 * 128;
            g0 += (size_t)8 * P.ldh;
            if (MODE == 0) g1 += (size_t)8 * P.ldh;
          }
        }
        if (qmine) {
          const __nv_bfloat16* g = P.qsrc + (size_t)(row0 + rq) * P.ldq + q0 + pq * 8;
          const uint32_t qb = sq + (pq >> 3) * MK * 128;
          for (int r = rq; r < MK; r += 4) {
            const bool ok = (row0 + r) < P.M;
            cp_async16(qb + r * 128 + (((pq & 7) ^ (r & 7)) << 4), ok ? g : P.qsrc, ok ? 16u : 0u);
            g += (size_t)4 * P.ldq;
          }
        }
        cp_async_arrive(&raw_full[s]);            // arrives when this thread's copies above have landed
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
    if (lane == 0) stamp(9);
  }
}

}  // namespace

extern unsigned long long* g_pw_trace;

// mode 0: hidden side = G = ca*t0 + cb*t1 + cc with t0 = dz, t1 = y_raw [M][N];  narrow side = `pure` = A [M][K]
// mode 1: hidden side = A = relu6(scale*t0 + shift) with t0 = d_raw [M][K];      narrow side = `pure` = G [M][N]
// Row packing for layers whose WIDE side is narrow too (hidden <= 64): r = 128 / hidden consecutive pixels, at most 4.
// One MMA then reduces over 16 r pixels instead of 16 -- these layers are bound by the per-instruction cost of
// tcgen05.mma (~120 ns each whatever the shape, scripts/trace_pw.py), not by bandwidth.
int pw_wgrad_pick_pack(int mode, int M, int N, int K) {
  const int hidden = mode == 0 ? N : K, narrow = mode == 0 ? K : N;
  for (int r = 4; r > 1; r >>= 1)
    if (hidden * r <= 128 && narrow * r <= 128 && M % r == 0) return r;
  return 1;
}

int pw_wgrad_launch(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0, const void* t1,
                    int ld_t, const float* coef, const float* asum, float* dw, int ldw, int max_ctas, cudaStream_t s,
                    int pack, const BnBwdSrc* src) {
  B200DP_REQUIRE(mode == 0 || mode == 1, "pw_wgrad: bad mode");
  B200DP_REQUIRE(M > 0 && N % 16 == 0 && K % 16 == 0, "pw_wgrad: M=%d N=%d K=%d invalid", M, N, K);
  B200DP_REQUIRE(pure && t0 && coef && dw && (mode == 1 || (t1 && asum)), "pw_wgrad: missing pointers");
  B200DP_REQUIRE(pack >= 1 && M % pack == 0, "pw_wgrad: pack=%d does not divide M=%d", pack, M);
  WgradParams p{};
  p.hidden_true = mode == 0 ? N : K;
  p.narrow_true = mode == 0 ? K : N;
  p.pack = pack;
  if (pack > 1) {
    B200DP_REQUIRE(ld_t == p.hidden_true && ld_pure == p.narrow_true, "pw_wgrad: row packing needs dense tensors");
    M /= pack; ld_t *= pack; ld_pure *= pack;
  }
  p.M = M;
  p.hidden = p.hidden_true * pack;
  p.narrow = p.narrow_true * pack;
  p.ldh = ld_t; p.ldq = ld_pure; p.ldw = ldw;
  p.coef = coef; p.dw = dw;
  p.hchunks = (p.hidden + 127) / 128;
  {
    // measured per layer (profile_step, B200DP_WGRAD_FUSEG=0/1): the transform warps are slower than the second MMA
    // everywhere except where the wider accumulator saves passes over the hidden tensors (last conv: 47 -> 36 us)
    const char* fg = getenv("B200DP_WGRAD_FUSEG");
    const bool want = fg != nullptr ? atoi(fg) != 0 : p.narrow > 160;
    p.fuse_g = (mode == 0 && want) ? 1 : 0;
  }
  const int qmax = (mode == 0 && !p.fuse_g) ? 128 : 256;   // the factored MODE 0 keeps two accumulators per stage
  p.qchunks = (p.narrow + qmax - 1) / qmax;
  while (p.narrow % p.qchunks != 0 || (p.narrow / p.qchunks) % 16 != 0) ++p.qchunks;
  p.nq = p.narrow / p.qchunks;
  p.qblks = (p.nq + 63) / 64;
  const int row_bytes = 256 * (mode == 0 ? 2 : 1) + p.qblks * 128;
  p.mk = row_bytes <= 384 ? 128 : 64;
  const size_t stage_bytes = (size_t)p.mk * row_bytes;
  const size_t fixed = ((size_t)(mode == 0 ? 3 : 2) * p.hidden + 4) * 4 + (3 * MAX_STAGES + 4) * 8 + 16 + 1024;
  const size_t budget = 220 * 1024;
  int stages = (int)((budget - fixed) / stage_bytes);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  B200DP_REQUIRE(stages >= 2, "pw_wgrad: shared memory budget exceeded");
  p.stages = stages;
  const size_t smem = fixed + stages * stage_bytes;
  const int total_chunks = (M + p.mk - 1) / p.mk;
  const int cap = max_ctas > 0 ? max_ctas : num_sms();
  const int hq = p.hchunks * p.qchunks;
  int msplit = (2 * cap + hq - 1) / hq;
  const int max_split = (total_chunks + 3) / 4;        // at least ~4 stages of work per item
  if (msplit > max_split) msplit = max_split;
  if (msplit < 1) msplit = 1;
  p.chunks_per_split = (total_chunks + msplit - 1) / msplit;
  p.msplit = (total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  p.t0 = static_cast<const __nv_bfloat16*>(t0);
  p.t1 = static_cast<const __nv_bfloat16*>(t1);
  p.qsrc = static_cast<const __nv_bfloat16*>(pure);
  p.asum = asum;
  p.trace = g_pw_trace;
  const int items = hq * p.msplit;
  const int grid = items < cap ? items : cap;
  if (mode == 0) {
    B200DP_CUDA(cudaFuncSetAttribute(pw_wgrad_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200DP_CUDA(launch_k(pw_wgrad_kernel<0>, dim3(grid), dim3(512), smem, s, p, src ? *src : BnBwdSrc{}));
  } else {
    B200DP_CUDA(cudaFuncSetAttribute(pw_wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200DP_CUDA(launch_k(pw_wgrad_kernel<1>, dim3(grid), dim3(512), smem, s, p, BnBwdSrc{}));
  }
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_pw_wgrad(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0,
                               const void* t1, int ld_t, const float* coef, const float* asum, float* dw, int ldw,
                               int max_ctas, void* stream) {
  return pw_wgrad_launch(mode, M, N, K, pure, ld_pure, t0, t1, ld_t, coef, asum, dw, ldw, max_ctas,
                         static_cast<cudaStream_t>(stream), 1, nullptr);
}
// Same with row packing (see pw_wgrad_pick_pack); pack = 0 lets the library choose.
extern "C" int b200dp_pw_wgrad_packed(int mode, int M, int N, int K, const void* pure, const void* t0, const void* t1,
                                      const float* coef, const float* asum, float* dw, int max_ctas, int pack,
                                      void* stream) {
  if (pack <= 0) pack = pw_wgrad_pick_pack(mode, M, N, K);
  return pw_wgrad_launch(mode, M, N, K, pure, mode == 0 ? K : N, t0, t1, mode == 0 ? N : K, coef, asum, dw, K, max_ctas,
                         static_cast<cudaStream_t>(stream), pack, nullptr);
}
