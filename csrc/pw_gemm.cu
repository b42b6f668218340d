// Pointwise (1x1) convolution as a persistent, warp-specialised tcgen05 GEMM.
//
//   Out[M, N] (bf16) = Aeff[M, K] (bf16) * W[N, K]^T,  fp32 accumulation in TMEM.
//
// M = batch*H*W rows of an NHWC activation, K / N = input / output channels.  The same kernel
// serves the forward convolutions and the data gradients (weights pre-transposed), selected by
//   AMODE  how the A operand reaches shared memory
//     A_TMA     TMA loads the bf16 tile as is                      (expand conv fwd, project dgrad)
//     A_BNRELU  producer warps load the *raw* output of the previous conv, apply its BatchNorm
//               scale/shift (+ReLU6) in registers, store the swizzled tile   (project conv fwd)
//     A_BNGRAD  producer warps combine (dz, x) into the BatchNorm input gradient
//               ca*dz + cb*x + cc in registers                                (expand conv dgrad)
//   EMODE  what the epilogue does besides the bf16 store
//     E_PLAIN      nothing
//     E_STATS      per-channel sum and sum of squares (the training-mode BatchNorm statistics of THIS
//                  conv's output -- no separate batch-norm pass ever reads the tensor)
//     E_MASKSTATS  ReLU6 backward mask recomputed from the raw forward activation, plus the two
//                  per-channel sums BatchNorm backward needs (sum dz, sum dz*x)
//
// Reference behaviour replaced: cuDNN 1x1 convolution + cuDNN batch-norm + ATen hardtanh kernels,
// each a full read+write of the activation (torchvision mobilenetv2.py via
// cifar10_mpi_mobilenet_224.py:177, :179).
//
// Roles (one CTA per SM, static round-robin over 128 x BLOCK_N tiles):
//   warp 0      TMA producer: weight tile (and the A tile in A_TMA mode) per 64-wide K chunk
//   warp 1      MMA issuer: one elected thread issues tcgen05.mma (M=128, N=BLOCK_N, K=16)
//   warp 2      TMEM allocator (2 accumulator stages x 128 columns)
//   warps 4-7   epilogue: tcgen05.ld -> bf16 -> swizzled smem -> coalesced pass (stats / mask) -> TMA store
//   warps 8-15  A producers (only in A_BNRELU / A_BNGRAD)
// Pipelines: smem full/empty mbarriers per stage, TMEM full/empty per accumulator stage.
#include "common.cuh"
#include "pw_gemm.h"
#include "sm100.cuh"

namespace b200dp {

namespace {

using namespace sm100;

constexpr int BM = 128;
constexpr int KCH = 64;
constexpr int A_STAGE = BM * 128;       // bytes
constexpr int OUT_CHUNK = BM * 128;     // one 64-column block of the output tile, bytes
constexpr int ACC_COLS = 128;           // TMEM columns per accumulator stage
constexpr int TMEM_COLS = 256;
constexpr int EPI_THREADS = 128;
constexpr int PROD_THREADS = 256;        // transform warps: 8 (one warp per scheduler cannot hide its own ALU latency)
constexpr int XROWS = PROD_THREADS / 8;  // rows covered by one sweep of the transform threads
constexpr int MAX_STAGES = 8;

struct Smem {
  uint8_t* stage_a[MAX_STAGES];
  uint8_t* stage_b[MAX_STAGES];
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                    pack_bf16x2(f[6], f[7]));
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

// Thread layout: 4 control warps + 2 epilogue groups of 4 warps; each group owns one TMEM accumulator stage, one
// staging buffer and (E_MASKSTATS) one aux buffer, so two tiles drain concurrently.  A_BNRELU / A_BNGRAD add 4
// transform warps: ALL global->shared traffic is TMA (deeply pipelined, no registers held across the latency); the
// transform warps rewrite the landed raw tile in place (shared -> registers -> BatchNorm math -> shared).
template <int AMODE>
struct Layout {
  static constexpr int kEpiGroups = 2;
  static constexpr int kProdWarp0 = 4 + 4 * kEpiGroups;
  static constexpr int kThreads = AMODE == A_TMA ? 384 : 384 + PROD_THREADS;
};

template <int AMODE, int EMODE>
__global__ void __launch_bounds__(Layout<AMODE>::kThreads, 1)
pw_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmAt, const __grid_constant__ CUtensorMap tmBt,
               const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmA2t,
               const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmX,
               const __grid_constant__ PwGemmParams P, const __grid_constant__ BnFwdSrc asrc,
               const __grid_constant__ BnBwdSrc gsrc) {
  using L = Layout<AMODE>;
  constexpr int NG = L::kEpiGroups;
  auto stamp = [&](int slot) {
    if (P.trace != nullptr && blockIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      P.trace[slot] = t;
    }
  };
  if (threadIdx.x == 0) stamp(0);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // ---- carve shared memory (tile buffers 1024-byte aligned for the 128-byte swizzle) ------------
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = P.stages;
  const int b_stage = P.block_n * 128;
  const int nch = (P.block_n + 63) >> 6;
  uint8_t* stage_a = base;
  uint8_t* stage_a2 = stage_a + (size_t)stages * A_STAGE;                      // A_BNGRAD: raw x tile next to raw dz
  uint8_t* stage_b = stage_a2 + (AMODE == A_BNGRAD ? (size_t)stages * A_STAGE : 0);
  uint8_t* out_buf = stage_b + (size_t)stages * b_stage;                       // 2 x nch x 16 KB
  uint8_t* aux_buf = out_buf + 2 * (size_t)nch * OUT_CHUNK;                    // 2 x nch x 16 KB (E_MASKSTATS)
  float* s_acoef = reinterpret_cast<float*>(aux_buf + (EMODE == E_MASKSTATS ? 2 * (size_t)nch * OUT_CHUNK : 0));
  const int n_acoef = (AMODE == A_BNRELU ? 2 : (AMODE == A_BNGRAD ? 3 : 0)) * P.K;
  float* s_ecoef = s_acoef + n_acoef;
  const int n_ecoef = (EMODE == E_MASKSTATS ? 2 : 0) * P.N;
  float* s_stats = s_ecoef + n_ecoef;                                          // [8 epilogue warps][2][128]
  const int n_stats = EMODE != E_PLAIN ? 4 * NG * 2 * 128 : 0;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_stats + ((n_stats + 1) & ~1));
  uint64_t* full = bars;
  uint64_t* empty = bars + MAX_STAGES;
  uint64_t* tmem_full = bars + 2 * MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* aux_full = tmem_empty + 2;
  uint64_t* aux_empty = aux_full + 2;
  uint64_t* raw_full = aux_empty + 2;                                          // [MAX_STAGES] raw A tile(s) landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(raw_full + MAX_STAGES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // A CTA keeps ONE block of output columns for its whole life (b % n_tiles) and walks the row tiles with stride
  // gridDim / n_tiles: the per-column sums stay in registers until the end, and CTAs b, b+1, ... work on the same
  // rows at the same time (the A tile is fetched from HBM once and hit in L2 by the others).
  const int n0 = ((int)blockIdx.x % P.n_tiles) * P.block_n;
  const int mt0 = (int)blockIdx.x / P.n_tiles, mt_step = (int)gridDim.x / P.n_tiles;

  // ---- one-time set-up ---------------------------------------------------------------------------
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmBt);
    prefetch_tmap(&tmO);
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmAt);
    if (AMODE == A_BNGRAD) { prefetch_tmap(&tmA2); prefetch_tmap(&tmA2t); }
    if (EMODE == E_MASKSTATS) prefetch_tmap(&tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], AMODE == A_TMA ? 1 : 1 + PROD_THREADS / 32);   // TMA (weights) + one arrival per transform warp
      mbar_init(&empty[s], 1);
      mbar_init(&raw_full[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], EPI_THREADS);
      mbar_init(&aux_full[a], 1);
      mbar_init(&aux_empty[a], EPI_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
  // everything above is on-chip set-up and overlaps the previous kernel's tail (PDL); global data from here on
  pdl_trigger();
  pdl_wait();
  // coefficient arrays have the unpacked lengths; packed column c belongs to channel c % true width
  if (AMODE == A_BNRELU && asrc.sums != nullptr) {
    // the A-side BatchNorm is finalised here (this GEMM is its first consumer); CTA 0 publishes the coefficients
    for (int c = threadIdx.x; c < P.K; c += blockDim.x) {
      float sc, sh;
      bn_fwd_channel(asrc, c % P.k_true, blockIdx.x == 0 && c < P.k_true, sc, sh);
      s_acoef[c] = sc;
      s_acoef[P.K + c] = sh;
    }
  } else if (AMODE == A_BNGRAD && gsrc.sums != nullptr) {
    // BatchNorm-backward coefficients straight from the producer's sums, with NO side effects: the weight-gradient
    // kernel of the same layer runs concurrently on another graph branch and is the designated writer
    for (int c = threadIdx.x; c < P.K; c += blockDim.x) {
      float ca, cb, cc;
      bn_bwd_channel(gsrc, c % P.k_true, false, ca, cb, cc);
      s_acoef[c] = ca;
      s_acoef[P.K + c] = cb;
      s_acoef[2 * P.K + c] = cc;
    }
  } else {
    for (int i = threadIdx.x; i < n_acoef; i += blockDim.x) {
      const int j = i / P.K, c = i - j * P.K;
      s_acoef[i] = P.acoef[j * P.k_true + c % P.k_true];
    }
  }
  for (int i = threadIdx.x; i < n_ecoef; i += blockDim.x) {
    const int j = i / P.N, c = i - j * P.N;
    s_ecoef[i] = P.ecoef[j * P.n_true + c % P.n_true];
  }
  for (int i = threadIdx.x; i < n_stats; i += blockDim.x) s_stats[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) stamp(1);

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int mt = mt0; mt < P.m_tiles; mt += mt_step, ++it) {
        const int m0 = mt * BM;
        if (EMODE == E_MASKSTATS) {     // raw forward activation tile for the ReLU6 mask, read by the epilogue
          const int ab = it & 1;
          mbar_wait(&aux_empty[ab], ((it >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&aux_full[ab], (uint32_t)nch * OUT_CHUNK);
          for (int c = 0; c < nch; ++c)
            tma_load_2d(aux_buf + ((size_t)ab * nch + c) * OUT_CHUNK, &tmX, &aux_full[ab], n0 + 64 * c, m0);
        }
        for (int kc = 0; kc < P.kc; ++kc) {
          // the last K chunk may be narrower than 64: its box is exactly that wide (a box that hangs over the
          // tensor edge makes TMA crawl), and lands in the matching 32B / 64B swizzle layout
          const bool last = kc == P.kc - 1;
          const int w = last ? P.tail_w : KCH;
          mbar_wait(&empty[s], ph ^ 1);
          if (AMODE == A_TMA) {
            mbar_arrive_expect_tx(&full[s], (uint32_t)(BM + P.block_n) * w * 2);
            tma_load_2d(stage_a + (size_t)s * A_STAGE, last ? &tmAt : &tmA, &full[s], kc * KCH, m0);
          } else {
            // raw tile(s) -> raw_full (consumed by the transform warps); weights -> full (consumed by the MMA)
            mbar_arrive_expect_tx(&raw_full[s], (uint32_t)BM * w * 2 * (AMODE == A_BNGRAD ? 2 : 1));
            tma_load_2d(stage_a + (size_t)s * A_STAGE, last ? &tmAt : &tmA, &raw_full[s], kc * KCH, m0);
            if (AMODE == A_BNGRAD)
              tma_load_2d(stage_a2 + (size_t)s * A_STAGE, last ? &tmA2t : &tmA2, &raw_full[s], kc * KCH, m0);
            mbar_arrive_expect_tx(&full[s], (uint32_t)P.block_n * w * 2);
          }
          tma_load_2d(stage_b + (size_t)s * b_stage, last ? &tmBt : &tmB, &full[s], kc * KCH, n0);
          if (it == 0 && kc == 0) stamp(2);
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(BM, P.block_n, 0, 0);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int mt = mt0; mt < P.m_tiles; mt += mt_step, ++it) {
        const int as = it & 1;
        mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * ACC_COLS;
        for (int kc = 0; kc < P.kc; ++kc) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          if (it == 0 && kc == 0) stamp(3);
          if (it == 0 && kc == 4) stamp(14);      // steady-state pace of the K loop: (stamp 15 - stamp 14) / 8 chunks
          if (it == 0 && kc == 12) stamp(15);
          const int w = kc == P.kc - 1 ? P.tail_w : KCH;
          const uint64_t da = umma_desc_kmajor(smem_u32(stage_a + (size_t)s * A_STAGE), w);
          const uint64_t db = umma_desc_kmajor(smem_u32(stage_b + (size_t)s * b_stage), w);
          const int ksteps = w >> 4;
          for (int k = 0; k < ksteps; ++k)
            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kc | k) != 0);
          umma_commit(&empty[s]);   // smem slot reusable once these MMAs have read it
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        umma_commit(&tmem_full[as]);  // accumulator complete
        if (it == 0) stamp(4);
      }
    }
  } else if (warp >= 4 && warp < L::kProdWarp0) {
    // =========================== epilogue (warp-local: no block-level barrier per tile) ===========================
    // Warp q of a group owns TMEM lanes / tile rows 32q..32q+31 end to end: TMEM -> bf16 -> swizzled staging rows
    // -> coalesced sweep over the same rows (statistics / mask) -> its own TMA store of a {64, 32} box.
    const int eg = (warp - 4) >> 2;            // epilogue group
    const int q = warp & 3;                    // TMEM lane quadrant this warp may read
    const int row0 = q * 32;
    const int row = row0 + lane;               // pass 1: one row per thread
    const int p = lane & 7, rr = lane >> 3;    // pass 2: 16-byte piece p of rows row0 + rr + 4 i
    float s0[2][8], s1[2][8];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) s0[c][j] = s1[c][j] = 0.f;
    float esc[2][8], esh[2][8];

    // my 32 rows' per-column sums -> my own slot of shared memory (plain stores: shared-memory atomics cost ~280
    // cycles apiece here and made this a 4.6 us stall per column-block change, see scripts/trace_pw.py); the slots
    // are added up and sent to global memory after the main loop
    auto flush = [&]() {
      if (EMODE == E_PLAIN) return;
      float* slot = s_stats + (size_t)(warp - 4) * 256;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float a[8], b[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a[j] = s0[c][j]; b[j] = s1[c][j];
          a[j] += __shfl_xor_sync(0xffffffffu, a[j], 8);
          a[j] += __shfl_xor_sync(0xffffffffu, a[j], 16);
          b[j] += __shfl_xor_sync(0xffffffffu, b[j], 8);
          b[j] += __shfl_xor_sync(0xffffffffu, b[j], 16);
        }
        if (lane < 8) {
          float4* d0 = reinterpret_cast<float4*>(slot + c * 64 + lane * 8);
          float4* d1 = reinterpret_cast<float4*>(slot + 128 + c * 64 + lane * 8);
          d0[0] = make_float4(a[0], a[1], a[2], a[3]); d0[1] = make_float4(a[4], a[5], a[6], a[7]);
          d1[0] = make_float4(b[0], b[1], b[2], b[3]); d1[1] = make_float4(b[4], b[5], b[6], b[7]);
        }
      }
    };
    if (EMODE == E_MASKSTATS) {
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int col = n0 + c * 64 + p * 8 + j;
          esc[c][j] = col < P.N ? s_ecoef[col] : 0.f;
          esh[c][j] = col < P.N ? s_ecoef[P.N + col] : 0.f;
        }
    }

    for (int it = eg, mt = mt0 + eg * mt_step; mt < P.m_tiles; it += NG, mt += NG * mt_step) {
      const int m0 = mt * BM;
      const int as = it & 1;
      uint8_t* ob = out_buf + (size_t)as * nch * OUT_CHUNK;
      const uint32_t ob_u32 = smem_u32(ob);
      mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      if (it == 0 && warp == 4 && lane == 0) stamp(5);
      // this warp's previous TMA store out of the same staging rows must have finished reading them
      if (lane == 0) {
        if (NG == 2) tma_store_wait_read<0>(); else tma_store_wait_read<1>();
      }
      __syncwarp();
      // ---- pass 1: TMEM -> registers -> bf16 -> swizzled staging tile (one row per thread)
      const uint32_t taddr = tmem_base + ((uint32_t)row0 << 16) + as * ACC_COLS;
      // two 16-column loads in flight per wait: the TMEM round trip, not the conversion, paces this loop
      auto stage16 = [&](int c16, const uint32_t (&r)[16]) {
        const uint4 lo = make_uint4(pack_bf16x2(__uint_as_float(r[0]), __uint_as_float(r[1])),
                                    pack_bf16x2(__uint_as_float(r[2]), __uint_as_float(r[3])),
                                    pack_bf16x2(__uint_as_float(r[4]), __uint_as_float(r[5])),
                                    pack_bf16x2(__uint_as_float(r[6]), __uint_as_float(r[7])));
        const uint4 hi = make_uint4(pack_bf16x2(__uint_as_float(r[8]), __uint_as_float(r[9])),
                                    pack_bf16x2(__uint_as_float(r[10]), __uint_as_float(r[11])),
                                    pack_bf16x2(__uint_as_float(r[12]), __uint_as_float(r[13])),
                                    pack_bf16x2(__uint_as_float(r[14]), __uint_as_float(r[15])));
        const uint32_t rowaddr = ob_u32 + (c16 >> 2) * OUT_CHUNK + row * 128;
        const int pc = (c16 & 3) * 2;
        sts16(rowaddr + (((pc) ^ (row & 7)) << 4), lo);
        sts16(rowaddr + (((pc + 1) ^ (row & 7)) << 4), hi);
      };
      const int n16 = P.block_n >> 4;
      for (int c16 = 0; c16 < n16; c16 += 2) {
        uint32_t ra[16], rb[16];
        tmem_ld_x16(taddr + c16 * 16, ra);
        const bool two = c16 + 1 < n16;
        if (two) tmem_ld_x16(taddr + (c16 + 1) * 16, rb);
        tmem_ld_wait();
        stage16(c16, ra);
        if (two) stage16(c16 + 1, rb);
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);   // accumulator stage free for the MMA warp
      __syncwarp();
      // ---- pass 2: coalesced sweep over this warp's 32 staged rows: statistics (and ReLU6 mask)
      if (EMODE != E_PLAIN) {
        uint32_t ax_u32 = 0;
        if (EMODE == E_MASKSTATS) {
          mbar_wait(&aux_full[as], (it >> 1) & 1);
          ax_u32 = smem_u32(aux_buf + (size_t)as * nch * OUT_CHUNK);
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c < nch) {
            const int col0 = n0 + c * 64 + p * 8;
            const bool colok = col0 < P.N && (c * 64 + p * 8) < P.block_n;
            if (!colok) continue;   // staging columns past the tile hold no data
            if (EMODE == E_STATS) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int r = row0 + rr + 4 * i;
                const uint32_t off = c * OUT_CHUNK + r * 128 + ((p ^ (r & 7)) << 4);
                float x[8];
                unpack8(lds16(ob_u32 + off), x);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  s0[c][j] += x[j];
                  s1[c][j] = fmaf(x[j], x[j], s1[c][j]);
                }
              }
            } else {
              // two rows per trip, loads before math before stores (volatile shared accesses keep their order)
#pragma unroll
              for (int i = 0; i < 8; i += 2) {
                uint32_t off[2];
                uint4 xv[2], av[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                  const int r = row0 + rr + 4 * (i + u);
                  off[u] = c * OUT_CHUNK + r * 128 + ((p ^ (r & 7)) << 4);
                  xv[u] = lds16(ob_u32 + off[u]);
                  av[u] = lds16(ax_u32 + off[u]);      // rows past M were zero-filled by TMA -> masked to 0
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                  float x[8], a[8];
                  unpack8(xv[u], x);
                  unpack8(av[u], a);
#pragma unroll
                  for (int j = 0; j < 8; ++j) {
                    const float z = fmaf(a[j], esc[c][j], esh[c][j]);
                    const float o = (z > 0.f && z < 6.f) ? x[j] : 0.f;
                    x[j] = o;
                    s0[c][j] += o;
                    s1[c][j] = fmaf(o, a[j], s1[c][j]);
                  }
                  xv[u] = pack8(x);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) sts16(ob_u32 + off[u], xv[u]);
              }
            }
          }
        }
        if (EMODE == E_MASKSTATS) mbar_arrive(&aux_empty[as]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        for (int c = 0; c < nch; ++c)
          tma_store_2d(&tmO, ob + (size_t)c * OUT_CHUNK + row0 * 128, n0 + c * 64, m0 + row0);
        tma_store_commit();
        if (it == 0 && warp == 4) stamp(6);
      }
    }
    if (warp == 4 && lane == 0) stamp(10);
    flush();
    if (warp == 4 && lane == 0) stamp(11);
    if (lane == 0) tma_store_wait_all<0>();
    if (warp == 4 && lane == 0) stamp(12);
  } else if (AMODE != A_TMA && warp >= L::kProdWarp0) {
    // =========================== transform warps: raw tile -> operand tile, in place ===========================
    const int pt = threadIdx.x - L::kProdWarp0 * 32;   // 0..127
    const int p = pt & 7, rb = pt >> 3;                // 16-byte piece p of rows rb + XROWS i
    int s = 0;
    uint32_t ph = 0;
    for (int mt = mt0; mt < P.m_tiles; mt += mt_step) {
      const int m0 = mt * BM;
      for (int kc = 0; kc < P.kc; ++kc) {
        const int w = kc == P.kc - 1 ? P.tail_w : KCH;   // chunk width: row pitch 2w bytes, w/8 pieces per row
        const int wsh = w == 64 ? 0 : (w == 32 ? 1 : 2);
        const int col = kc * KCH + p * 8;
        const bool colok = p * 8 < w;
        // my piece's coefficients as packed fp32 pairs (three 16-byte loads per array instead of 24 scalar ones)
        f2 k0[4], k1[4], k2[4];
        if (colok) {
          const uint32_t ca = smem_u32(s_acoef + col);
          const uint4 a0 = lds16(ca), a1 = lds16(ca + 16);
          const uint4 b0 = lds16(ca + 4 * P.K), b1 = lds16(ca + 4 * P.K + 16);
          k0[0] = f2_pair(a0.x, a0.y); k0[1] = f2_pair(a0.z, a0.w); k0[2] = f2_pair(a1.x, a1.y); k0[3] = f2_pair(a1.z, a1.w);
          k1[0] = f2_pair(b0.x, b0.y); k1[1] = f2_pair(b0.z, b0.w); k1[2] = f2_pair(b1.x, b1.y); k1[3] = f2_pair(b1.z, b1.w);
          if (AMODE == A_BNGRAD) {
            const uint4 c0 = lds16(ca + 8 * P.K), c1 = lds16(ca + 8 * P.K + 16);
            k2[0] = f2_pair(c0.x, c0.y); k2[1] = f2_pair(c0.z, c0.w); k2[2] = f2_pair(c1.x, c1.y); k2[3] = f2_pair(c1.z, c1.w);
          }
        }
        mbar_wait(&raw_full[s], ph);
        if (pt == 0 && mt == mt0 && kc == 4) stamp(16);     // transform warps: raw chunk 4 / 12 available
        if (pt == 0 && mt == mt0 && kc == 12) stamp(17);
        const uint32_t sa = smem_u32(stage_a + (size_t)s * A_STAGE);
        const uint32_t sx = smem_u32(stage_a2 + (size_t)s * A_STAGE);
        if (colok) {
          const bool full_tile = m0 + BM <= P.M;       // rows past M were zero-filled by TMA and must stay zero
          // all loads of a batch first, then the math, then the stores: the shared-memory round trip (not the ~30
          // instructions per piece) is what one piece costs, and the volatile asm accesses do not reorder themselves
          constexpr int NPIECE = BM / XROWS;                 // pieces per thread per chunk
          constexpr int NB = NPIECE;                         // all of them in one batch
#pragma unroll
          for (int b0 = 0; b0 < NPIECE; b0 += NB) {
            uint4 xv[NB], gv[NB];
            uint32_t off[NB];
#pragma unroll
            for (int i = 0; i < NB; ++i) {
              const int r = rb + XROWS * (b0 + i);
              off[i] = r * 2 * w + ((p ^ ((r >> wsh) & ((8 >> wsh) - 1))) << 4);
              xv[i] = lds16(sa + off[i]);
              if (AMODE == A_BNGRAD) gv[i] = lds16(sx + off[i]);          // sa = dz, sx = raw activation
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) {
              const uint32_t xw[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
              uint32_t yw[4];
              if (AMODE == A_BNRELU) {
                // packed fp32 FMA, one rounding to bf16, clamp on the packed pair (0 and 6 are exact in bf16)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const uint32_t z = f2_to_bf16x2(f2_fma(f2_from_bf16x2(xw[q]), k0[q], k1[q]));
                  yw[q] = P.a_relu6 ? bf16x2_relu6(z) : z;
                }
              } else {
                const uint32_t gw[4] = {gv[i].x, gv[i].y, gv[i].z, gv[i].w};
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  yw[q] = f2_to_bf16x2(f2_fma(k0[q], f2_from_bf16x2(xw[q]), f2_fma(k1[q], f2_from_bf16x2(gw[q]), k2[q])));
              }
              if (!full_tile && (m0 + rb + XROWS * (b0 + i)) >= P.M) yw[0] = yw[1] = yw[2] = yw[3] = 0u;
              xv[i] = make_uint4(yw[0], yw[1], yw[2], yw[3]);
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) sts16(sa + off[i], xv[i]);
          }
        }
        // one arrival per warp: 128 per-thread arrivals on one mbarrier serialise in shared memory, once per K chunk
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    }
  }

  // ---- teardown: per-channel sums shared -> global, TMEM release ------------------------------------
  if (threadIdx.x == 0) stamp(7);
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) stamp(8);
  if (EMODE != E_PLAIN) {
    for (int i = threadIdx.x; i < 2 * P.block_n; i += blockDim.x) {
      const int h = i / P.block_n, col = i - h * P.block_n;
      if (n0 + col < P.N) {
        float a = 0.f;
#pragma unroll
        for (int wq = 0; wq < 4 * NG; ++wq) a += s_stats[wq * 256 + h * 128 + col];
        atomicAdd(P.stats + (size_t)h * P.stats_stride + (n0 + col) % P.n_true, a);
      }
    }
  }
  if (warp == 2) {
    if (lane == 0) stamp(13);
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
    if (lane == 0) stamp(9);
  }
}

template <int AMODE, int EMODE>
int launch_t(const PwGemmPlan& pl, cudaStream_t s) {
  auto kern = pw_gemm_kernel<AMODE, EMODE>;
  // always the architectural maximum: the attribute is function-global state, and CUDA-graph replays must not
  // depend on which layer set it last
  B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  B200DP_CUDA(launch_k(kern, dim3(pl.grid), dim3(Layout<AMODE>::kThreads), pl.smem_bytes, s, pl.tmA, pl.tmB, pl.tmAt, pl.tmBt,
                       pl.tmA2, pl.tmA2t, pl.tmO, pl.tmX, pl.p, pl.asrc, pl.gsrc));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_cols, uint32_t box_rows, int swizzle_bytes) {
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
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE,
                      swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for rows=%llu cols=%llu ld=%llu box=%ux%u base=%p", (int)r,
              (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_cols, box_rows, base);
    return B200DP_ECUDA;
  }
  return B200DP_OK;
}

// Row packing pays when a side of the GEMM is narrower than a 128-byte row: r = 64 / min(K, N), at most 4, as long
// as the packed K still splits into 64-wide chunks plus a 16/32 tail and the rows divide evenly.
int pw_gemm_pick_pack(int M, int N, int K) {
  const int narrow = K < N ? K : N;
  for (int r = 4; r > 1; r >>= 1) {
    if (narrow * r > 64 || M % r != 0) continue;
    const int tail = (K * r) % KCH;
    if (tail != 0 && tail != 16 && tail != 32) continue;
    return r;
  }
  return 1;
}

int pw_gemm_plan(PwGemmPlan* pl, int amode, int emode, int M, int N, int K, const void* a_tma, int lda,
                 const void* a0, const void* a1, const float* acoef, int a_relu6, const void* w, void* out, int ldo,
                 const void* aux, const float* ecoef, float* stats, int stats_stride, int max_ctas, int pack) {
  B200DP_REQUIRE(M > 0 && N > 0 && K > 0 && (N % 16) == 0 && (K % 16) == 0, "pw_gemm: M=%d N=%d K=%d invalid", M, N, K);
  B200DP_REQUIRE(pack >= 1 && pack <= 8 && M % pack == 0, "pw_gemm: pack=%d does not divide M=%d", pack, M);
  const int n_true = N, k_true = K;
  if (pack > 1) {
    B200DP_REQUIRE(lda == K && ldo == N, "pw_gemm: row packing needs dense tensors (lda=%d K=%d ldo=%d N=%d)", lda, K, ldo, N);
    M /= pack; N *= pack; K *= pack; lda *= pack; ldo *= pack;
  }
  B200DP_REQUIRE((lda % 8) == 0 && (ldo % 8) == 0, "pw_gemm: row pitches must be multiples of 8 elements");
  B200DP_REQUIRE(amode >= 0 && amode <= 2 && emode >= 0 && emode <= 2, "pw_gemm: bad mode");
  PwGemmParams& p = pl->p;
  p = PwGemmParams{};
  pl->asrc = BnFwdSrc{};
  pl->gsrc = BnBwdSrc{};
  p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldo = ldo;
  p.n_true = n_true; p.k_true = k_true;
  p.block_n = N < 128 ? N : 128;
  p.m_tiles = (M + BM - 1) / BM;
  // few row tiles (7x7 layers): 128-wide column blocks leave SMs idle and make every CTA stream a 128-column weight
  // tile per K chunk through its own ~27 B/clk of L2 bandwidth; 64-wide blocks use more SMs and halve that tile
  if (N > 64 && p.m_tiles * ((N + 127) / 128) < num_sms()) p.block_n = 64;
  // two raw A tiles per stage plus four epilogue buffers per 64 columns: this combination only fits with 64-wide blocks
  if (N > 64 && amode == A_BNGRAD && emode == E_MASKSTATS) p.block_n = 64;
  p.n_tiles = (N + p.block_n - 1) / p.block_n;
  p.kc = (K + KCH - 1) / KCH;
  p.a0 = static_cast<const __nv_bfloat16*>(a0);
  p.a1 = static_cast<const __nv_bfloat16*>(a1);
  p.acoef = acoef; p.a_relu6 = a_relu6;
  p.aux = static_cast<const __nv_bfloat16*>(aux);
  p.ecoef = ecoef; p.stats = stats; p.stats_stride = stats_stride;
  if (amode == A_TMA) B200DP_REQUIRE(a_tma != nullptr, "pw_gemm: A pointer missing");
  if (amode != A_TMA) B200DP_REQUIRE(a0 != nullptr && acoef != nullptr, "pw_gemm: register-path A needs a0 + acoef");
  if (amode == A_BNGRAD) B200DP_REQUIRE(a1 != nullptr, "pw_gemm: A_BNGRAD needs a1");
  if (emode != E_PLAIN) B200DP_REQUIRE(stats != nullptr, "pw_gemm: stats pointer missing");
  if (emode == E_MASKSTATS) B200DP_REQUIRE(aux != nullptr && ecoef != nullptr, "pw_gemm: E_MASKSTATS needs aux + ecoef");
  const int nch = (p.block_n + 63) / 64;
  const size_t fixed = (emode == E_MASKSTATS ? 4 : 2) * (size_t)nch * OUT_CHUNK +
                       ((amode == A_BNRELU ? 2 : amode == A_BNGRAD ? 3 : 0) * (size_t)K +
                        (emode == E_MASKSTATS ? 2 : 0) * (size_t)N + (emode != E_PLAIN ? 2048 : 0) + 8) * 4 +
                       (3 * MAX_STAGES + 8) * 8 + 16 + 1024;
  const size_t per_stage = (amode == A_BNGRAD ? 2 : 1) * (size_t)A_STAGE + (size_t)p.block_n * 128;
  const size_t budget = 220 * 1024;
  B200DP_REQUIRE(fixed + 2 * per_stage <= budget, "pw_gemm: shared memory budget exceeded (K=%d N=%d)", K, N);
  int stages = (int)((budget - fixed) / per_stage);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  p.stages = stages;
  pl->smem_bytes = fixed + stages * per_stage;
  pl->amode = amode; pl->emode = emode;
  const int cap = max_ctas > 0 ? max_ctas : num_sms();
  int groups = cap / p.n_tiles;                // CTAs per column block
  if (groups < 1) groups = 1;
  if (groups > p.m_tiles) groups = p.m_tiles;
  pl->grid = groups * p.n_tiles;
  const int tail = K % KCH;
  B200DP_REQUIRE(tail == 0 || tail == 16 || tail == 32, "pw_gemm: K %% 64 must be 0, 16 or 32 (K=%d)", K);
  p.tail_w = tail == 0 ? KCH : tail;
  const uint32_t wfull = K >= KCH ? KCH : p.tail_w;          // box width of the full-chunk maps (unused when K < 64)
  const void* a_first = amode == A_TMA ? a_tma : a0;       // register modes: raw tiles also arrive by TMA
  if (int rc = make_tmap_2d_bf16(&pl->tmA, a_first, M, K, lda, wfull, BM, wfull * 2)) return rc;
  if (int rc = make_tmap_2d_bf16(&pl->tmAt, a_first, M, K, lda, p.tail_w, BM, p.tail_w * 2)) return rc;
  if (amode == A_BNGRAD) {
    if (int rc = make_tmap_2d_bf16(&pl->tmA2, a1, M, K, lda, wfull, BM, wfull * 2)) return rc;
    if (int rc = make_tmap_2d_bf16(&pl->tmA2t, a1, M, K, lda, p.tail_w, BM, p.tail_w * 2)) return rc;
  } else {
    pl->tmA2 = CUtensorMap{};
    pl->tmA2t = CUtensorMap{};
  }
  if (int rc = make_tmap_2d_bf16(&pl->tmB, w, N, K, K, wfull, p.block_n, wfull * 2)) return rc;
  if (int rc = make_tmap_2d_bf16(&pl->tmBt, w, N, K, K, p.tail_w, p.block_n, p.tail_w * 2)) return rc;
  if (int rc = make_tmap_2d_bf16(&pl->tmO, out, M, N, ldo, 64, 32, 128)) return rc;     // each epilogue warp stores its 32 rows
  if (emode == E_MASKSTATS) {
    B200DP_REQUIRE(amode == A_TMA || amode == A_BNGRAD, "pw_gemm: E_MASKSTATS is implemented for the TMA and BN-backward A paths");
    if (int rc = make_tmap_2d_bf16(&pl->tmX, aux, M, N, ldo, 64, BM, 128)) return rc;
  } else {
    pl->tmX = CUtensorMap{};
  }
  return B200DP_OK;
}

unsigned long long* g_pw_trace = nullptr;

int pw_gemm_launch(const PwGemmPlan& pl_in, cudaStream_t s) {
  PwGemmPlan pl = pl_in;
  pl.p.trace = g_pw_trace;
  switch (pl.amode * 3 + pl.emode) {
    case 0: return launch_t<A_TMA, E_PLAIN>(pl, s);
    case 1: return launch_t<A_TMA, E_STATS>(pl, s);
    case 2: return launch_t<A_TMA, E_MASKSTATS>(pl, s);
    case 3: return launch_t<A_BNRELU, E_PLAIN>(pl, s);
    case 4: return launch_t<A_BNRELU, E_STATS>(pl, s);
    case 5: return launch_t<A_BNRELU, E_MASKSTATS>(pl, s);
    case 6: return launch_t<A_BNGRAD, E_PLAIN>(pl, s);
    case 7: return launch_t<A_BNGRAD, E_STATS>(pl, s);
    default: return launch_t<A_BNGRAD, E_MASKSTATS>(pl, s);
  }
}

}  // namespace b200dp

using namespace b200dp;

// Stand-alone entry point (tests, micro-benchmarks): builds the plan (tensor maps) and launches once.
extern "C" int b200dp_pw_gemm(int amode, int emode, int M, int N, int K, const void* a_tma, int lda,
                              const void* a0, const void* a1, const float* acoef, int a_relu6, const void* w,
                              void* out, int ldo, const void* aux, const float* ecoef, float* stats,
                              int stats_stride, int max_ctas, void* stream) {
  PwGemmPlan pl;
  if (int rc = pw_gemm_plan(&pl, amode, emode, M, N, K, a_tma, lda, a0, a1, acoef, a_relu6, w, out, ldo, aux, ecoef,
                            stats, stats_stride, max_ctas))
    return rc;
  return pw_gemm_launch(pl, static_cast<cudaStream_t>(stream));
}

// Same with row packing: `w` is blockdiag_pack(W), [pack*N][pack*K]; everything else as above (true sizes).
extern "C" int b200dp_pw_gemm_packed(int amode, int emode, int M, int N, int K, const void* a_tma, const void* a0,
                                     const void* a1, const float* acoef, int a_relu6, const void* w_blockdiag,
                                     void* out, const void* aux, const float* ecoef, float* stats, int stats_stride,
                                     int max_ctas, int pack, void* stream) {
  PwGemmPlan pl;
  if (int rc = pw_gemm_plan(&pl, amode, emode, M, N, K, a_tma, K, a0, a1, acoef, a_relu6, w_blockdiag, out, N, aux,
                            ecoef, stats, stats_stride, max_ctas, pack))
    return rc;
  return pw_gemm_launch(pl, static_cast<cudaStream_t>(stream));
}
extern "C" int b200dp_pw_gemm_pick_pack(int M, int N, int K) { return pw_gemm_pick_pack(M, N, K); }

extern "C" void b200dp_debug_pw_trace(void* buf) { b200dp::g_pw_trace = static_cast<unsigned long long*>(buf); }
