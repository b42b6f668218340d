// Pointwise-convolution weight gradient on tcgen05 (SURVEY W5 "wgrad", K1).
//
//   dW[n][k] += sum_m G[m][n] * A[m][k]          (fp32, accumulated into the flat gradient arena)
//
// Both operands are [rows m][channels] NHWC tensors, i.e. the reduction index m is the *slow* index in
// memory: they are MN-major UMMA operands.  A tile of 64 rows x 64 channels lands in shared memory as
// 64 lines of 128 bytes with the 128-byte swizzle (exactly what a TMA box {64 ch, 64 rows} writes), and
// tcgen05.mma consumes it with the "MN-major" descriptor bits set -- no transpose pass, no im2col.
//
//   D[k (128 TMEM lanes)][n (<= 256 columns)] += P[m x 128 k-channels]^T * Q[m x n-channels]
//   P = A side (conv input),  Q = G side (gradient of the conv output)
//
// One side of every wgrad needs a BatchNorm transform that cannot be materialised cheaply:
//   MODE 0 (expand / last conv):  P = finished activation, pure, via TMA;   Q = ca*dz + cb*y_raw + cc on load
//   MODE 1 (project conv):        P = relu6(scale*d_raw + shift) on load;   Q = materialised gradient, via TMA
// The transformed side is produced by 8 warps (global -> registers -> swizzled smem), the other by TMA.
// Work items = (128-wide k chunk) x (n chunk) x (split of the m range); each item's 128 x n fp32 tile is
// added to dW with coalesced red.global.add.f32 (lane = k is the contiguous index of dW).
// Reference behaviour replaced: cuDNN convolution-backward-filter for the 34 1x1 convs.
#include "common.cuh"
#include "pw_gemm.h"
#include "sm100.cuh"

namespace b200dp {
namespace {

using namespace sm100;

constexpr int MK = 64;                 // m rows per pipeline stage
constexpr int BLK = MK * 128;          // one 64-channel block of a stage: 64 lines x 128 B
constexpr int P_BLKS = 2;              // 128 k-channels
constexpr int MAXQ_BLKS = 4;           // up to 256 n-channels
constexpr int MAX_STAGES = 6;
constexpr int PROD_THREADS = 256;
constexpr int TMEM_COLS = 512;         // 2 accumulator stages x 256 columns

struct WgradParams {
  int M, N, K;                 // rows, G channels (cout), A channels (cin)
  int lda, ldg, ldw;           // row pitches (elements) of A-side, G-side tensors and of dW
  int nq, qblks;               // n-chunk width (multiple of 16, <= 256) and its number of 64-wide blocks
  int kchunks, nchunks, msplit, chunks_per_split, stages;
  const __nv_bfloat16* t0;     // transformed side, first source  (MODE 0: dz      MODE 1: d_raw)
  const __nv_bfloat16* t1;     // transformed side, second source (MODE 0: y_raw)
  const float* coef;           // MODE 0: [3][N] ca, cb, cc     MODE 1: [2][K] scale, shift
  float* dw;                   // [N][ldw] fp32
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ void sts16(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(512, 1)
pw_wgrad_kernel(const __grid_constant__ CUtensorMap tmPure, const __grid_constant__ WgradParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = P.stages;
  const int stage_bytes = (P_BLKS + P.qblks) * BLK;
  float* s_coef = reinterpret_cast<float*>(base + (size_t)stages * stage_bytes);
  const int ncoef = MODE == 0 ? 3 * P.N : 2 * P.K;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_coef + ((ncoef + 3) & ~3));
  uint64_t* full = bars;
  uint64_t* empty = bars + MAX_STAGES;
  uint64_t* tmem_full = bars + 2 * MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = P.kchunks * P.nchunks * P.msplit;
  const int total_chunks = (P.M + MK - 1) / MK;

  if (warp == 0 && lane == 0) prefetch_tmap(&tmPure);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1 + PROD_THREADS);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
  for (int i = threadIdx.x; i < ncoef; i += blockDim.x) s_coef[i] = P.coef[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // item -> (k chunk, n chunk, m split); k fastest so concurrently running CTAs share the Q-side rows in L2
  auto decode = [&](int item, int& k0, int& n0, int& c_lo, int& c_hi) {
    const int kc = item % P.kchunks;
    const int nc = (item / P.kchunks) % P.nchunks;
    const int ms = item / (P.kchunks * P.nchunks);
    k0 = kc * 128;
    n0 = nc * P.nq;
    c_lo = ms * P.chunks_per_split;
    c_hi = min(c_lo + P.chunks_per_split, total_chunks);
  };

  if (warp == 0) {
    // =========================== TMA producer (pure side) ===========================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      const uint32_t tx = (MODE == 0 ? P_BLKS : P.qblks) * BLK;
      for (int item = blockIdx.x; item < items; item += gridDim.x) {
        int k0, n0, c_lo, c_hi;
        decode(item, k0, n0, c_lo, c_hi);
        for (int c = c_lo; c < c_hi; ++c) {
          mbar_wait(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], tx);
          uint8_t* st = base + (size_t)s * stage_bytes;
          if (MODE == 0) {
            for (int b = 0; b < P_BLKS; ++b) tma_load_2d(st + b * BLK, &tmPure, &full[s], k0 + 64 * b, c * MK);
          } else {
            for (int b = 0; b < P.qblks; ++b)
              tma_load_2d(st + (P_BLKS + b) * BLK, &tmPure, &full[s], n0 + 64 * b, c * MK);
          }
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, P.nq, 1, 1);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
        int k0, n0, c_lo, c_hi;
        decode(item, k0, n0, c_lo, c_hi);
        const int as = it & 1;
        mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int c = c_lo; c < c_hi; ++c) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t st = smem_u32(base + (size_t)s * stage_bytes);
          const uint64_t dp = umma_desc_mnmajor_sw128(st, BLK);
          const uint64_t dq = umma_desc_mnmajor_sw128(st + P_BLKS * BLK, BLK);
#pragma unroll
          for (int k = 0; k < MK / 16; ++k)   // 16 m-rows per MMA = 2 swizzle groups = 2048 B
            umma_bf16(d_tmem, dp + (uint64_t)(128 * k), dq + (uint64_t)(128 * k), idesc, (c > c_lo) | (k > 0));
          umma_commit(&empty[s]);
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        umma_commit(&tmem_full[as]);
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =========================== epilogue: TMEM -> red.global.add ===========================
    const int q = warp & 3;
    int it = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x, ++it) {
      int k0, n0, c_lo, c_hi;
      decode(item, k0, n0, c_lo, c_hi);
      const int as = it & 1;
      mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      const int k = k0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * 256;
      for (int c16 = 0; c16 < (P.nq >> 4); ++c16) {
        uint32_t r[16];
        tmem_ld_x16(taddr + c16 * 16, r);
        tmem_ld_wait();
        if (k < P.K && c_hi > c_lo) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n0 + c16 * 16 + j;
            if (n < P.N) atomicAdd(P.dw + (size_t)n * P.ldw + k, __uint_as_float(r[j]));
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
    }
  } else if (warp >= 8) {
    // =========================== producers (transformed side) ===========================
    const int pt = threadIdx.x - 256;
    const int p = pt & 7, rb = pt >> 3;          // 16-byte piece p of rows rb, rb + 32
    const int nblk = MODE == 0 ? P.qblks : P_BLKS;
    const int ld = MODE == 0 ? P.ldg : P.lda;
    const int cdim = MODE == 0 ? P.N : P.K;
    int s = 0;
    uint32_t ph = 0;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      int k0, n0, c_lo, c_hi;
      decode(item, k0, n0, c_lo, c_hi);
      const int ch0 = MODE == 0 ? n0 : k0;
      const int chw = MODE == 0 ? P.nq : 128;
      for (int c = c_lo; c < c_hi; ++c) {
        mbar_wait(&empty[s], ph ^ 1);
        const uint32_t st = smem_u32(base + (size_t)s * stage_bytes) + (MODE == 0 ? P_BLKS * BLK : 0);
        for (int b = 0; b < nblk; ++b) {
          const int lc = b * 64 + p * 8;                 // channel offset inside the chunk
          const int col = ch0 + lc;
          const bool colok = lc < chw && col < cdim;
          float c0[8], c1[8], c2[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            c0[j] = colok ? s_coef[col + j] : 0.f;
            c1[j] = colok ? s_coef[cdim + col + j] : 0.f;
            c2[j] = (MODE == 0 && colok) ? s_coef[2 * cdim + col + j] : 0.f;
          }
          uint4 v0[2], v1[2];
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int gm = c * MK + rb + 32 * i;
            v0[i] = make_uint4(0u, 0u, 0u, 0u);
            v1[i] = make_uint4(0u, 0u, 0u, 0u);
            if (colok && gm < P.M) {
              v0[i] = ldg_stream(P.t0 + (size_t)gm * ld + col);
              if (MODE == 0) v1[i] = ldg_stream(P.t1 + (size_t)gm * ld + col);
            }
          }
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int r = rb + 32 * i;
            const bool ok = colok && (c * MK + r) < P.M;
            float x[8], g[8], y[8];
            unpack8(v0[i], x);
            unpack8(v1[i], g);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (MODE == 0) y[j] = ok ? fmaf(c0[j], x[j], fmaf(c1[j], g[j], c2[j])) : 0.f;
              else y[j] = ok ? relu6f(fmaf(x[j], c0[j], c1[j])) : 0.f;
            }
            sts16(st + b * BLK + r * 128 + ((p ^ (r & 7)) << 4),
                  make_uint4(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]), pack_bf16x2(y[4], y[5]),
                             pack_bf16x2(y[6], y[7])));
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&full[s]);
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace

// mode 0: a_pure = finished input activation [M][K] (TMA); g0 = dz, g1 = y_raw [M][N]; coef = [3][N]
// mode 1: g_pure = materialised output gradient [M][N] (TMA); a0 = d_raw [M][K]; coef = [2][K]
int pw_wgrad_launch(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0, const void* t1,
                    int ld_t, const float* coef, float* dw, int ldw, int max_ctas, cudaStream_t s) {
  B200DP_REQUIRE(mode == 0 || mode == 1, "pw_wgrad: bad mode");
  B200DP_REQUIRE(M > 0 && N % 16 == 0 && K % 16 == 0, "pw_wgrad: M=%d N=%d K=%d invalid", M, N, K);
  B200DP_REQUIRE(pure && t0 && coef && dw && (mode == 1 || t1), "pw_wgrad: missing pointers");
  WgradParams p{};
  p.M = M; p.N = N; p.K = K;
  p.lda = mode == 0 ? ld_pure : ld_t;
  p.ldg = mode == 0 ? ld_t : ld_pure;
  p.ldw = ldw;
  p.nchunks = (N + 255) / 256;
  p.nq = ((N + p.nchunks - 1) / p.nchunks + 15) / 16 * 16;
  p.qblks = (p.nq + 63) / 64;
  p.kchunks = (K + 127) / 128;
  const int total_chunks = (M + MK - 1) / MK;
  const int cap = max_ctas > 0 ? max_ctas : num_sms();
  int msplit = (2 * cap + p.kchunks * p.nchunks - 1) / (p.kchunks * p.nchunks);
  const int max_split = (total_chunks + 3) / 4;        // at least 4 stages of work per item
  if (msplit > max_split) msplit = max_split;
  if (msplit < 1) msplit = 1;
  p.chunks_per_split = (total_chunks + msplit - 1) / msplit;
  p.msplit = (total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  p.t0 = static_cast<const __nv_bfloat16*>(t0);
  p.t1 = static_cast<const __nv_bfloat16*>(t1);
  p.coef = coef; p.dw = dw;
  const size_t stage_bytes = (size_t)(P_BLKS + p.qblks) * BLK;
  const size_t fixed = ((mode == 0 ? 3 * (size_t)N : 2 * (size_t)K) + 4) * 4 + (2 * MAX_STAGES + 4) * 8 + 16 + 1024;
  const size_t budget = 220 * 1024;
  int stages = (int)((budget - fixed) / stage_bytes);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  B200DP_REQUIRE(stages >= 2, "pw_wgrad: shared memory budget exceeded");
  p.stages = stages;
  const size_t smem = fixed + stages * stage_bytes;
  CUtensorMap tm;
  if (mode == 0) {
    if (int rc = make_tmap_2d_bf16(&tm, pure, M, K, ld_pure, 64, MK)) return rc;
  } else {
    if (int rc = make_tmap_2d_bf16(&tm, pure, M, N, ld_pure, 64, MK)) return rc;
  }
  const int items = p.kchunks * p.nchunks * p.msplit;
  const int grid = items < cap ? items : cap;
  if (mode == 0) {
    B200DP_CUDA(cudaFuncSetAttribute(pw_wgrad_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    pw_wgrad_kernel<0><<<grid, 512, smem, s>>>(tm, p);
  } else {
    B200DP_CUDA(cudaFuncSetAttribute(pw_wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    pw_wgrad_kernel<1><<<grid, 512, smem, s>>>(tm, p);
  }
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_pw_wgrad(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0,
                               const void* t1, int ld_t, const float* coef, float* dw, int ldw, int max_ctas,
                               void* stream) {
  return pw_wgrad_launch(mode, M, N, K, pure, ld_pure, t0, t1, ld_t, coef, dw, ldw, max_ctas,
                         static_cast<cudaStream_t>(stream));
}
