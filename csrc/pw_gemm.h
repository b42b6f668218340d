// Internal interface of the tcgen05 pointwise-convolution GEMM (csrc/pw_gemm.cu).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stddef.h>

#include "bn_inline.cuh"

namespace b200dp {

enum { A_TMA = 0, A_BNRELU = 1, A_BNGRAD = 2 };
enum { E_PLAIN = 0, E_STATS = 1, E_MASKSTATS = 2 };

struct PwGemmParams {
  int M, N, K;                  // rows, output channels, input channels (channels are multiples of 16) -- AFTER
                                // row packing: `pack` consecutive rows are treated as one row of pack*K channels
  int n_true, k_true;           // channel counts before packing (coefficient / statistic arrays have these lengths)
  int lda, ldo;                 // row pitches (elements) of the register-path A sources / of out and aux
  int block_n, n_tiles, m_tiles, kc, stages;
  int tail_w;                   // width of the last K chunk: 16, 32 or 64 elements
  const __nv_bfloat16* a0;      // A_BNRELU: raw activation          A_BNGRAD: dz
  const __nv_bfloat16* a1;      //                                    A_BNGRAD: raw activation x
  const float* acoef;           // A_BNRELU: [2][K] scale, shift      A_BNGRAD: [3][K] ca, cb, cc
  int a_relu6;                  // A_BNRELU: clamp to [0, 6] after the affine
  const __nv_bfloat16* aux;     // E_MASKSTATS: raw forward activation of the tensor being differentiated [M][ldo]
  const float* ecoef;           // E_MASKSTATS: [2][N] scale, shift of that activation's BatchNorm
  float* stats;                 // E_STATS: sum y, sum y^2   E_MASKSTATS: sum dz, sum dz*aux   ([2] x stats_stride)
  int stats_stride;
  unsigned long long* trace;    // debugging: CTA 0 writes %globaltimer at its pipeline milestones (null = off)
};

struct PwGemmPlan {
  CUtensorMap tmA, tmB, tmAt, tmBt, tmA2, tmA2t, tmO, tmX;   // At/Bt: maps for the (possibly narrower) last K chunk
  PwGemmParams p;
  BnFwdSrc asrc;                // A_BNRELU: finalise the A-side BatchNorm from its producer's sums in the prologue
  BnBwdSrc gsrc;                // A_BNGRAD: compute (ca, cb, cc) from the sums in the prologue, side-effect free
  int amode, emode, grid;
  size_t smem_bytes;
};

// `pack` = r > 1: narrow layers (K or N < 64) are run as the GEMM [M/r x rK] x blockdiag_r(W)^T -> [M/r x rN] over
// the same memory (r consecutive pixels form one row), so that TMA moves 128-byte rows and one tcgen05.mma covers r
// times as many pixels.  Requires dense tensors (lda == K, ldo == N), M % r == 0, and `w` = the block-diagonal
// [rN][rK] matrix (weight_prep builds it).  M, N, K, coefficient and statistic arrays are the TRUE sizes.
int pw_gemm_plan(PwGemmPlan* pl, int amode, int emode, int M, int N, int K, const void* a_tma, int lda,
                 const void* a0, const void* a1, const float* acoef, int a_relu6, const void* w, void* out, int ldo,
                 const void* aux, const float* ecoef, float* stats, int stats_stride, int max_ctas, int pack = 1);
int pw_gemm_pick_pack(int M, int N, int K);     // packing factor the executor uses for a layer (1, 2 or 4)
int pw_gemm_launch(const PwGemmPlan& pl, cudaStream_t s);

// 2-D bf16 row-major tensor map [rows][cols] (row pitch `ld` elements), box {box_cols, box_rows}, 128-byte swizzle.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_cols, uint32_t box_rows, int swizzle_bytes = 128);

}  // namespace b200dp
