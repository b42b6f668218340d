// Internal launch interface of every kernel family, used by the executor (csrc/engine.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "bn_inline.cuh"
#include "pw_gemm.h"

namespace b200dp {

// dwconv.cu
int dw_fwd_launch(const void* x, const float* scale, const float* shift, const float* w, void* y, float* stats,
                  int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s, const BnFwdSrc* src = nullptr);
int dw_bwd_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                  const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                  int H, int W, int C, int stride, cudaStream_t s, const BnBwdSrc* src = nullptr);
// dw_tile.cu: stride-1 forward, TMA halo tiles in shared memory + register-window stencil on the CUDA cores
bool dw_tile_supported(int B, int H, int W, int C, int stride);
bool dw_tile_fwd_supported(int B, int H, int W, int C, int stride);    // the forward kernel also does stride 2
int dw_tile_fwd_launch(const void* x, const float* scale, const float* shift, const float* w, void* y, float* stats,
                       int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s,
                       const BnFwdSrc* src = nullptr);
int dw_tile_bwd_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                       const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                       int H, int W, int C, cudaStream_t s, const BnBwdSrc* src = nullptr);
bool dw_tile_bwd_supported(int B, int H, int W, int C, int stride);
int dw_tile_bwd2_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                        const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                        int H, int W, int C, cudaStream_t s, const BnBwdSrc* src = nullptr);
// stem.cu
int stem_fwd_launch(const void* x, const float* w, void* y, float* stats, int stats_stride, int B, int H, int W,
                    cudaStream_t s);
int stem_wgrad_launch(const void* x, const void* dz, const void* yraw, const float* coef, float* dw, int B, int H,
                      int W, cudaStream_t s);
// the stem as a GEMM: im2col [B*Ho*Wo][32] (+ column sums), bf16 weight layout, gradient fold-back
int stem_im2col_launch(const void* x, void* col, float* asum, int B, int H, int W, cudaStream_t s);
int stem_weight_prep_launch(const float* w, void* wout, int pack, cudaStream_t s);
int stem_dw_fold_launch(const float* scratch, float* dw, cudaStream_t s);
// pw_wgrad.cu
int pw_wgrad_launch(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0, const void* t1,
                    int ld_t, const float* coef, const float* asum, float* dw, int ldw, int max_ctas, cudaStream_t s,
                    int pack = 1, const BnBwdSrc* src = nullptr);
int pw_wgrad_pick_pack(int mode, int M, int N, int K);
// bn_ops.cu
int bn_fwd_finalize_launch(const float* sums, int stride, const float* gamma, const float* beta, float count,
                           float eps, float momentum, float* rmean, float* rvar, float* coef, float* saved, int C,
                           cudaStream_t s);
int bn_eval_coef_launch(const float* gamma, const float* beta, const float* rmean, const float* rvar, float eps,
                        float* coef, int C, cudaStream_t s);
int bn_bwd_finalize_launch(const float* sums, int stride, const float* saved, const float* gamma, float count,
                           float* coef3, float* dgamma, float* dbeta, int C, cudaStream_t s);
// `src` (optional, here and below): finalise the BatchNorm from the producer's sums in this kernel's prologue
// (bn_inline.cuh) instead of reading the coefficient array
int affine_launch(const void* a, const void* b, const float* coef, int three, void* out, long long rows, int C,
                  cudaStream_t s, const BnBwdSrc* src = nullptr);
int block_out_launch(const void* raw, const void* res, const float* coef, void* out, float* colsum, long long rows,
                     int C, cudaStream_t s, const BnFwdSrc* src = nullptr);
int grad_join_stats_launch(const void* g1, const void* g2, const void* x, void* dout, float* sums, int stride,
                           long long rows, int C, cudaStream_t s);
int weight_prep_launch(const float* master, void* w, void* wt, void* wpack, const void* table_dev, int nlayers,
                       cudaStream_t s);
// head.cu
int head_pool_launch(const void* raw, const float* coef, float* pooled, float* drop, int B, int HW, int C,
                     float p_drop, uint32_t seed, const int* step, cudaStream_t s, const BnFwdSrc* src = nullptr);
int head_fc_launch(const float* pooled, const float* w, const float* bias, float* logits, int B, int C, int classes,
                   cudaStream_t s);
int head_bwd_launch(const float* dlogits, const float* pooled, const float* drop, const float* w, float* dpool,
                    float* dw, float* dbias, const void* raw, const float* coef, void* dz, float* sums, int stride,
                    int B, int HW, int C, int classes, cudaStream_t s);

}  // namespace b200dp
