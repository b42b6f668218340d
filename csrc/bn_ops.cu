// BatchNorm plumbing and the small elementwise passes over the *bottleneck* tensors (K4, K6).
//
// Training-mode BatchNorm never runs as a pass over an expanded activation here: producers emit per-channel
// sums from their epilogues and consumers apply scale/shift (+ReLU6) while loading.  What remains are
//   * per-channel finalisation kernels (C <= 1280 values): sums -> mean / inv-std / scale / shift / running
//     statistics (forward), sums -> (ca, cb, cc, dgamma, dbeta) (backward);
//   * passes over the narrow block outputs (16..320 channels, 8.6 % of the activation volume):
//     y = scale*p_raw + shift (+ residual)            block output (linear bottleneck, K6 residual add)
//     dout = g1 (+ g2), sums for the projection BN     gradient join of the residual branch
//     dp = ca*dout + cb*p_raw + cc                     BatchNorm backward of the projection conv
// Reference behaviour replaced: cuDNN batch-norm forward/backward and ATen add kernels.
#include "bn_inline.cuh"
#include "common.cuh"

namespace b200dp {
namespace {

// ---- forward finalisation -----------------------------------------------------------------------------
__global__ void bn_fwd_finalize_kernel(const float* __restrict__ sums, int stride, const float* __restrict__ gamma,
                                       const float* __restrict__ beta, float count, float eps, float momentum,
                                       float* __restrict__ rmean, float* __restrict__ rvar,
                                       float* __restrict__ coef /*[2][C] scale, shift*/,
                                       float* __restrict__ saved /*[2][C] mean, invstd*/, int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = sums[c] / count;
  const float var = fmaxf(sums[stride + c] / count - mean * mean, 0.f);
  const float inv = rsqrtf(var + eps);
  const float sc = gamma[c] * inv;
  coef[c] = sc;
  coef[C + c] = beta[c] - mean * sc;
  saved[c] = mean;
  saved[C + c] = inv;
  if (rmean != nullptr) {   // torch: running = (1-m) running + m batch, variance unbiased
    rmean[c] = (1.f - momentum) * rmean[c] + momentum * mean;
    const float unb = count > 1.f ? var * count / (count - 1.f) : var;
    rvar[c] = (1.f - momentum) * rvar[c] + momentum * unb;
  }
}

__global__ void bn_eval_coef_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const float* __restrict__ rmean, const float* __restrict__ rvar, float eps,
                                    float* __restrict__ coef, int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = gamma[c] * rsqrtf(rvar[c] + eps);
  coef[c] = sc;
  coef[C + c] = beta[c] - rmean[c] * sc;
}

// ---- backward finalisation ----------------------------------------------------------------------------
// sums = (sum dz, sum dz*x) over the batch.  dx = ca*dz + cb*x + cc.
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ sums, int stride, const float* __restrict__ saved,
                                       const float* __restrict__ gamma, float count, float* __restrict__ coef3,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta, int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = saved[c], r = saved[C + c], g = gamma[c];
  const float db = sums[c];
  const float dg = r * (sums[stride + c] - mean * db);
  const float gr = g * r;
  coef3[c] = gr;
  const float cb = -gr * r * dg / count;
  coef3[C + c] = cb;
  coef3[2 * C + c] = -gr * db / count - cb * mean;
  dgamma[c] += dg;
  dbeta[c] += db;
}

// ---- passes over bottleneck tensors -------------------------------------------------------------------
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                    pack_bf16x2(f[6], f[7]));
}

// The three passes below share one shape.  A thread owns one 8-channel group for its whole life (its coefficients
// and partial sums stay in registers), a block owns a contiguous range of rows, and every trip issues FOUR
// independent 16-byte loads per operand before touching any of them -- these tensors are small (<= 51 MB), the
// kernels are latency-bound, and bytes in flight are what buys bandwidth.  Per-channel sums are reduced through a
// [row slot][C] staging array in shared memory and one atomic per channel per block (the earlier version did
// blockDim.x shared atomics onto C <= 32 addresses, a 64- to 128-way serialisation per block).
constexpr int NP_UNROLL = 4;

struct RowRange { long long begin, end; };
__device__ __forceinline__ RowRange block_rows(long long rows) {
  const long long per = (rows + gridDim.x - 1) / gridDim.x;
  RowRange r;
  r.begin = (long long)blockIdx.x * per;
  r.end = r.begin + per < rows ? r.begin + per : rows;
  return r;
}

// s_part[ti][C] holds this block's partial sums for `nstat` statistics; add them to dst[k * stride + c]
__device__ __forceinline__ void flush_partials(const float* s_part, int nstat, int ipb, int C, float* dst, int stride) {
  __syncthreads();
  for (int i = threadIdx.x; i < nstat * C; i += blockDim.x) {
    const int k = i / C, c = i - k * C;
    const float* p = s_part + (size_t)k * ipb * C + c;
    float a = 0.f;
    for (int t = 0; t < ipb; ++t) a += p[(size_t)t * C];
    atomicAdd(dst + (size_t)k * stride + c, a);
  }
}

// out = c0*a + c1*b + c2             (BatchNorm backward apply of the projection conv)
__global__ void __launch_bounds__(256)
affine_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
              const float* __restrict__ coef, int three, __nv_bfloat16* __restrict__ out, long long rows, int C,
              const BnBwdSrc src) {
  pdl_trigger();
  pdl_wait();
  const int CG = C >> 3;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  float k0[8], k1[8], k2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (src.sums != nullptr) {   // BatchNorm backward finalised here (first consumer); block 0 publishes it
      bn_bwd_channel(src, cg * 8 + j, blockIdx.x == 0 && ti == 0, k0[j], k1[j], k2[j]);
    } else {
      k0[j] = coef[cg * 8 + j];
      k1[j] = coef[C + cg * 8 + j];
      k2[j] = three ? coef[2 * C + cg * 8 + j] : 0.f;
    }
  }
  const bool hasb = b != nullptr;
  const RowRange rr = block_rows(rows);
  for (long long r0 = rr.begin + ti; r0 < rr.end; r0 += (long long)ipb * NP_UNROLL) {
    uint4 va[NP_UNROLL], vb[NP_UNROLL];
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        va[u] = ldg_stream(a + r * C + cg * 8);
        if (hasb) vb[u] = ldg_stream(b + r * C + cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        float x[8], y[8], o[8];
        unpack8(va[u], x);
        if (hasb) unpack8(vb[u], y);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (three) o[j] = fmaf(k0[j], x[j], fmaf(k1[j], y[j], k2[j]));
          else o[j] = fmaf(k0[j], x[j], k1[j]) + (hasb ? y[j] : 0.f);
        }
        stg_stream(out + r * C + cg * 8, pack8(o));
      }
    }
  }
}

// dout = g1 (+ g2) [written only when g2 != null], sums += (sum dout, sum dout*x) per channel.
__global__ void __launch_bounds__(256)
grad_join_stats_kernel(const __nv_bfloat16* __restrict__ g1, const __nv_bfloat16* __restrict__ g2,
                       const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ dout,
                       float* __restrict__ sums, int stride, long long rows, int C) {
  extern __shared__ float s_part[];   // [2][ipb][C]
  pdl_trigger();
  pdl_wait();
  const int CG = C >> 3;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  float s0[8], s1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.f;
  const bool join = g2 != nullptr;
  const RowRange rr = block_rows(rows);
  for (long long r0 = rr.begin + ti; r0 < rr.end; r0 += (long long)ipb * NP_UNROLL) {
    uint4 va[NP_UNROLL], vx[NP_UNROLL], vb[NP_UNROLL];
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        va[u] = ldg_stream(g1 + r * C + cg * 8);
        vx[u] = ldg_stream(x + r * C + cg * 8);
        if (join) vb[u] = ldg_stream(g2 + r * C + cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        float a[8], xv[8];
        unpack8(va[u], a);
        unpack8(vx[u], xv);
        if (join) {
          float bq[8];
          unpack8(vb[u], bq);
#pragma unroll
          for (int j = 0; j < 8; ++j) a[j] += bq[j];
          const uint4 pk = pack8(a);
          stg_stream(dout + r * C + cg * 8, pk);
          unpack8(pk, a);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0[j] += a[j]; s1[j] = fmaf(a[j], xv[j], s1[j]); }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s_part[(size_t)ti * C + cg * 8 + j] = s0[j];
    s_part[(size_t)(ipb + ti) * C + cg * 8 + j] = s1[j];
  }
  flush_partials(s_part, 2, ipb, C, sums, stride);
}

// Block output: out = scale*raw + shift (+ res), and the per-channel column sums of the ROUNDED result.  The next
// expand conv's weight gradient factors its BatchNorm-backward affine out of the reduction and needs sum_m x[m][k].
__global__ void __launch_bounds__(256)
block_out_kernel(const __nv_bfloat16* __restrict__ raw, const __nv_bfloat16* __restrict__ res,
                 const float* __restrict__ coef, __nv_bfloat16* __restrict__ out, float* __restrict__ colsum,
                 long long rows, int C, const BnFwdSrc src) {
  extern __shared__ float s_part[];   // [ipb][C]
  pdl_trigger();
  pdl_wait();
  const int CG = C >> 3;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  float k0[8], k1[8], acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (src.sums != nullptr) bn_fwd_channel(src, cg * 8 + j, blockIdx.x == 0 && ti == 0, k0[j], k1[j]);
    else { k0[j] = coef[cg * 8 + j]; k1[j] = coef[C + cg * 8 + j]; }
    acc[j] = 0.f;
  }
  const bool hasr = res != nullptr;
  const RowRange rr = block_rows(rows);
  for (long long r0 = rr.begin + ti; r0 < rr.end; r0 += (long long)ipb * NP_UNROLL) {
    uint4 va[NP_UNROLL], vr[NP_UNROLL];
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        va[u] = ldg_stream(raw + r * C + cg * 8);
        if (hasr) vr[u] = ldg_stream(res + r * C + cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < NP_UNROLL; ++u) {
      const long long r = r0 + (long long)u * ipb;
      if (r < rr.end) {
        float x[8], y[8], o[8];
        unpack8(va[u], x);
        if (hasr) unpack8(vr[u], y);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf(k0[j], x[j], k1[j]) + (hasr ? y[j] : 0.f);
        const uint4 pk = pack8(o);
        stg_stream(out + r * C + cg * 8, pk);
        if (colsum != nullptr) {
          unpack8(pk, o);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += o[j];
        }
      }
    }
  }
  if (colsum != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) s_part[(size_t)ti * C + cg * 8 + j] = acc[j];
    flush_partials(s_part, 1, ipb, C, colsum, 0);
  }
}

// Weight preparation: fp32 master [N][K] -> bf16 [N][K] and bf16 transposed [K][N] for every pointwise conv.
// For row-packed layers (pw_gemm.h) also the block-diagonal copies blockdiag_r(W) [rN][rK] (forward) and
// blockdiag_r(W^T) [rK][rN] (data gradient); their off-diagonal blocks are zero from allocation and never written.
struct WPrepEntry { long long off; int n, k; int rf, rd; long long off_pf, off_pd; };

__global__ void __launch_bounds__(256)
weight_prep_kernel(const float* __restrict__ master, __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wt,
                   __nv_bfloat16* __restrict__ wpack, const WPrepEntry* __restrict__ table) {
  pdl_trigger();
  pdl_wait();
  const WPrepEntry e = table[blockIdx.y];
  const int total = e.n * e.k;
  // every destination is written in ITS OWN contiguous order (the strided side is the fp32 read, which hits L2)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i / e.k, k = i - n * e.k;               // (n, k) order: W and its block-diagonal copy
    const __nv_bfloat16 v = __float2bfloat16(master[e.off + i]);
    if (w != nullptr) w[e.off + i] = v;
    for (int p = 0; p < e.rf && e.rf > 1; ++p)
      wpack[e.off_pf + (long long)(p * e.n + n) * (e.rf * e.k) + p * e.k + k] = v;
    const int k2 = i / e.n, n2 = i - k2 * e.n;            // (k, n) order: W^T and its block-diagonal copy
    const __nv_bfloat16 vt = __float2bfloat16(master[e.off + (long long)n2 * e.k + k2]);
    wt[e.off + i] = vt;
    for (int p = 0; p < e.rd && e.rd > 1; ++p)
      wpack[e.off_pd + (long long)(p * e.k + k2) * (e.rd * e.n) + p * e.n + n2] = vt;
  }
}

}  // namespace

int bn_fwd_finalize_launch(const float* sums, int stride, const float* gamma, const float* beta, float count,
                           float eps, float momentum, float* rmean, float* rvar, float* coef, float* saved, int C,
                           cudaStream_t s) {
  B200DP_CUDA(launch_k(bn_fwd_finalize_kernel, dim3((C + 127) / 128), dim3(128), 0, s, sums, stride, gamma, beta, count, eps, momentum, rmean, rvar,
                                                         coef, saved, C));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int bn_eval_coef_launch(const float* gamma, const float* beta, const float* rmean, const float* rvar, float eps,
                        float* coef, int C, cudaStream_t s) {
  B200DP_CUDA(launch_k(bn_eval_coef_kernel, dim3((C + 127) / 128), dim3(128), 0, s, gamma, beta, rmean, rvar, eps, coef, C));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int bn_bwd_finalize_launch(const float* sums, int stride, const float* saved, const float* gamma, float count,
                           float* coef3, float* dgamma, float* dbeta, int C, cudaStream_t s) {
  B200DP_CUDA(launch_k(bn_bwd_finalize_kernel, dim3((C + 127) / 128), dim3(128), 0, s, sums, stride, saved, gamma, count, coef3, dgamma, dbeta, C));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

// grid: enough blocks for two per SM, fewer when the tensor is tiny (every block pays one reduction tail)
static int narrow_grid(long long rows, int ipb) {
  long long blocks = (rows + (long long)ipb * NP_UNROLL - 1) / ((long long)ipb * NP_UNROLL);
  const long long cap = (long long)num_sms() * 2;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

int affine_launch(const void* a, const void* b, const float* coef, int three, void* out, long long rows, int C,
                  cudaStream_t s, const BnBwdSrc* src) {
  B200DP_REQUIRE(C % 8 == 0 && C / 8 <= 256, "affine: C=%d unsupported", C);
  const int CG = C / 8;
  const int threads = (256 / CG) * CG;
  const int ipb = threads / CG;
  B200DP_CUDA(launch_k(affine_kernel, dim3((unsigned)narrow_grid(rows, ipb)), dim3(threads), 0, s,
                       static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), coef, three,
                       static_cast<__nv_bfloat16*>(out), rows, C, src ? *src : BnBwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int block_out_launch(const void* raw, const void* res, const float* coef, void* out, float* colsum, long long rows,
                     int C, cudaStream_t s, const BnFwdSrc* src) {
  B200DP_REQUIRE(C % 8 == 0 && C / 8 <= 256, "block_out: C=%d unsupported", C);
  const int CG = C / 8;
  const int threads = (256 / CG) * CG;
  const int ipb = threads / CG;
  B200DP_CUDA(launch_k(block_out_kernel, dim3((unsigned)narrow_grid(rows, ipb)), dim3(threads),
                       (size_t)ipb * C * sizeof(float), s, static_cast<const __nv_bfloat16*>(raw),
                       static_cast<const __nv_bfloat16*>(res), coef, static_cast<__nv_bfloat16*>(out), colsum, rows, C,
                       src ? *src : BnFwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int grad_join_stats_launch(const void* g1, const void* g2, const void* x, void* dout, float* sums, int stride,
                           long long rows, int C, cudaStream_t s) {
  B200DP_REQUIRE(C % 8 == 0 && C / 8 <= 256, "grad_join: C=%d unsupported", C);
  const int CG = C / 8;
  const int threads = (256 / CG) * CG;
  const int ipb = threads / CG;
  B200DP_CUDA(launch_k(grad_join_stats_kernel, dim3((unsigned)narrow_grid(rows, ipb)), dim3(threads),
                       2 * (size_t)ipb * C * sizeof(float), s, static_cast<const __nv_bfloat16*>(g1),
                       static_cast<const __nv_bfloat16*>(g2), static_cast<const __nv_bfloat16*>(x),
                       static_cast<__nv_bfloat16*>(dout), sums, stride, rows, C));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int weight_prep_launch(const float* master, void* w, void* wt, void* wpack, const void* table_dev, int nlayers,
                       cudaStream_t s) {
  dim3 grid(32, nlayers);
  B200DP_CUDA(launch_k(weight_prep_kernel, grid, dim3(256), 0, s, master, static_cast<__nv_bfloat16*>(w),
                       static_cast<__nv_bfloat16*>(wt), static_cast<__nv_bfloat16*>(wpack),
                       static_cast<const WPrepEntry*>(table_dev)));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_bn_fwd_finalize(const float* sums, int stride, const float* gamma, const float* beta,
                                      float count, float eps, float momentum, float* rmean, float* rvar, float* coef,
                                      float* saved, int C, void* stream) {
  return bn_fwd_finalize_launch(sums, stride, gamma, beta, count, eps, momentum, rmean, rvar, coef, saved, C,
                                static_cast<cudaStream_t>(stream));
}
extern "C" int b200dp_bn_bwd_finalize(const float* sums, int stride, const float* saved, const float* gamma,
                                      float count, float* coef3, float* dgamma, float* dbeta, int C, void* stream) {
  return bn_bwd_finalize_launch(sums, stride, saved, gamma, count, coef3, dgamma, dbeta, C,
                                static_cast<cudaStream_t>(stream));
}
extern "C" int b200dp_affine(const void* a, const void* b, const float* coef, int three, void* out, long long rows,
                             int C, void* stream) {
  return affine_launch(a, b, coef, three, out, rows, C, static_cast<cudaStream_t>(stream), nullptr);
}
extern "C" int b200dp_grad_join_stats(const void* g1, const void* g2, const void* x, void* dout, float* sums,
                                      int stride, long long rows, int C, void* stream) {
  return grad_join_stats_launch(g1, g2, x, dout, sums, stride, rows, C, static_cast<cudaStream_t>(stream));
}
extern "C" int b200dp_block_out(const void* raw, const void* res, const float* coef, void* out, float* colsum,
                                long long rows, int C, void* stream) {
  return block_out_launch(raw, res, coef, out, colsum, rows, C, static_cast<cudaStream_t>(stream), nullptr);
}
