// Stem convolution 3x3 stride 2, 3 -> 32 channels, NHWC bf16.
//
// Forward : reads the pre-processed image (bf16 NHWC with 4 channels, the 4th zero, so a pixel is one
//           8-byte load), writes the raw conv output [B,112,112,32] and the per-channel sum / sum of squares
//           for its BatchNorm.  One thread per output pixel, 32 accumulators, weights broadcast from smem.
// Backward: only the weight gradient exists (the input image needs no gradient).  dy = ca*dz + cb*y_raw + cc is
//           formed on load (BatchNorm backward), pixels are staged 32 at a time per warp in shared memory and
//           each lane owns a 4-tap x 8-channel block of the [27][32] gradient.
// K = 27 is too small for a tcgen05 tile to pay off (10.8 of the 299.5 MMAC/image); this stays on CUDA cores.
// Reference behaviour replaced: cuDNN conv fwd + wgrad for features.0.0 (cifar10_mpi_mobilenet_224.py:177-179).
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int CO = 32;
constexpr int KT = 27;

struct StemFwdParams {
  const __nv_bfloat16* x;   // [B,H,W,4]
  const float* w;           // [27][32] fp32, (ky,kx,ci)-major
  __nv_bfloat16* y;         // [B,Ho,Wo,32]
  float* stats;             // [2][stats_stride] or null
  int B, H, W, Ho, Wo, stats_stride;
};

__global__ void __launch_bounds__(256)
stem_fwd_kernel(const StemFwdParams P) {
  pdl_trigger();
  pdl_wait();
  __shared__ __align__(16) float s_w[KT * CO];
  __shared__ float s_red[2 * CO];
  for (int i = threadIdx.x; i < KT * CO; i += blockDim.x) s_w[i] = P.w[i];
  if (threadIdx.x < 2 * CO) s_red[threadIdx.x] = 0.f;
  __syncthreads();
  float s0[CO], s1[CO];
#pragma unroll
  for (int c = 0; c < CO; ++c) s0[c] = s1[c] = 0.f;
  const long long npix = (long long)P.B * P.Ho * P.Wo;
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < npix;
       pix += (long long)gridDim.x * blockDim.x) {
    const int wo = (int)(pix % P.Wo);
    const int ho = (int)((pix / P.Wo) % P.Ho);
    const int n = (int)(pix / ((long long)P.Wo * P.Ho));
    float acc[CO];
#pragma unroll
    for (int c = 0; c < CO; ++c) acc[c] = 0.f;
    const __nv_bfloat16* xin = P.x + (size_t)n * P.H * P.W * 4;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = 2 * ho + ky - 1;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = 2 * wo + kx - 1;
        if (iy < 0 || iy >= P.H || ix < 0 || ix >= P.W) continue;
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(xin + ((size_t)iy * P.W + ix) * 4));
        const float xi[3] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y)};
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const float4* wrow = reinterpret_cast<const float4*>(s_w + ((ky * 3 + kx) * 3 + ci) * CO);
#pragma unroll
          for (int c4 = 0; c4 < CO / 4; ++c4) {
            const float4 w4 = wrow[c4];
            acc[c4 * 4 + 0] = fmaf(xi[ci], w4.x, acc[c4 * 4 + 0]);
            acc[c4 * 4 + 1] = fmaf(xi[ci], w4.y, acc[c4 * 4 + 1]);
            acc[c4 * 4 + 2] = fmaf(xi[ci], w4.z, acc[c4 * 4 + 2]);
            acc[c4 * 4 + 3] = fmaf(xi[ci], w4.w, acc[c4 * 4 + 3]);
          }
        }
      }
    }
    uint4* yo = reinterpret_cast<uint4*>(P.y + (size_t)pix * CO);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 o = make_uint4(pack_bf16x2(acc[q * 8 + 0], acc[q * 8 + 1]), pack_bf16x2(acc[q * 8 + 2], acc[q * 8 + 3]),
                                 pack_bf16x2(acc[q * 8 + 4], acc[q * 8 + 5]), pack_bf16x2(acc[q * 8 + 6], acc[q * 8 + 7]));
      yo[q] = o;
      const float r[8] = {bf16_lo(o.x), bf16_hi(o.x), bf16_lo(o.y), bf16_hi(o.y),
                          bf16_lo(o.z), bf16_hi(o.z), bf16_lo(o.w), bf16_hi(o.w)};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s0[q * 8 + j] += r[j];
        s1[q * 8 + j] = fmaf(r[j], r[j], s1[q * 8 + j]);
      }
    }
  }
  if (P.stats != nullptr) {
#pragma unroll
    for (int c = 0; c < CO; ++c) {
      const float a = warp_sum(s0[c]), b = warp_sum(s1[c]);
      if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_red[c], a);
        atomicAdd(&s_red[CO + c], b);
      }
    }
    __syncthreads();
    if (threadIdx.x < CO) {
      atomicAdd(P.stats + threadIdx.x, s_red[threadIdx.x]);
      atomicAdd(P.stats + P.stats_stride + threadIdx.x, s_red[CO + threadIdx.x]);
    }
  }
}

struct StemWgradParams {
  const __nv_bfloat16* x;    // [B,H,W,4]
  const __nv_bfloat16* dz;   // [B,Ho,Wo,32] gradient w.r.t. the stem BatchNorm output (ReLU6-masked)
  const __nv_bfloat16* yraw; // [B,Ho,Wo,32] raw stem conv output
  const float* coef;         // [3][32] ca, cb, cc
  float* dw;                 // [27][32] accumulated
  int B, H, W, Ho, Wo;
};

constexpr int WG_WARPS = 4;
constexpr int XS_LD = 28;    // 27 taps + 1 zero pad -> float4 loads
constexpr int DY_LD = 36;    // 32 channels + 4 pad: conflict-light float4 rows

__global__ void __launch_bounds__(WG_WARPS * 32)
stem_wgrad_kernel(const StemWgradParams P) {
  pdl_trigger();
  pdl_wait();
  __shared__ __align__(16) float s_x[WG_WARPS][32][XS_LD];
  __shared__ __align__(16) float s_dy[WG_WARPS][32][DY_LD];
  __shared__ float s_dw[28 * CO];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 28 * CO; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  float ca[CO], cb[CO], cc[CO];
#pragma unroll
  for (int c = 0; c < CO; ++c) { ca[c] = P.coef[c]; cb[c] = P.coef[CO + c]; cc[c] = P.coef[2 * CO + c]; }
  const int kg = lane % 7, cg = lane / 7;          // lanes 0..27: taps 4kg..4kg+3, channels 8cg..8cg+7
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  const long long npix = (long long)P.B * P.Ho * P.Wo;
  const long long nchunks = (npix + 31) / 32;
  for (long long ch = (long long)blockIdx.x * WG_WARPS + warp; ch < nchunks; ch += (long long)gridDim.x * WG_WARPS) {
    const long long pix = ch * 32 + lane;
    const bool ok = pix < npix;
    // ---- stage this lane's pixel: 27 inputs and 32 transformed output gradients
    float xv[XS_LD];
#pragma unroll
    for (int k = 0; k < XS_LD; ++k) xv[k] = 0.f;
    if (ok) {
      const int wo = (int)(pix % P.Wo);
      const int ho = (int)((pix / P.Wo) % P.Ho);
      const int n = (int)(pix / ((long long)P.Wo * P.Ho));
      const __nv_bfloat16* xin = P.x + (size_t)n * P.H * P.W * 4;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int iy = 2 * ho + ky - 1, ix = 2 * wo + kx - 1;
          if (iy >= 0 && iy < P.H && ix >= 0 && ix < P.W) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(xin + ((size_t)iy * P.W + ix) * 4));
            xv[(ky * 3 + kx) * 3 + 0] = bf16_lo(v.x);
            xv[(ky * 3 + kx) * 3 + 1] = bf16_hi(v.x);
            xv[(ky * 3 + kx) * 3 + 2] = bf16_lo(v.y);
          }
        }
    }
#pragma unroll
    for (int k4 = 0; k4 < XS_LD / 4; ++k4)
      *reinterpret_cast<float4*>(&s_x[warp][lane][k4 * 4]) = make_float4(xv[k4 * 4], xv[k4 * 4 + 1], xv[k4 * 4 + 2], xv[k4 * 4 + 3]);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float d[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = 0.f;
      if (ok) {
        const uint4 g = ldg_stream(P.dz + (size_t)pix * CO + q * 8);
        const uint4 y = ldg_stream(P.yraw + (size_t)pix * CO + q * 8);
        const float gf[8] = {bf16_lo(g.x), bf16_hi(g.x), bf16_lo(g.y), bf16_hi(g.y), bf16_lo(g.z), bf16_hi(g.z), bf16_lo(g.w), bf16_hi(g.w)};
        const float yf[8] = {bf16_lo(y.x), bf16_hi(y.x), bf16_lo(y.y), bf16_hi(y.y), bf16_lo(y.z), bf16_hi(y.z), bf16_lo(y.w), bf16_hi(y.w)};
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = fmaf(ca[q * 8 + j], gf[j], fmaf(cb[q * 8 + j], yf[j], cc[q * 8 + j]));
      }
      *reinterpret_cast<float4*>(&s_dy[warp][lane][q * 8]) = make_float4(d[0], d[1], d[2], d[3]);
      *reinterpret_cast<float4*>(&s_dy[warp][lane][q * 8 + 4]) = make_float4(d[4], d[5], d[6], d[7]);
    }
    __syncwarp();
    if (lane < 28) {
#pragma unroll 4
      for (int p = 0; p < 32; ++p) {
        const float4 xk = *reinterpret_cast<const float4*>(&s_x[warp][p][kg * 4]);
        const float4 d0 = *reinterpret_cast<const float4*>(&s_dy[warp][p][cg * 8]);
        const float4 d1 = *reinterpret_cast<const float4*>(&s_dy[warp][p][cg * 8 + 4]);
        const float xs[4] = {xk.x, xk.y, xk.z, xk.w};
        const float ds[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(xs[i], ds[j], acc[i][j]);
      }
    }
    __syncwarp();
  }
  if (lane < 28) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&s_dw[(kg * 4 + i) * CO + cg * 8 + j], acc[i][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < KT * CO; i += blockDim.x) atomicAdd(P.dw + i, s_dw[i]);
}


// ---------------------------------------------------------------------------------------------------------------
// The stem as a GEMM.  K = 27 is too small for the tensor cores only as long as one row is one pixel: with the
// row packing of pw_gemm / pw_wgrad (r pixels per row against a block-diagonal weight matrix) a [M, 32] im2col
// matrix is a perfectly good operand, and both the convolution and its weight gradient become the same tcgen05
// kernels every pointwise layer uses.  What remains here is the im2col pass itself (51 MB in, 103 MB out), the
// weight layout change and folding the [32][32] gradient back into the [27][32] arena slot.
struct StemColParams {
  const __nv_bfloat16* x;   // [B,H,W,4]
  __nv_bfloat16* col;       // [B*Ho*Wo][32]: k = (ky*3 + kx)*3 + ci for k < 27, zero above
  float* asum;              // [32] column sums over all rows (for the factored weight gradient), or null
  int B, H, W, Ho, Wo;
};

__global__ void __launch_bounds__(256)
stem_im2col_kernel(const StemColParams P) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_sum[32];
  if (threadIdx.x < 32) s_sum[threadIdx.x] = 0.f;
  __syncthreads();
  float cs[KT];
#pragma unroll
  for (int k = 0; k < KT; ++k) cs[k] = 0.f;
  const long long npix = (long long)P.B * P.Ho * P.Wo;
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < npix;
       pix += (long long)gridDim.x * blockDim.x) {
    const int wo = (int)(pix % P.Wo);
    const int ho = (int)((pix / P.Wo) % P.Ho);
    const int n = (int)(pix / ((long long)P.Wo * P.Ho));
    const __nv_bfloat16* xin = P.x + (size_t)n * P.H * P.W * 4;
    uint32_t h[32];              // one bf16 per k in the low half
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int iy = 2 * ho + t / 3 - 1, ix = 2 * wo + t % 3 - 1;
      uint2 v = make_uint2(0u, 0u);
      if (iy >= 0 && iy < P.H && ix >= 0 && ix < P.W) v = __ldg(reinterpret_cast<const uint2*>(xin + ((size_t)iy * P.W + ix) * 4));
      h[3 * t + 0] = v.x & 0xffffu;
      h[3 * t + 1] = v.x >> 16;
      h[3 * t + 2] = v.y & 0xffffu;
    }
#pragma unroll
    for (int k = KT; k < 32; ++k) h[k] = 0u;
    uint4* dst = reinterpret_cast<uint4*>(P.col + (size_t)pix * 32);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      dst[q] = make_uint4(h[8 * q + 0] | (h[8 * q + 1] << 16), h[8 * q + 2] | (h[8 * q + 3] << 16),
                          h[8 * q + 4] | (h[8 * q + 5] << 16), h[8 * q + 6] | (h[8 * q + 7] << 16));
    if (P.asum != nullptr) {
#pragma unroll
      for (int k = 0; k < KT; ++k) cs[k] += __uint_as_float(h[k] << 16);
    }
  }
  if (P.asum != nullptr) {
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      const float a = warp_sum(cs[k]);
      if ((threadIdx.x & 31) == 0) atomicAdd(&s_sum[k], a);
    }
    __syncthreads();
    if (threadIdx.x < KT) atomicAdd(P.asum + threadIdx.x, s_sum[threadIdx.x]);
  }
}

// fp32 arena weights [27][32] (k, cout) -> bf16 GEMM weights: plain [32 cout][32 k] at wout, and the block-diagonal
// [32 r][32 r] copy at wout + 1024 when r > 1 (k >= 27 and off-diagonal blocks stay zero from allocation)
__global__ void stem_weight_prep_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wout, int r) {
  pdl_trigger();
  pdl_wait();
  for (int i = threadIdx.x; i < KT * CO; i += blockDim.x) {
    const int k = i / CO, n = i - k * CO;
    const __nv_bfloat16 v = __float2bfloat16(w[i]);
    wout[n * 32 + k] = v;
    for (int p = 0; p < r && r > 1; ++p) wout[1024 + (p * 32 + n) * (32 * r) + p * 32 + k] = v;
  }
}

// grads[k][n] += scratch[n][k] for k < 27 (the GEMM produces dW as [cout][k])
__global__ void stem_dw_fold_kernel(const float* __restrict__ scratch, float* __restrict__ dw) {
  pdl_trigger();
  pdl_wait();
  for (int i = threadIdx.x; i < KT * CO; i += blockDim.x) {
    const int k = i / CO, n = i - k * CO;
    dw[i] += scratch[n * 32 + k];
  }
}

}  // namespace

int stem_fwd_launch(const void* x, const float* w, void* y, float* stats, int stats_stride, int B, int H, int W,
                    cudaStream_t s) {
  StemFwdParams P;
  P.x = static_cast<const __nv_bfloat16*>(x); P.w = w; P.y = static_cast<__nv_bfloat16*>(y);
  P.stats = stats; P.stats_stride = stats_stride; P.B = B; P.H = H; P.W = W;
  P.Ho = (H - 1) / 2 + 1; P.Wo = (W - 1) / 2 + 1;
  const long long npix = (long long)B * P.Ho * P.Wo;
  long long blocks = (npix + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  B200DP_CUDA(launch_k(stem_fwd_kernel, dim3((unsigned)blocks), dim3(256), 0, s, P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int stem_wgrad_launch(const void* x, const void* dz, const void* yraw, const float* coef, float* dw, int B, int H,
                      int W, cudaStream_t s) {
  StemWgradParams P;
  P.x = static_cast<const __nv_bfloat16*>(x); P.dz = static_cast<const __nv_bfloat16*>(dz);
  P.yraw = static_cast<const __nv_bfloat16*>(yraw); P.coef = coef; P.dw = dw;
  P.B = B; P.H = H; P.W = W; P.Ho = (H - 1) / 2 + 1; P.Wo = (W - 1) / 2 + 1;
  const long long nchunks = ((long long)B * P.Ho * P.Wo + 31) / 32;
  long long blocks = (nchunks + WG_WARPS - 1) / WG_WARPS;
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  B200DP_CUDA(launch_k(stem_wgrad_kernel, dim3((unsigned)blocks), dim3(WG_WARPS * 32), 0, s, P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_stem_fwd(const void* x, const float* w, void* y, float* stats, int stats_stride, int B, int H,
                               int W, void* stream) {
  return stem_fwd_launch(x, w, y, stats, stats_stride, B, H, W, static_cast<cudaStream_t>(stream));
}

extern "C" int b200dp_stem_wgrad(const void* x, const void* dz, const void* yraw, const float* coef, float* dw,
                                 int B, int H, int W, void* stream) {
  return stem_wgrad_launch(x, dz, yraw, coef, dw, B, H, W, static_cast<cudaStream_t>(stream));
}

namespace b200dp {

int stem_im2col_launch(const void* x, void* col, float* asum, int B, int H, int W, cudaStream_t s) {
  StemColParams P;
  P.x = static_cast<const __nv_bfloat16*>(x); P.col = static_cast<__nv_bfloat16*>(col); P.asum = asum;
  P.B = B; P.H = H; P.W = W; P.Ho = (H - 1) / 2 + 1; P.Wo = (W - 1) / 2 + 1;
  const long long npix = (long long)B * P.Ho * P.Wo;
  long long blocks = (npix + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  B200DP_CUDA(launch_k(stem_im2col_kernel, dim3((unsigned)blocks), dim3(256), 0, s, P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int stem_weight_prep_launch(const float* w, void* wout, int pack, cudaStream_t s) {
  B200DP_CUDA(launch_k(stem_weight_prep_kernel, dim3(1), dim3(256), 0, s, w, static_cast<__nv_bfloat16*>(wout), pack));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int stem_dw_fold_launch(const float* scratch, float* dw, cudaStream_t s) {
  B200DP_CUDA(launch_k(stem_dw_fold_kernel, dim3(1), dim3(256), 0, s, scratch, dw));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

extern "C" int b200dp_stem_im2col(const void* x, void* col, float* asum, int B, int H, int W, void* stream) {
  return b200dp::stem_im2col_launch(x, col, asum, B, H, W, static_cast<cudaStream_t>(stream));
}
