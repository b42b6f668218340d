// Depthwise 3x3 convolution, NHWC bf16, fused with the surrounding BatchNorm / ReLU6 (K2, K4, K5).
//
// Forward : y_raw = dwconv3x3( relu6(scale * x_raw + shift) )      + per-channel sum(y), sum(y^2)
//           The producer's BatchNorm + ReLU6 are applied while loading (the raw expand-conv output is
//           read exactly once), and the statistics of THIS conv's BatchNorm come out of the same pass.
// Backward: one pass computes, per input pixel,
//             dd   = ca*dz_d + cb*d_raw + cc                       (BatchNorm-backward of the dw output, on load)
//             da   = dwconv3x3_transposed(dd)                      (data gradient)
//             dz_e = da * [0 < scale*e_raw+shift < 6]              (ReLU6 backward of the dw input) -> stored
//             sum(dz_e), sum(dz_e*e_raw)                           (statistics for the expand BatchNorm backward)
//             dW[ky][kx] += relu6(scale*e_raw+shift) * dd          (weight gradient)
// Reference behaviour replaced: cuDNN/ATen depthwise conv fwd + dgrad + wgrad, cuDNN batch-norm fwd/bwd and
// hardtanh fwd/bwd as separate full passes over the expanded tensors (torchvision mobilenetv2.py ConvBNReLU;
// cifar10_mpi_mobilenet_224.py:177-179).
//
// Mapping: a thread owns 4 channels (8-byte vectors) and sweeps TH=2 rows along W with a 3-column sliding
// window in registers (each input vector is loaded once per thread); consecutive threads take consecutive
// channel groups, so a warp reads contiguous NHWC bytes.  Per-channel sums are reduced thread -> shared -> global.
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int TH = 2;
constexpr int VEC = 4;

struct DwFwdParams {
  const __nv_bfloat16* x;
  const float* scale;
  const float* shift;
  const float* w;        // [9][C] fp32 (tap-major)
  __nv_bfloat16* y;
  float* stats;          // [2][stats_stride]; may be null (inference)
  int B, H, W, C, Ho, Wo, nseg, seglen, stats_stride;
};

__device__ __forceinline__ void unpack4(const uint2& v, float (&f)[4]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
}
__device__ __forceinline__ uint2 ldg8(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint2*>(p)); }

template <int S>
__global__ void __launch_bounds__(256)
dw_fwd_kernel(const DwFwdParams P) {
  constexpr int R = (TH - 1) * S + 3;
  extern __shared__ float s_red[];                 // [2][C]
  const int CG = P.C / VEC;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  const int c0 = cg * VEC;
  for (int i = threadIdx.x; i < 2 * P.C; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  float s0[VEC] = {0.f, 0.f, 0.f, 0.f}, s1[VEC] = {0.f, 0.f, 0.f, 0.f};
  if (ti < ipb) {
    float wr[9][VEC], sc[VEC], sh[VEC];
#pragma unroll
    for (int k = 0; k < 9; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) wr[k][j] = P.w[k * P.C + c0 + j];
#pragma unroll
    for (int j = 0; j < VEC; ++j) { sc[j] = P.scale[c0 + j]; sh[j] = P.shift[c0 + j]; }
    const int rblocks = (P.Ho + TH - 1) / TH;
    const long long items = (long long)P.B * rblocks * P.nseg;
    for (long long item = (long long)blockIdx.x * ipb + ti; item < items; item += (long long)gridDim.x * ipb) {
      const int seg = (int)(item % P.nseg);
      const int yb = (int)((item / P.nseg) % rblocks);
      const int n = (int)(item / ((long long)P.nseg * rblocks));
      const int y0 = yb * TH, xs = seg * P.seglen, xe = min(xs + P.seglen, P.Wo);
      const int iy0 = y0 * S - 1;
      const __nv_bfloat16* xin = P.x + (size_t)n * P.H * P.W * P.C + c0;
      __nv_bfloat16* yout = P.y + (size_t)n * P.Ho * P.Wo * P.C + c0;
      auto loadcol = [&](int xi, float (&col)[R][VEC]) {
        const bool xok = xi >= 0 && xi < P.W;
        uint2 raw[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int iy = iy0 + r;
          raw[r] = make_uint2(0u, 0u);
          if (xok && iy >= 0 && iy < P.H) raw[r] = ldg8(xin + ((size_t)iy * P.W + xi) * P.C);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const bool ok = xok && (iy0 + r) >= 0 && (iy0 + r) < P.H;
          float f[VEC];
          unpack4(raw[r], f);
#pragma unroll
          for (int j = 0; j < VEC; ++j) col[r][j] = ok ? relu6f(fmaf(f[j], sc[j], sh[j])) : 0.f;   // zero padding is post-activation
        }
      };
      float ca[R][VEC], cb[R][VEC], cc[R][VEC];
      if (S == 1) { loadcol(xs - 1, ca); loadcol(xs, cb); } else { loadcol(2 * xs - 1, ca); }
      for (int xo = xs; xo < xe; ++xo) {
        if (S == 1) { loadcol(xo + 1, cc); } else { loadcol(2 * xo, cb); loadcol(2 * xo + 1, cc); }
#pragma unroll
        for (int t = 0; t < TH; ++t) {
          if (y0 + t < P.Ho) {
            float acc[VEC] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
              for (int j = 0; j < VEC; ++j) {
                acc[j] = fmaf(wr[ky * 3 + 0][j], ca[t * S + ky][j], acc[j]);
                acc[j] = fmaf(wr[ky * 3 + 1][j], cb[t * S + ky][j], acc[j]);
                acc[j] = fmaf(wr[ky * 3 + 2][j], cc[t * S + ky][j], acc[j]);
              }
            const uint2 o = make_uint2(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]));
            *reinterpret_cast<uint2*>(yout + ((size_t)(y0 + t) * P.Wo + xo) * P.C) = o;
            float q[VEC];
            unpack4(o, q);      // statistics of the values the consumer will actually read
#pragma unroll
            for (int j = 0; j < VEC; ++j) { s0[j] += q[j]; s1[j] = fmaf(q[j], q[j], s1[j]); }
          }
        }
        if (S == 1) {
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int j = 0; j < VEC; ++j) { ca[r][j] = cb[r][j]; cb[r][j] = cc[r][j]; }
        } else {
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int j = 0; j < VEC; ++j) ca[r][j] = cc[r][j];
        }
      }
    }
    if (P.stats != nullptr) {
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        atomicAdd(&s_red[c0 + j], s0[j]);
        atomicAdd(&s_red[P.C + c0 + j], s1[j]);
      }
    }
  }
  __syncthreads();
  if (P.stats != nullptr)
    for (int i = threadIdx.x; i < P.C; i += blockDim.x) {
      atomicAdd(P.stats + i, s_red[i]);
      atomicAdd(P.stats + P.stats_stride + i, s_red[P.C + i]);
    }
}

struct DwBwdParams {
  const __nv_bfloat16* e_raw;   // [B,H,W,C] raw input-side activation (expand-conv / stem output)
  const float* escale;          // its BatchNorm scale / shift
  const float* eshift;
  const __nv_bfloat16* dzd;     // [B,Ho,Wo,C] gradient w.r.t. the dw BatchNorm output (after its ReLU6 mask)
  const __nv_bfloat16* d_raw;   // [B,Ho,Wo,C] raw dw-conv output
  const float* dcoef;           // [3][C] ca, cb, cc of the dw BatchNorm backward
  const float* w;               // [9][C] fp32
  __nv_bfloat16* dze;           // [B,H,W,C] out: gradient w.r.t. the input-side BatchNorm output (masked)
  float* stats;                 // [2][stats_stride]: sum dze, sum dze*e_raw
  float* dw;                    // [9][C] fp32 weight gradient (accumulated)
  int B, H, W, C, Ho, Wo, nseg, seglen, stats_stride;
};

template <int S>
__global__ void __launch_bounds__(256)
dw_bwd_kernel(const DwBwdParams P) {
  extern __shared__ float s_red[];                 // [2][C] stats + [9][C] dW
  const int CG = P.C / VEC;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  const int c0 = cg * VEC;
  for (int i = threadIdx.x; i < 11 * P.C; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  if (ti < ipb) {
    float wr[9][VEC], gw[9][VEC], sc[VEC], sh[VEC], ka[VEC], kb[VEC], kc[VEC];
    float s0[VEC] = {0.f, 0.f, 0.f, 0.f}, s1[VEC] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 9; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) { wr[k][j] = P.w[k * P.C + c0 + j]; gw[k][j] = 0.f; }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      sc[j] = P.escale[c0 + j]; sh[j] = P.eshift[c0 + j];
      ka[j] = P.dcoef[c0 + j]; kb[j] = P.dcoef[P.C + c0 + j]; kc[j] = P.dcoef[2 * P.C + c0 + j];
    }
    const int rblocks = (P.H + TH - 1) / TH;
    const long long items = (long long)P.B * rblocks * P.nseg;
    for (long long item = (long long)blockIdx.x * ipb + ti; item < items; item += (long long)gridDim.x * ipb) {
      const int seg = (int)(item % P.nseg);
      const int hb = (int)((item / P.nseg) % rblocks);
      const int n = (int)(item / ((long long)P.nseg * rblocks));
      const int h0 = hb * TH, ws = seg * P.seglen, we = min(ws + P.seglen, P.W);
      const __nv_bfloat16* ein = P.e_raw + (size_t)n * P.H * P.W * P.C + c0;
      __nv_bfloat16* dout = P.dze + (size_t)n * P.H * P.W * P.C + c0;
      const __nv_bfloat16* gz = P.dzd + (size_t)n * P.Ho * P.Wo * P.C + c0;
      const __nv_bfloat16* gr = P.d_raw + (size_t)n * P.Ho * P.Wo * P.C + c0;
      // dd(ho, wo) = ca*dz_d + cb*d_raw + cc inside the output grid, 0 outside
      auto dd_at = [&](int ho, int wo, float (&d)[VEC]) {
        const bool ok = ho >= 0 && ho < P.Ho && wo >= 0 && wo < P.Wo;
        uint2 a = make_uint2(0u, 0u), b = make_uint2(0u, 0u);
        if (ok) {
          const size_t off = ((size_t)ho * P.Wo + wo) * P.C;
          a = ldg8(gz + off);
          b = ldg8(gr + off);
        }
        float fa[VEC], fb[VEC];
        unpack4(a, fa);
        unpack4(b, fb);
#pragma unroll
        for (int j = 0; j < VEC; ++j) d[j] = ok ? fmaf(ka[j], fa[j], fmaf(kb[j], fb[j], kc[j])) : 0.f;
      };
      // finish one input pixel: mask, store, statistics, weight-gradient contribution is done by the caller
      auto finish = [&](int h, int w, const float (&da)[VEC], const float (&e)[VEC], const bool (&keep)[VEC]) {
        float o[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) o[j] = keep[j] ? da[j] : 0.f;
        const uint2 pk = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
        *reinterpret_cast<uint2*>(dout + ((size_t)h * P.W + w) * P.C) = pk;
        float q[VEC];
        unpack4(pk, q);
#pragma unroll
        for (int j = 0; j < VEC; ++j) { s0[j] += q[j]; s1[j] = fmaf(q[j], e[j], s1[j]); }
      };
      if (S == 1) {
        // sliding window over output columns wo = w-1, w, w+1; rows ho = h0-1 .. h0+TH
        constexpr int R = TH + 2;
        float dp[R][VEC], dc[R][VEC], dn[R][VEC];
#pragma unroll
        for (int r = 0; r < R; ++r) { dd_at(h0 - 1 + r, ws - 1, dp[r]); dd_at(h0 - 1 + r, ws, dc[r]); }
        for (int w = ws; w < we; ++w) {
#pragma unroll
          for (int r = 0; r < R; ++r) dd_at(h0 - 1 + r, w + 1, dn[r]);
#pragma unroll
          for (int t = 0; t < TH; ++t) {
            const int h = h0 + t;
            if (h < P.H) {
              float e[VEC], a[VEC], da[VEC] = {0.f, 0.f, 0.f, 0.f};
              bool keep[VEC];
              unpack4(ldg8(ein + ((size_t)h * P.W + w) * P.C), e);
#pragma unroll
              for (int j = 0; j < VEC; ++j) {
                const float z = fmaf(e[j], sc[j], sh[j]);
                keep[j] = z > 0.f && z < 6.f;
                a[j] = relu6f(z);
              }
#pragma unroll
              for (int ky = 0; ky < 3; ++ky) {
                const int r = t + 2 - ky;        // ho = h + 1 - ky  ->  row index ho - (h0 - 1)
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                  da[j] = fmaf(dn[r][j], wr[ky * 3 + 0][j], da[j]);   // kx = 0 -> wo = w + 1
                  da[j] = fmaf(dc[r][j], wr[ky * 3 + 1][j], da[j]);   // kx = 1 -> wo = w
                  da[j] = fmaf(dp[r][j], wr[ky * 3 + 2][j], da[j]);   // kx = 2 -> wo = w - 1
                  gw[ky * 3 + 0][j] = fmaf(a[j], dn[r][j], gw[ky * 3 + 0][j]);
                  gw[ky * 3 + 1][j] = fmaf(a[j], dc[r][j], gw[ky * 3 + 1][j]);
                  gw[ky * 3 + 2][j] = fmaf(a[j], dp[r][j], gw[ky * 3 + 2][j]);
                }
              }
              finish(h, w, da, e, keep);
            }
          }
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int j = 0; j < VEC; ++j) { dp[r][j] = dc[r][j]; dc[r][j] = dn[r][j]; }
        }
      } else {
        // stride 2: gather the (at most 4) output positions that read this input pixel
        for (int w = ws; w < we; ++w) {
#pragma unroll
          for (int t = 0; t < TH; ++t) {
            const int h = h0 + t;
            if (h < P.H) {
              float e[VEC], a[VEC], da[VEC] = {0.f, 0.f, 0.f, 0.f};
              bool keep[VEC];
              unpack4(ldg8(ein + ((size_t)h * P.W + w) * P.C), e);
#pragma unroll
              for (int j = 0; j < VEC; ++j) {
                const float z = fmaf(e[j], sc[j], sh[j]);
                keep[j] = z > 0.f && z < 6.f;
                a[j] = relu6f(z);
              }
#pragma unroll
              for (int ky = 0; ky < 3; ++ky) {
                const int hy = h + 1 - ky;
                if (hy < 0 || (hy & 1)) continue;
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                  const int wx = w + 1 - kx;
                  if (wx < 0 || (wx & 1)) continue;
                  float d[VEC];
                  dd_at(hy >> 1, wx >> 1, d);
#pragma unroll
                  for (int j = 0; j < VEC; ++j) {
                    da[j] = fmaf(d[j], wr[ky * 3 + kx][j], da[j]);
                    gw[ky * 3 + kx][j] = fmaf(a[j], d[j], gw[ky * 3 + kx][j]);
                  }
                }
              }
              finish(h, w, da, e, keep);
            }
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      atomicAdd(&s_red[c0 + j], s0[j]);
      atomicAdd(&s_red[P.C + c0 + j], s1[j]);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) atomicAdd(&s_red[(2 + k) * P.C + c0 + j], gw[k][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < P.C; i += blockDim.x) {
    atomicAdd(P.stats + i, s_red[i]);
    atomicAdd(P.stats + P.stats_stride + i, s_red[P.C + i]);
  }
  for (int i = threadIdx.x; i < 9 * P.C; i += blockDim.x) atomicAdd(P.dw + i, s_red[2 * P.C + i]);
}

void pick_segments(int rows, int width, int B, int C, int* nseg, int* seglen) {
  // enough independent strips to fill the machine (~300k threads), segments no shorter than 7 pixels
  const long long base = (long long)B * ((rows + TH - 1) / TH) * (C / VEC);
  int n = 1;
  while (base * n < 300000 && width / (n * 2) >= 7) n *= 2;
  *nseg = n;
  *seglen = (width + n - 1) / n;
}

}  // namespace

int dw_fwd_launch(const void* x, const float* scale, const float* shift, const float* w, void* y, float* stats,
                  int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s) {
  B200DP_REQUIRE(C % 16 == 0 && C <= 1024, "dw_fwd: C=%d unsupported", C);
  B200DP_REQUIRE(stride == 1 || stride == 2, "dw_fwd: stride must be 1 or 2");
  DwFwdParams P;
  P.x = static_cast<const __nv_bfloat16*>(x); P.scale = scale; P.shift = shift; P.w = w;
  P.y = static_cast<__nv_bfloat16*>(y); P.stats = stats; P.stats_stride = stats_stride;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.Ho = (H - 1) / stride + 1; P.Wo = (W - 1) / stride + 1;
  pick_segments(P.Ho, P.Wo, B, C, &P.nseg, &P.seglen);
  const int CG = C / VEC;
  const int threads = CG >= 256 ? 256 : (256 / CG) * CG;
  B200DP_REQUIRE(CG <= 256, "dw_fwd: C too large");
  const int ipb = threads / CG;
  const long long items = (long long)B * ((P.Ho + TH - 1) / TH) * P.nseg;
  long long blocks = (items + ipb - 1) / ipb;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const size_t smem = 2 * (size_t)C * sizeof(float);
  if (stride == 1) dw_fwd_kernel<1><<<(int)blocks, threads, smem, s>>>(P);
  else dw_fwd_kernel<2><<<(int)blocks, threads, smem, s>>>(P);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int dw_bwd_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                  const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                  int H, int W, int C, int stride, cudaStream_t s) {
  B200DP_REQUIRE(C % 16 == 0 && C <= 1024, "dw_bwd: C=%d unsupported", C);
  B200DP_REQUIRE(stride == 1 || stride == 2, "dw_bwd: stride must be 1 or 2");
  DwBwdParams P;
  P.e_raw = static_cast<const __nv_bfloat16*>(e_raw); P.escale = escale; P.eshift = eshift;
  P.dzd = static_cast<const __nv_bfloat16*>(dzd); P.d_raw = static_cast<const __nv_bfloat16*>(d_raw);
  P.dcoef = dcoef; P.w = w; P.dze = static_cast<__nv_bfloat16*>(dze); P.stats = stats; P.stats_stride = stats_stride;
  P.dw = dw; P.B = B; P.H = H; P.W = W; P.C = C;
  P.Ho = (H - 1) / stride + 1; P.Wo = (W - 1) / stride + 1;
  pick_segments(H, W, B, C, &P.nseg, &P.seglen);
  const int CG = C / VEC;
  B200DP_REQUIRE(CG <= 256, "dw_bwd: C too large");
  const int threads = (256 / CG) * CG;
  const int ipb = threads / CG;
  const long long items = (long long)B * ((H + TH - 1) / TH) * P.nseg;
  long long blocks = (items + ipb - 1) / ipb;
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  const size_t smem = 11 * (size_t)C * sizeof(float);
  if (stride == 1) dw_bwd_kernel<1><<<(int)blocks, threads, smem, s>>>(P);
  else dw_bwd_kernel<2><<<(int)blocks, threads, smem, s>>>(P);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_dw_fwd(const void* x, const float* scale, const float* shift, const float* w, void* y,
                             float* stats, int stats_stride, int B, int H, int W, int C, int stride, void* stream) {
  return dw_fwd_launch(x, scale, shift, w, y, stats, stats_stride, B, H, W, C, stride, static_cast<cudaStream_t>(stream));
}

extern "C" int b200dp_dw_bwd(const void* e_raw, const float* escale, const float* eshift, const void* dzd,
                             const void* d_raw, const float* dcoef, const float* w, void* dze, float* stats,
                             int stats_stride, float* dw, int B, int H, int W, int C, int stride, void* stream) {
  return dw_bwd_launch(e_raw, escale, eshift, dzd, d_raw, dcoef, w, dze, stats, stats_stride, dw, B, H, W, C, stride,
                       static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant: depthwise conv + the BatchNorm coefficients of its output, in one call.
//
// b200dp_dw_fwd accumulates the per-channel sums of its output into a buffer the caller owns, and the
// caller then runs b200dp_bn_fwd_finalize on it.  That buffer is pure scratch: nobody wants the sums, only
// the scale / shift they turn into.  Some callers cannot provide it, because they are not allowed to
// allocate device memory where they sit: a plug-in running inside another runtime's enqueue hook, a serving
// loop whose memory pool is sized up front, a latency-critical path that must never enter the allocator, or
// a process that has handed all free memory to a framework allocator.  For them
// the scratch lives in the library image itself: a file-scope __device__ array, 8 KB on every GPU that loads
// this library, held until the process exits.
//
// Rules that follow from the scratch being shared and static:
//   * at most B200DP_DW_NOALLOC_MAX_C channels; larger C is rejected with B200DP_EINVAL;
//   * the sums must start at zero, so every call clears the part it uses first (a memset node on `stream`);
//   * all calls share the one array per GPU: two calls that can overlap on the device (different streams,
//     no dependency between them) race on it and give wrong coefficients -- serialise them;
//   * the array's name is only a host-side handle.  Its device address is looked up once per device with
//     cudaGetSymbolAddress while that device is current.  The first call on a device does this look-up, so
//     make that first call outside any stream capture.
// No cudaMalloc / cudaFree / allocating library routine is called here or below.
namespace b200dp {
namespace {

constexpr int kMaxDevices = 64;
static __device__ __align__(256) float g_dw_noalloc_sums[2 * B200DP_DW_NOALLOC_MAX_C];
float* g_dw_noalloc_ptr[kMaxDevices] = {nullptr};

int dw_noalloc_scratch(float** out) {
  int dev = -1;
  B200DP_CUDA(cudaGetDevice(&dev));
  B200DP_REQUIRE(dev >= 0 && dev < kMaxDevices, "dw_fwd_bn_noalloc: device ordinal %d unsupported", dev);
  if (g_dw_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    B200DP_CUDA(cudaGetSymbolAddress(&p, g_dw_noalloc_sums));   // device address of the symbol ON THIS DEVICE
    B200DP_REQUIRE(p != nullptr, "dw_fwd_bn_noalloc: cudaGetSymbolAddress returned a null address");
    g_dw_noalloc_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_dw_noalloc_ptr[dev];
  return B200DP_OK;
}

}  // namespace

int bn_fwd_finalize_launch(const float* sums, int stride, const float* gamma, const float* beta, float count,
                           float eps, float momentum, float* rmean, float* rvar, float* coef, float* saved, int C,
                           cudaStream_t s);

}  // namespace b200dp

extern "C" int b200dp_dw_fwd_bn_noalloc(const void* x, const float* in_scale, const float* in_shift, const float* w,
                                        const float* gamma, const float* beta, float eps, void* y, float* coef_out,
                                        float* saved_out, int B, int H, int W, int C, int stride, void* stream) {
  B200DP_REQUIRE(x && in_scale && in_shift && w && gamma && beta && y && coef_out && saved_out,
                 "dw_fwd_bn_noalloc: null pointer argument");
  B200DP_REQUIRE(C > 0 && C <= B200DP_DW_NOALLOC_MAX_C,
                 "dw_fwd_bn_noalloc: C=%d exceeds the static scratch (at most %d channels); use b200dp_dw_fwd with "
                 "your own sums buffer", C, B200DP_DW_NOALLOC_MAX_C);
  B200DP_REQUIRE(B > 0 && H > 0 && W > 0, "dw_fwd_bn_noalloc: empty tensor");
  float* sums = nullptr;
  if (int rc = dw_noalloc_scratch(&sums)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the previous call's sums are still in the array: clear the 2*C entries this call accumulates into
  B200DP_CUDA(cudaMemsetAsync(sums, 0, 2 * (size_t)C * sizeof(float), s));
  if (int rc = dw_fwd_launch(x, in_scale, in_shift, w, y, sums, C, B, H, W, C, stride, s)) return rc;
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  return bn_fwd_finalize_launch(sums, C, gamma, beta, (float)B * Ho * Wo, eps, 0.f, nullptr, nullptr, coef_out,
                                saved_out, C, s);
}
