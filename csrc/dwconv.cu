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
#include "bn_inline.cuh"
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int TH = 2;
constexpr int VEC = 2;     // channels per thread: one bf16x2 word, a warp covers 64 contiguous channels (128 B)

struct DwFwdParams {
  const __nv_bfloat16* x;
  const float* scale;
  const float* shift;
  const float* w;        // [9][C] fp32 (tap-major)
  __nv_bfloat16* y;
  float* stats;          // [2][stats_stride]; may be null (inference)
  int B, H, W, C, Ho, Wo, nseg, seglen, stats_stride, cgb;   // cgb: channel pairs handled by one block
};

__device__ __forceinline__ uint32_t ldg4(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint32_t*>(p)); }

// Forward.  Per output column a thread loads S new input columns of R rows (4-byte words), already issued one
// iteration ahead (register prefetch), so global latency overlaps the FMAs of the current column.
template <int S>
__global__ void __launch_bounds__(256, 2)
dw_fwd_kernel(const DwFwdParams P, const BnFwdSrc bsrc) {
  pdl_trigger();
  pdl_wait();
  constexpr int R = (TH - 1) * S + 3;
  extern __shared__ float s_red[];                 // [2][C]
  const int CG = P.cgb;                            // channel pairs of this block: [blockIdx.y * cgb, +cgb)
  const int ipb = blockDim.x / CG;
  const int cg = blockIdx.y * CG + threadIdx.x % CG, ti = threadIdx.x / CG;
  const int c0 = cg * VEC;
  const int cb0 = blockIdx.y * CG * VEC, cbn = CG * VEC;   // channel range flushed by this block
  for (int i = threadIdx.x; i < 2 * P.C; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;
  if (ti < ipb) {
    float wa[9], wb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) { wa[k] = P.w[k * P.C + c0]; wb[k] = P.w[k * P.C + c0 + 1]; }
    float sca, scb, sha, shb;
    if (bsrc.sums != nullptr) {     // producer's BatchNorm finalised here; (block row 0, first strip slot) publishes it
      const bool wr = blockIdx.x == 0 && ti == 0;
      bn_fwd_channel(bsrc, c0, wr, sca, sha);
      bn_fwd_channel(bsrc, c0 + 1, wr, scb, shb);
    } else {
      sca = P.scale[c0]; scb = P.scale[c0 + 1]; sha = P.shift[c0]; shb = P.shift[c0 + 1];
    }
    const int rblocks = (P.Ho + TH - 1) / TH;
    const long long items = (long long)P.B * rblocks * P.nseg;
    for (long long item = (long long)blockIdx.x * ipb + ti; item < items; item += (long long)gridDim.x * ipb) {
      const int seg = (int)(item % P.nseg);
      const int yb = (int)((item / P.nseg) % rblocks);
      const int n = (int)(item / ((long long)P.nseg * rblocks));
      const int y0 = yb * TH, xs = seg * P.seglen, xe = min(xs + P.seglen, P.Wo);
      if (xs >= xe) continue;
      const int iy0 = y0 * S - 1;
      const __nv_bfloat16* xin = P.x + (size_t)n * P.H * P.W * P.C + c0;
      __nv_bfloat16* yout = P.y + (size_t)n * P.Ho * P.Wo * P.C + c0;
      // row pointers and 0/1 row masks are loop invariants of the sweep (out-of-image rows read row 0 and are
      // multiplied by 0): the column loop is left with one IMAD per load and no per-element predicates
      const __nv_bfloat16* rp[R];
      float rm[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int iy = iy0 + r;
        const bool ok = iy >= 0 && iy < P.H;
        rp[r] = xin + (size_t)(ok ? iy : 0) * P.W * P.C;
        rm[r] = ok ? 1.f : 0.f;
      }
      auto fetch = [&](int xi, uint32_t (&raw)[R]) {
        const int xc = min(max(xi, 0), P.W - 1);
        const size_t co = (size_t)xc * P.C;
#pragma unroll
        for (int r = 0; r < R; ++r) raw[r] = ldg4(rp[r] + co);
      };
      // zero padding is post-activation: out-of-image taps contribute 0, not relu6(shift)
      auto xform = [&](int xi, const uint32_t (&raw)[R], float (&a)[R], float (&b)[R]) {
        const float cm = (xi >= 0 && xi < P.W) ? 1.f : 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float m = rm[r] * cm;
          a[r] = relu6f(fmaf(bf16_lo(raw[r]), sca, sha)) * m;
          b[r] = relu6f(fmaf(bf16_hi(raw[r]), scb, shb)) * m;
        }
      };
      __nv_bfloat16* yrow[TH];
#pragma unroll
      for (int t = 0; t < TH; ++t) yrow[t] = yout + (size_t)min(y0 + t, P.Ho - 1) * P.Wo * P.C;
      float pa[R], pb[R], ca[R], cb[R], na[R], nb[R];     // previous / centre / next input column (2 channels)
      uint32_t n0[R], n1[R];
      if (S == 1) {
        fetch(xs - 1, n0); xform(xs - 1, n0, pa, pb);
        fetch(xs, n0);     xform(xs, n0, ca, cb);
        fetch(xs + 1, n0);
      } else {
        fetch(2 * xs - 1, n0); xform(2 * xs - 1, n0, pa, pb);
        fetch(2 * xs, n0);
        fetch(2 * xs + 1, n1);
      }
      for (int xo = xs; xo < xe; ++xo) {
        if (S == 1) {
          xform(xo + 1, n0, na, nb);
          if (xo + 1 < xe) fetch(xo + 2, n0);
        } else {
          xform(2 * xo, n0, ca, cb);
          xform(2 * xo + 1, n1, na, nb);
          if (xo + 1 < xe) { fetch(2 * xo + 2, n0); fetch(2 * xo + 3, n1); }
        }
#pragma unroll
        for (int t = 0; t < TH; ++t) {
          if (y0 + t < P.Ho) {
            float acca = 0.f, accb = 0.f;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
              const int r = t * S + ky;
              acca = fmaf(wa[ky * 3 + 0], pa[r], acca); accb = fmaf(wb[ky * 3 + 0], pb[r], accb);
              acca = fmaf(wa[ky * 3 + 1], ca[r], acca); accb = fmaf(wb[ky * 3 + 1], cb[r], accb);
              acca = fmaf(wa[ky * 3 + 2], na[r], acca); accb = fmaf(wb[ky * 3 + 2], nb[r], accb);
            }
            const uint32_t o = pack_bf16x2(acca, accb);
            *reinterpret_cast<uint32_t*>(yrow[t] + (size_t)xo * P.C) = o;
            const float qa = bf16_lo(o), qb = bf16_hi(o);   // statistics of the values the consumer will read
            s0a += qa; s0b += qb;
            s1a = fmaf(qa, qa, s1a); s1b = fmaf(qb, qb, s1b);
          }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if (S == 1) { pa[r] = ca[r]; pb[r] = cb[r]; ca[r] = na[r]; cb[r] = nb[r]; }
          else { pa[r] = na[r]; pb[r] = nb[r]; }
        }
      }
    }
    if (P.stats != nullptr) {
      atomicAdd(&s_red[c0], s0a); atomicAdd(&s_red[c0 + 1], s0b);
      atomicAdd(&s_red[P.C + c0], s1a); atomicAdd(&s_red[P.C + c0 + 1], s1b);
    }
  }
  __syncthreads();
  if (P.stats != nullptr)
    for (int i = cb0 + threadIdx.x; i < cb0 + cbn; i += blockDim.x) {
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
  int B, H, W, C, Ho, Wo, nseg, seglen, stats_stride, cgb;
};

// Backward.  Stride 1: sliding window over the three output columns wo = w-1, w, w+1 that read input column w.
// Stride 2: input pixels are handled as 2x2 blocks (rows 2r, 2r+1; columns 2c, 2c+1), which read exactly the
// output positions (r..r+1, c..c+1) -- nine (tap, position) pairs per block, column c+1 carried to the next block.
// Either way each dd = ca*dz_d + cb*d_raw + cc is formed once per thread and raw words are prefetched one
// iteration ahead.
template <int S>
__global__ void __launch_bounds__(256, 2)
dw_bwd_kernel(const DwBwdParams P, const BnBwdSrc bsrc) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float s_red[];                 // [2][C] stats + [9][C] dW
  const int CG = P.cgb;                            // channel pairs of this block: [blockIdx.y * cgb, +cgb)
  const int ipb = blockDim.x / CG;
  const int cg = blockIdx.y * CG + threadIdx.x % CG, ti = threadIdx.x / CG;
  const int c0 = cg * VEC;
  const int cb0 = blockIdx.y * CG * VEC, cbn = CG * VEC;   // channel range flushed by this block
  for (int i = threadIdx.x; i < 11 * P.C; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  if (ti < ipb) {
    float wa[9], wb[9], ga[9], gb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      wa[k] = P.w[k * P.C + c0]; wb[k] = P.w[k * P.C + c0 + 1];
      ga[k] = 0.f; gb[k] = 0.f;
    }
    const float sca = P.escale[c0], scb = P.escale[c0 + 1], sha = P.eshift[c0], shb = P.eshift[c0 + 1];
    float kaa, kab, kba, kbb, kca, kcb;
    if (bsrc.sums != nullptr) {
      const bool wr = blockIdx.x == 0 && ti == 0;
      bn_bwd_channel(bsrc, c0, wr, kaa, kba, kca);
      bn_bwd_channel(bsrc, c0 + 1, wr, kab, kbb, kcb);
    } else {
      kaa = P.dcoef[c0]; kab = P.dcoef[c0 + 1];
      kba = P.dcoef[P.C + c0]; kbb = P.dcoef[P.C + c0 + 1];
      kca = P.dcoef[2 * P.C + c0]; kcb = P.dcoef[2 * P.C + c0 + 1];
    }
    float s0a = 0.f, s0b = 0.f, s1a = 0.f, s1b = 0.f;
    const int rblocks = (P.H + TH - 1) / TH;
    const long long items = (long long)P.B * rblocks * P.nseg;
    for (long long item = (long long)blockIdx.x * ipb + ti; item < items; item += (long long)gridDim.x * ipb) {
      const int seg = (int)(item % P.nseg);
      const int hb = (int)((item / P.nseg) % rblocks);
      const int n = (int)(item / ((long long)P.nseg * rblocks));
      const int h0 = hb * TH, ws = seg * P.seglen, we = min(ws + P.seglen, P.W);
      if (ws >= we) continue;
      const __nv_bfloat16* ein = P.e_raw + (size_t)n * P.H * P.W * P.C + c0;
      __nv_bfloat16* dout = P.dze + (size_t)n * P.H * P.W * P.C + c0;
      const __nv_bfloat16* gz = P.dzd + (size_t)n * P.Ho * P.Wo * P.C + c0;
      const __nv_bfloat16* gr = P.d_raw + (size_t)n * P.Ho * P.Wo * P.C + c0;
      // one input pixel: mask, store, statistics (the tap sums da and the weight-gradient terms come from the caller)
      auto finish = [&](int h, int w, uint32_t eraw, float daa, float dab) {
        const float ea = bf16_lo(eraw), eb = bf16_hi(eraw);
        const float za = fmaf(ea, sca, sha), zb = fmaf(eb, scb, shb);
        const float oa = (za > 0.f && za < 6.f) ? daa : 0.f;
        const float ob = (zb > 0.f && zb < 6.f) ? dab : 0.f;
        const uint32_t pk = pack_bf16x2(oa, ob);
        *reinterpret_cast<uint32_t*>(dout + ((size_t)h * P.W + w) * P.C) = pk;
        const float qa = bf16_lo(pk), qb = bf16_hi(pk);
        s0a += qa; s0b += qb;
        s1a = fmaf(qa, ea, s1a); s1b = fmaf(qb, eb, s1b);
      };
      if (S == 1) {
        constexpr int R = TH + 2;                   // output rows ho = h0-1 .. h0+TH
        const __nv_bfloat16 *zp[R], *gp[R], *ep[TH];
        float rm[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int ho = h0 - 1 + r;
          const bool ok = ho >= 0 && ho < P.Ho;
          const size_t ro = (size_t)(ok ? ho : 0) * P.Wo * P.C;
          zp[r] = gz + ro;
          gp[r] = gr + ro;
          rm[r] = ok ? 1.f : 0.f;
        }
#pragma unroll
        for (int t = 0; t < TH; ++t) ep[t] = ein + (size_t)min(h0 + t, P.H - 1) * P.W * P.C;
        auto fetch = [&](int wo, uint32_t (&z)[R], uint32_t (&g)[R]) {
          const size_t co = (size_t)min(max(wo, 0), P.Wo - 1) * P.C;
#pragma unroll
          for (int r = 0; r < R; ++r) { z[r] = ldg4(zp[r] + co); g[r] = ldg4(gp[r] + co); }
        };
        auto xform = [&](int wo, const uint32_t (&z)[R], const uint32_t (&g)[R], float (&a)[R], float (&b)[R]) {
          const float cm = (wo >= 0 && wo < P.Wo) ? 1.f : 0.f;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const float m = rm[r] * cm;
            a[r] = fmaf(kaa, bf16_lo(z[r]), fmaf(kba, bf16_lo(g[r]), kca)) * m;
            b[r] = fmaf(kab, bf16_hi(z[r]), fmaf(kbb, bf16_hi(g[r]), kcb)) * m;
          }
        };
        auto fetch_e = [&](int w, uint32_t (&e)[TH]) {
          const size_t co = (size_t)min(w, P.W - 1) * P.C;
#pragma unroll
          for (int t = 0; t < TH; ++t) e[t] = ldg4(ep[t] + co);
        };
        float pa[R], pb[R], ca[R], cb[R], na[R], nb[R];
        uint32_t nz[R], ng[R], ne[TH], ec[TH];
        fetch(ws - 1, nz, ng); xform(ws - 1, nz, ng, pa, pb);
        fetch(ws, nz, ng);     xform(ws, nz, ng, ca, cb);
        fetch(ws + 1, nz, ng);
        fetch_e(ws, ne);
        for (int w = ws; w < we; ++w) {
          xform(w + 1, nz, ng, na, nb);
#pragma unroll
          for (int t = 0; t < TH; ++t) ec[t] = ne[t];
          if (w + 1 < we) { fetch(w + 2, nz, ng); fetch_e(w + 1, ne); }
#pragma unroll
          for (int t = 0; t < TH; ++t) {
            const int h = h0 + t;
            if (h < P.H) {
              const float aa = relu6f(fmaf(bf16_lo(ec[t]), sca, sha));
              const float ab = relu6f(fmaf(bf16_hi(ec[t]), scb, shb));
              float daa = 0.f, dab = 0.f;
#pragma unroll
              for (int ky = 0; ky < 3; ++ky) {
                const int r = t + 2 - ky;              // ho = h + 1 - ky  ->  row index ho - (h0 - 1)
                daa = fmaf(na[r], wa[ky * 3 + 0], daa); dab = fmaf(nb[r], wb[ky * 3 + 0], dab);   // kx = 0 -> wo = w + 1
                daa = fmaf(ca[r], wa[ky * 3 + 1], daa); dab = fmaf(cb[r], wb[ky * 3 + 1], dab);   // kx = 1 -> wo = w
                daa = fmaf(pa[r], wa[ky * 3 + 2], daa); dab = fmaf(pb[r], wb[ky * 3 + 2], dab);   // kx = 2 -> wo = w - 1
                ga[ky * 3 + 0] = fmaf(aa, na[r], ga[ky * 3 + 0]); gb[ky * 3 + 0] = fmaf(ab, nb[r], gb[ky * 3 + 0]);
                ga[ky * 3 + 1] = fmaf(aa, ca[r], ga[ky * 3 + 1]); gb[ky * 3 + 1] = fmaf(ab, cb[r], gb[ky * 3 + 1]);
                ga[ky * 3 + 2] = fmaf(aa, pa[r], ga[ky * 3 + 2]); gb[ky * 3 + 2] = fmaf(ab, pb[r], gb[ky * 3 + 2]);
              }
              finish(h, w, ec[t], daa, dab);
            }
          }
#pragma unroll
          for (int r = 0; r < R; ++r) { pa[r] = ca[r]; pb[r] = cb[r]; ca[r] = na[r]; cb[r] = nb[r]; }
        }
      } else {
        // h0 is even (TH == 2): rows h0 = 2r, h0+1; output rows r, r+1.  Column pairs (2c, 2c+1), output cols c, c+1.
        const int r0 = h0 >> 1;
        const int cs = ws >> 1, ce = (we + 1) >> 1;     // segments start at even columns (seglen is even)
        const __nv_bfloat16 *zp[2], *gp[2], *ep[2];
        float rm[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const bool ok = (r0 + r) < P.Ho;
          const size_t ro = (size_t)(ok ? r0 + r : 0) * P.Wo * P.C;
          zp[r] = gz + ro;
          gp[r] = gr + ro;
          rm[r] = ok ? 1.f : 0.f;
          ep[r] = ein + (size_t)min(h0 + r, P.H - 1) * P.W * P.C;
        }
        auto fetch = [&](int wo, uint32_t (&z)[2], uint32_t (&g)[2]) {
          const size_t co = (size_t)min(max(wo, 0), P.Wo - 1) * P.C;
#pragma unroll
          for (int r = 0; r < 2; ++r) { z[r] = ldg4(zp[r] + co); g[r] = ldg4(gp[r] + co); }
        };
        auto xform = [&](int wo, const uint32_t (&z)[2], const uint32_t (&g)[2], float (&a)[2], float (&b)[2]) {
          const float cm = (wo >= 0 && wo < P.Wo) ? 1.f : 0.f;
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const float m = rm[r] * cm;
            a[r] = fmaf(kaa, bf16_lo(z[r]), fmaf(kba, bf16_lo(g[r]), kca)) * m;
            b[r] = fmaf(kab, bf16_hi(z[r]), fmaf(kbb, bf16_hi(g[r]), kcb)) * m;
          }
        };
        auto fetch_e = [&](int c, uint32_t (&e)[4]) {    // e[2*dy + dx]; out-of-image pixels are masked where used
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int w = min(2 * c + (q & 1), P.W - 1);
            e[q] = ldg4(ep[q >> 1] + (size_t)w * P.C);
          }
        };
        float d0a[2], d0b[2], d1a[2], d1b[2];
        uint32_t nz[2], ng[2], ne[4], ec[4];
        fetch(cs, nz, ng); xform(cs, nz, ng, d0a, d0b);
        fetch(cs + 1, nz, ng);
        fetch_e(cs, ne);
        for (int c = cs; c < ce; ++c) {
          xform(c + 1, nz, ng, d1a, d1b);
#pragma unroll
          for (int q = 0; q < 4; ++q) ec[q] = ne[q];
          if (c + 1 < ce) { fetch(c + 2, nz, ng); fetch_e(c + 1, ne); }
          float aa[4], ab[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int hq = h0 + (q >> 1), wq = 2 * c + (q & 1);
            const bool v = hq < P.H && wq >= ws && wq < we;       // outside the image / segment: no contribution
            aa[q] = v ? relu6f(fmaf(bf16_lo(ec[q]), sca, sha)) : 0.f;
            ab[q] = v ? relu6f(fmaf(bf16_hi(ec[q]), scb, shb)) : 0.f;
          }
          // (even h, even w): tap (1,1) <- (r, c)
          float da0a = d0a[0] * wa[4], da0b = d0b[0] * wb[4];
          ga[4] = fmaf(aa[0], d0a[0], ga[4]); gb[4] = fmaf(ab[0], d0b[0], gb[4]);
          // (even h, odd w): (1,0) <- (r, c+1), (1,2) <- (r, c)
          float da1a = fmaf(d1a[0], wa[3], d0a[0] * wa[5]), da1b = fmaf(d1b[0], wb[3], d0b[0] * wb[5]);
          ga[3] = fmaf(aa[1], d1a[0], ga[3]); gb[3] = fmaf(ab[1], d1b[0], gb[3]);
          ga[5] = fmaf(aa[1], d0a[0], ga[5]); gb[5] = fmaf(ab[1], d0b[0], gb[5]);
          // (odd h, even w): (0,1) <- (r+1, c), (2,1) <- (r, c)
          float da2a = fmaf(d0a[1], wa[1], d0a[0] * wa[7]), da2b = fmaf(d0b[1], wb[1], d0b[0] * wb[7]);
          ga[1] = fmaf(aa[2], d0a[1], ga[1]); gb[1] = fmaf(ab[2], d0b[1], gb[1]);
          ga[7] = fmaf(aa[2], d0a[0], ga[7]); gb[7] = fmaf(ab[2], d0b[0], gb[7]);
          // (odd h, odd w): (0,0) <- (r+1, c+1), (0,2) <- (r+1, c), (2,0) <- (r, c+1), (2,2) <- (r, c)
          float da3a = fmaf(d1a[1], wa[0], fmaf(d0a[1], wa[2], fmaf(d1a[0], wa[6], d0a[0] * wa[8])));
          float da3b = fmaf(d1b[1], wb[0], fmaf(d0b[1], wb[2], fmaf(d1b[0], wb[6], d0b[0] * wb[8])));
          ga[0] = fmaf(aa[3], d1a[1], ga[0]); gb[0] = fmaf(ab[3], d1b[1], gb[0]);
          ga[2] = fmaf(aa[3], d0a[1], ga[2]); gb[2] = fmaf(ab[3], d0b[1], gb[2]);
          ga[6] = fmaf(aa[3], d1a[0], ga[6]); gb[6] = fmaf(ab[3], d1b[0], gb[6]);
          ga[8] = fmaf(aa[3], d0a[0], ga[8]); gb[8] = fmaf(ab[3], d0b[0], gb[8]);
          const int w0 = 2 * c;
          if (w0 >= ws && w0 < we) {
            if (h0 < P.H) finish(h0, w0, ec[0], da0a, da0b);
            if (h0 + 1 < P.H) finish(h0 + 1, w0, ec[2], da2a, da2b);
          }
          if (w0 + 1 >= ws && w0 + 1 < we) {
            if (h0 < P.H) finish(h0, w0 + 1, ec[1], da1a, da1b);
            if (h0 + 1 < P.H) finish(h0 + 1, w0 + 1, ec[3], da3a, da3b);
          }
#pragma unroll
          for (int r = 0; r < 2; ++r) { d0a[r] = d1a[r]; d0b[r] = d1b[r]; }
        }
      }
    }
    atomicAdd(&s_red[c0], s0a); atomicAdd(&s_red[c0 + 1], s0b);
    atomicAdd(&s_red[P.C + c0], s1a); atomicAdd(&s_red[P.C + c0 + 1], s1b);
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      atomicAdd(&s_red[(2 + k) * P.C + c0], ga[k]);
      atomicAdd(&s_red[(2 + k) * P.C + c0 + 1], gb[k]);
    }
  }
  __syncthreads();
  for (int i = cb0 + threadIdx.x; i < cb0 + cbn; i += blockDim.x) {
    atomicAdd(P.stats + i, s_red[i]);
    atomicAdd(P.stats + P.stats_stride + i, s_red[P.C + i]);
#pragma unroll
    for (int k = 0; k < 9; ++k) atomicAdd(P.dw + k * P.C + i, s_red[(2 + k) * P.C + i]);
  }
}

void pick_segments(int rows, int width, int B, int C, int* nseg, int* seglen) {
  // enough independent strips to fill the machine (~300k threads), segments no shorter than 7 pixels
  const long long base = (long long)B * ((rows + TH - 1) / TH) * (C / VEC);
  int n = 1;
  while (base * n < 600000 && width / (n * 2) >= 7) n *= 2;
  *nseg = n;
  *seglen = ((width + n - 1) / n + 1) & ~1;      // even: the stride-2 backward walks columns in pairs
}

}  // namespace

int dw_fwd_launch(const void* x, const float* scale, const float* shift, const float* w, void* y, float* stats,
                  int stats_stride, int B, int H, int W, int C, int stride, cudaStream_t s, const BnFwdSrc* src) {
  B200DP_REQUIRE(C % 16 == 0 && C <= 1024, "dw_fwd: C=%d unsupported", C);
  B200DP_REQUIRE(stride == 1 || stride == 2, "dw_fwd: stride must be 1 or 2");
  DwFwdParams P;
  P.x = static_cast<const __nv_bfloat16*>(x); P.scale = scale; P.shift = shift; P.w = w;
  P.y = static_cast<__nv_bfloat16*>(y); P.stats = stats; P.stats_stride = stats_stride;
  P.B = B; P.H = H; P.W = W; P.C = C;
  P.Ho = (H - 1) / stride + 1; P.Wo = (W - 1) / stride + 1;
  pick_segments(P.Ho, P.Wo, B, C, &P.nseg, &P.seglen);
  const int CG = C / VEC;
  const int csplit = (CG + 255) / 256;
  B200DP_REQUIRE(CG % csplit == 0, "dw_fwd: C=%d cannot be split evenly over blocks", C);
  P.cgb = CG / csplit;
  const int threads = (256 / P.cgb) * P.cgb;
  const int ipb = threads / P.cgb;
  const long long items = (long long)B * ((P.Ho + TH - 1) / TH) * P.nseg;
  long long blocks = (items + ipb - 1) / ipb;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const size_t smem = 2 * (size_t)C * sizeof(float);
  const dim3 grid((unsigned)blocks, (unsigned)csplit);
  if (stride == 1) B200DP_CUDA(launch_k(dw_fwd_kernel<1>, grid, dim3(threads), smem, s, P, src ? *src : BnFwdSrc{}));
  else B200DP_CUDA(launch_k(dw_fwd_kernel<2>, grid, dim3(threads), smem, s, P, src ? *src : BnFwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int dw_bwd_launch(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                  const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                  int H, int W, int C, int stride, cudaStream_t s, const BnBwdSrc* src) {
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
  const int csplit = (CG + 255) / 256;
  B200DP_REQUIRE(CG % csplit == 0, "dw_bwd: C=%d cannot be split evenly over blocks", C);
  P.cgb = CG / csplit;
  const int threads = (256 / P.cgb) * P.cgb;
  const int ipb = threads / P.cgb;
  const long long items = (long long)B * ((H + TH - 1) / TH) * P.nseg;
  long long blocks = (items + ipb - 1) / ipb;
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  const size_t smem = 11 * (size_t)C * sizeof(float);
  const dim3 grid((unsigned)blocks, (unsigned)csplit);
  if (stride == 1) B200DP_CUDA(launch_k(dw_bwd_kernel<1>, grid, dim3(threads), smem, s, P, src ? *src : BnBwdSrc{}));
  else B200DP_CUDA(launch_k(dw_bwd_kernel<2>, grid, dim3(threads), smem, s, P, src ? *src : BnBwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_dw_fwd(const void* x, const float* scale, const float* shift, const float* w, void* y,
                             float* stats, int stats_stride, int B, int H, int W, int C, int stride, void* stream) {
  return dw_fwd_launch(x, scale, shift, w, y, stats, stats_stride, B, H, W, C, stride, static_cast<cudaStream_t>(stream),
                       nullptr);
}

extern "C" int b200dp_dw_bwd(const void* e_raw, const float* escale, const float* eshift, const void* dzd,
                             const void* d_raw, const float* dcoef, const float* w, void* dze, float* stats,
                             int stats_stride, float* dw, int B, int H, int W, int C, int stride, void* stream) {
  return dw_bwd_launch(e_raw, escale, eshift, dzd, d_raw, dcoef, w, dze, stats, stats_stride, dw, B, H, W, C, stride,
                       static_cast<cudaStream_t>(stream), nullptr);
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
  if (int rc = dw_fwd_launch(x, in_scale, in_shift, w, y, sums, C, B, H, W, C, stride, s, nullptr)) return rc;
  const int Ho = (H - 1) / stride + 1, Wo = (W - 1) / stride + 1;
  return bn_fwd_finalize_launch(sums, C, gamma, beta, (float)B * Ho * Wo, eps, 0.f, nullptr, nullptr, coef_out,
                                saved_out, C, s);
}
