// Input pipeline on the device (SURVEY W11 + W12): uint8 32x32x3 -> normalised bf16 NHWC 224x224.
//
// Reference behaviour replaced (cifar10_mpi_mobilenet_224.py:72-89, :174): seven PIL transforms per
// image in DataLoader worker processes, then a blocking 77 MB fp32 H2D copy per step.  Here 393 KB
// of raw bytes cross PCIe and one kernel composes, per output pixel:
//   Resize(224) bilinear -> RandomResizedCrop box + bilinear resize -> HFlip -> ColorJitter (4 ops in
//   the sample's own order) -> Rotation (nearest, zero fill) -> ToTensor -> Normalize(ImageNet).
// The random draws arrive as one 12-float row per sample (data/transforms.py), so the CPU
// implementation (apply_pipeline_torch) and this kernel consume identical parameters.
//
// Augmenting kernel: one CTA per image.  Phase 0 builds the 224x224x3 resized image in shared
// memory (147 KB of the 227 KB), phase 1 reduces the grey mean the contrast op needs, phase 2
// writes the output.  Plain (test-time) kernel: a flat grid, 4-tap bilinear straight from a
// shared-memory copy of the 3 KB source image.
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int SRC = 32;
constexpr int NP = B200DP_AUG_NPARAM;
__constant__ float c_mean[3] = {0.485f * 255.f, 0.456f * 255.f, 0.406f * 255.f};
__constant__ float c_istd[3] = {1.f / (0.229f * 255.f), 1.f / (0.224f * 255.f), 1.f / (0.225f * 255.f)};
// lexicographic permutations of (brightness, contrast, saturation, hue) -- same table as transforms._PERMS
__constant__ unsigned char c_perm[24][4] = {
    {0, 1, 2, 3}, {0, 1, 3, 2}, {0, 2, 1, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {0, 3, 2, 1},
    {1, 0, 2, 3}, {1, 0, 3, 2}, {1, 2, 0, 3}, {1, 2, 3, 0}, {1, 3, 0, 2}, {1, 3, 2, 0},
    {2, 0, 1, 3}, {2, 0, 3, 1}, {2, 1, 0, 3}, {2, 1, 3, 0}, {2, 3, 0, 1}, {2, 3, 1, 0},
    {3, 0, 1, 2}, {3, 0, 2, 1}, {3, 1, 0, 2}, {3, 1, 2, 0}, {3, 2, 0, 1}, {3, 2, 1, 0}};

__device__ __forceinline__ float round_u8(float v) { return fminf(fmaxf(floorf(v + 0.5f), 0.f), 255.f); }

// torch/PIL bilinear upsampling with half-pixel centres, source clamped to [0, 31]
__device__ __forceinline__ void src_tap(int d, float scale, int& i0, int& i1, float& f) {
  float s = fmaxf(((float)d + 0.5f) * scale - 0.5f, 0.f);
  i0 = (int)s;
  if (i0 > SRC - 1) i0 = SRC - 1;
  i1 = min(i0 + 1, SRC - 1);
  f = s - (float)i0;
}

__device__ __forceinline__ float3 resized_px(const unsigned char* s_src, int y, int x, float scale) {
  int y0, y1, x0, x1;
  float fy, fx;
  src_tap(y, scale, y0, y1, fy);
  src_tap(x, scale, x0, x1, fx);
  const unsigned char* p00 = s_src + (y0 * SRC + x0) * 3;
  const unsigned char* p01 = s_src + (y0 * SRC + x1) * 3;
  const unsigned char* p10 = s_src + (y1 * SRC + x0) * 3;
  const unsigned char* p11 = s_src + (y1 * SRC + x1) * 3;
  float r[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // same association as ATen's upsample_bilinear2d: rows blended first, then columns
    const float top = (float)p00[c] * (1.f - fx) + (float)p01[c] * fx;
    const float bot = (float)p10[c] * (1.f - fx) + (float)p11[c] * fx;
    r[c] = round_u8(top * (1.f - fy) + bot * fy);
  }
  return make_float3(r[0], r[1], r[2]);
}

template <int OC>
__device__ __forceinline__ void store_px(__nv_bfloat16* out, size_t pix, float3 v) {
  const float a = (v.x - c_mean[0]) * c_istd[0];
  const float b = (v.y - c_mean[1]) * c_istd[1];
  const float c = (v.z - c_mean[2]) * c_istd[2];
  if constexpr (OC == 4) {
    *reinterpret_cast<uint2*>(out + pix * 4) = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, 0.f));
  } else {
    __nv_bfloat16* o = out + pix * 3;
    o[0] = __float2bfloat16(a);
    o[1] = __float2bfloat16(b);
    o[2] = __float2bfloat16(c);
  }
}

template <int OC>
__global__ void __launch_bounds__(256)
preprocess_plain_kernel(const uint8_t* __restrict__ images, __nv_bfloat16* __restrict__ out, int size,
                        int rows_per_cta) {
  pdl_trigger();
  pdl_wait();
  __shared__ unsigned char s_src[SRC * SRC * 3];
  const int img = blockIdx.y;
  const uint32_t* g = reinterpret_cast<const uint32_t*>(images + (size_t)img * SRC * SRC * 3);
  for (int i = threadIdx.x; i < SRC * SRC * 3 / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(s_src)[i] = g[i];
  __syncthreads();
  const float scale = (float)SRC / (float)size;
  const int y_lo = blockIdx.x * rows_per_cta;
  const int y_hi = min(y_lo + rows_per_cta, size);
  for (int p = y_lo * size + threadIdx.x; p < y_hi * size; p += blockDim.x) {
    const int y = p / size, x = p - y * size;
    store_px<OC>(out, (size_t)img * size * size + p, resized_px(s_src, y, x, scale));
  }
}

// ---- colour ops on 0..255 floats (match data/transforms.py) -------------------------------------
__device__ __forceinline__ float gray_of(float3 v) { return 0.299f * v.x + 0.587f * v.y + 0.114f * v.z; }
__device__ __forceinline__ float clamp255(float v) { return fminf(fmaxf(v, 0.f), 255.f); }
__device__ __forceinline__ float3 blend(float3 a, float3 b, float f) {
  return make_float3(clamp255(a.x + f * (b.x - a.x)), clamp255(a.y + f * (b.y - a.y)),
                     clamp255(a.z + f * (b.z - a.z)));
}

__device__ __forceinline__ float3 hue_shift(float3 v, float dh) {
  const float r = v.x, g = v.y, b = v.z;
  const float mx = fmaxf(fmaxf(r, g), b), mn = fminf(fminf(r, g), b);
  const float d = mx - mn;
  const float s = mx > 0.f ? d / fmaxf(mx, 1e-12f) : 0.f;
  const float dz = fmaxf(d, 1e-12f);
  float h;
  if (mx == r) {
    h = (g - b) / dz;
    h = h - 6.f * floorf(h / 6.f);
  } else if (mx == g) {
    h = (b - r) / dz + 2.f;
  } else {
    h = (r - g) / dz + 4.f;
  }
  h = d > 0.f ? h / 6.f : 0.f;
  h += dh;
  h = h - floorf(h);
  const float h6 = h * 6.f;
  const float fi = floorf(h6);
  const float f = h6 - fi;
  const float p = mx * (1.f - s), q = mx * (1.f - s * f), t = mx * (1.f - s * (1.f - f));
  int i = ((int)fi) % 6;
  switch (i) {
    case 0: return make_float3(mx, t, p);
    case 1: return make_float3(q, mx, p);
    case 2: return make_float3(p, mx, t);
    case 3: return make_float3(p, q, mx);
    case 4: return make_float3(t, p, mx);
    default: return make_float3(mx, p, q);
  }
}

struct AugP {
  float top, left, h, w, flip, fb, fc, fs, fh, angle;
  int order;
};

// crop-resize sample of the resized image held in shared memory
__device__ __forceinline__ float3 crop_px(const unsigned char* s_res, int S, const AugP& a, int v, int u) {
  if (a.flip > 0.f) u = S - 1 - u;
  float py = a.top + ((float)v + 0.5f) * (a.h / (float)S) - 0.5f;
  float px = a.left + ((float)u + 0.5f) * (a.w / (float)S) - 0.5f;
  const float ylo = a.top, yhi = a.top + a.h - 1.f, xlo = a.left, xhi = a.left + a.w - 1.f;
  py = fminf(fmaxf(py, ylo), yhi);
  px = fminf(fmaxf(px, xlo), xhi);
  const float y0f = floorf(py), x0f = floorf(px);
  const float fy = py - y0f, fx = px - x0f;
  const int y0 = (int)y0f, x0 = (int)x0f;
  const int y1 = (int)fminf(y0f + 1.f, yhi), x1 = (int)fminf(x0f + 1.f, xhi);
  const unsigned char* p00 = s_res + (y0 * S + x0) * 3;
  const unsigned char* p01 = s_res + (y0 * S + x1) * 3;
  const unsigned char* p10 = s_res + (y1 * S + x0) * 3;
  const unsigned char* p11 = s_res + (y1 * S + x1) * 3;
  float r[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float top = (float)p00[c] * (1.f - fx) + (float)p01[c] * fx;
    const float bot = (float)p10[c] * (1.f - fx) + (float)p11[c] * fx;
    r[c] = round_u8(top * (1.f - fy) + bot * fy);
  }
  return make_float3(r[0], r[1], r[2]);
}

// run the colour chain; `upto` = number of leading ops to apply (4 = all).  `mean` is the grey mean
// used by the contrast op.
__device__ __forceinline__ float3 color_chain(float3 v, const AugP& a, int upto, float mean) {
  for (int pos = 0; pos < upto; ++pos) {
    const int op = c_perm[a.order][pos];
    if (op == 0) {
      v = make_float3(clamp255(v.x * a.fb), clamp255(v.y * a.fb), clamp255(v.z * a.fb));
    } else if (op == 1) {
      v = blend(make_float3(mean, mean, mean), v, a.fc);
    } else if (op == 2) {
      const float g = gray_of(v);
      v = blend(make_float3(g, g, g), v, a.fs);
    } else {
      v = hue_shift(v, a.fh);
    }
  }
  return v;
}

template <int OC>
__global__ void __launch_bounds__(1024)
preprocess_aug_kernel(const uint8_t* __restrict__ images, const float* __restrict__ params,
                      __nv_bfloat16* __restrict__ out, int S) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ unsigned char smem[];
  unsigned char* s_src = smem;                      // 3072 B
  unsigned char* s_res = smem + SRC * SRC * 3;      // S*S*3 B
  __shared__ float s_red[32];
  __shared__ float s_mean;
  const int img = blockIdx.x;
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t* g = reinterpret_cast<const uint32_t*>(images + (size_t)img * SRC * SRC * 3);
  for (int i = tid; i < SRC * SRC * 3 / 4; i += nt) reinterpret_cast<uint32_t*>(s_src)[i] = g[i];
  const float* pr = params + (size_t)img * NP;
  AugP a;
  a.top = pr[0]; a.left = pr[1]; a.h = pr[2]; a.w = pr[3]; a.flip = pr[4];
  a.fb = pr[5]; a.fc = pr[6]; a.fs = pr[7]; a.fh = pr[8];
  a.order = min(max((int)pr[9], 0), 23);
  a.angle = pr[10];
  const bool enabled = pr[11] > 0.f;
  __syncthreads();
  const float scale = (float)SRC / (float)S;
  const int npx = S * S;
  if (!enabled) {  // test-transform row inside an augmenting batch
    for (int p = tid; p < npx; p += nt)
      store_px<OC>(out, (size_t)img * npx + p, resized_px(s_src, p / S, p % S, scale));
    return;
  }
  // phase 0: resized image -> shared memory
  for (int p = tid; p < npx; p += nt) {
    const float3 v = resized_px(s_src, p / S, p % S, scale);
    s_res[p * 3 + 0] = (unsigned char)v.x;
    s_res[p * 3 + 1] = (unsigned char)v.y;
    s_res[p * 3 + 2] = (unsigned char)v.z;
  }
  __syncthreads();
  // phase 1: grey mean of the image as it enters the contrast op
  int cpos = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (c_perm[a.order][i] == 1) cpos = i;
  float part = 0.f;
  for (int p = tid; p < npx; p += nt)
    part += gray_of(color_chain(crop_px(s_res, S, a, p / S, p % S), a, cpos, 0.f));
  part = warp_sum(part);
  if ((tid & 31) == 0) s_red[tid >> 5] = part;
  __syncthreads();
  if (tid < 32) {
    float t = tid < (nt >> 5) ? s_red[tid] : 0.f;
    t = warp_sum(t);
    if (tid == 0) s_mean = floorf(t / (float)npx + 0.5f);
  }
  __syncthreads();
  const float mean = s_mean;
  // phase 2: rotate (inverse map, nearest, zero fill) -> crop/flip -> colour -> normalise
  const float ang = -a.angle * 0.017453292519943295f;
  const float ca = cosf(ang), sa = sinf(ang);
  const float cx = (float)S * 0.5f;
  const float ra = ca, rb = sa, rd = -sa, re = ca;
  const float rc = cx - (ra * cx + rb * cx);
  const float rf = cx - (rd * cx + re * cx);
  for (int p = tid; p < npx; p += nt) {
    const int y = p / S, x = p - y * S;
    const float xin = ra * ((float)x + 0.5f) + rb * ((float)y + 0.5f) + rc;
    const float yin = rd * ((float)x + 0.5f) + re * ((float)y + 0.5f) + rf;
    const int u = (int)floorf(xin), v = (int)floorf(yin);
    float3 val = make_float3(0.f, 0.f, 0.f);
    if (u >= 0 && u < S && v >= 0 && v < S) val = color_chain(crop_px(s_res, S, a, v, u), a, 4, mean);
    store_px<OC>(out, (size_t)img * npx + p, val);
  }
}


// =====================================================================================================================
// Fused input pipeline for the training / evaluation step: uint8 32x32x3 -> the stem convolution's im2col matrix.
//
// The stem (3x3, stride 2, 3 -> 32) runs as a tcgen05 GEMM over rows of 27 taps + 5 zeros (csrc/stem.cu).  Writing the
// normalised 224x224 image (51 MB at batch 128) only to read it back and scatter it into those rows was a 194 us + 56 us
// pair of kernels; this kernel produces the rows directly and never materialises the image:
//   * one CTA per image, 1024 threads; source and resized image live in shared memory as PACKED pixels (one 32-bit word
//     r | g << 8 | b << 16 per pixel: a bilinear tap is one LDS.32 + three byte->float conversions instead of three
//     byte loads), 4 KB + 196 KB;
//   * pixel coordinates advance incrementally (no integer division by the runtime image size), and both interpolation
//     chains are separable, so their taps (two indices + a fraction per output row / column) are tabulated once per
//     image instead of being recomputed per pixel;
//   * the output is produced in bands of 4 im2col rows = 8 new image rows: the augmented, normalised pixels of a band go
//     to a 9-row ring in shared memory (the row shared with the previous band is kept), then every thread assembles
//     half an im2col row (32 bytes) from it -- a warp stores 1 KB contiguous;
//   * the column sums of the matrix (needed by the factored stem weight gradient) are accumulated on the way.
// The arithmetic per pixel is the same as in the kernels above (same device functions, same association), so the rows
// equal im2col(b200dp_preprocess(...)) bit for bit.
// =====================================================================================================================
constexpr int BAND = 4;                    // im2col rows (of Wo pixels) per band
constexpr int RING = 2 * BAND + 1;         // image rows held at a time

__device__ __forceinline__ float3 unpack_rgb(uint32_t v) {
  return make_float3((float)(v & 0xffu), (float)((v >> 8) & 0xffu), (float)((v >> 16) & 0xffu));
}
__device__ __forceinline__ uint32_t pack_rgb(float3 v) {
  return (uint32_t)v.x | ((uint32_t)v.y << 8) | ((uint32_t)v.z << 16);
}
// rows blended first, then columns, each channel rounded to an 8-bit level (same association as resized_px / crop_px)
// (a convex combination of 8-bit levels stays inside [0, 255]: the clamp of round_u8 is not needed here)
__device__ __forceinline__ float round_level(float v) { return floorf(v + 0.5f); }
__device__ __forceinline__ float3 bilerp_u8(uint32_t q00, uint32_t q01, uint32_t q10, uint32_t q11, float fx, float fy) {
  const float3 a = unpack_rgb(q00), b = unpack_rgb(q01), c = unpack_rgb(q10), d = unpack_rgb(q11);
  const float gx = 1.f - fx, gy = 1.f - fy;
  return make_float3(round_level((a.x * gx + b.x * fx) * gy + (c.x * gx + d.x * fx) * fy),
                     round_level((a.y * gx + b.y * fx) * gy + (c.y * gx + d.y * fx) * fy),
                     round_level((a.z * gx + b.z * fx) * gy + (c.z * gx + d.z * fx) * fy));
}
// A bilinear tap along one axis, tabulated per output coordinate: both interpolation chains of this pipeline are
// separable (resize: source row / column per output row / column; crop: row of the resized image per crop row, column
// per crop column with the flip folded in), so the coordinate arithmetic is done S times per axis instead of per pixel.
struct Tap { int i0, i1; float f; };
__device__ __forceinline__ uint2 tap_pack(int i0, int i1, float f) { return make_uint2((uint32_t)i0 | ((uint32_t)i1 << 16), __float_as_uint(f)); }
__device__ __forceinline__ Tap tap_unpack(uint2 t) { return Tap{(int)(t.x & 0xffffu), (int)(t.x >> 16), __uint_as_float(t.y)}; }
__device__ __forceinline__ float3 sample_tab(const uint32_t* img, int pitch, uint2 ty, uint2 tx) {
  const Tap y = tap_unpack(ty), x = tap_unpack(tx);
  return bilerp_u8(img[y.i0 * pitch + x.i0], img[y.i0 * pitch + x.i1], img[y.i1 * pitch + x.i0], img[y.i1 * pitch + x.i1], x.f, y.f);
}
// the colour chain with the permutation held in a register (four 2-bit op codes) instead of constant-memory look-ups
__device__ __forceinline__ float3 color_chain_r(float3 v, const AugP& a, uint32_t perm, int upto, float mean) {
  for (int pos = 0; pos < upto; ++pos) {
    const uint32_t op = (perm >> (2 * pos)) & 3u;
    if (op == 0) {
      v = make_float3(clamp255(v.x * a.fb), clamp255(v.y * a.fb), clamp255(v.z * a.fb));
    } else if (op == 1) {
      v = blend(make_float3(mean, mean, mean), v, a.fc);
    } else if (op == 2) {
      const float g = gray_of(v);
      v = blend(make_float3(g, g, g), v, a.fs);
    } else {
      v = hue_shift(v, a.fh);
    }
  }
  return v;
}
__device__ __forceinline__ uint2 norm_px(float3 v) {
  return make_uint2(pack_bf16x2((v.x - c_mean[0]) * c_istd[0], (v.y - c_mean[1]) * c_istd[1]),
                    pack_bf16x2((v.z - c_mean[2]) * c_istd[2], 0.f));
}

struct ColParams {
  const uint8_t* images;
  const float* params;       // null: test transform for every image
  __nv_bfloat16* col;        // [B * Ho * Wo][32]
  float* asum;               // [32] += column sums (null: not wanted)
  int S;
};

__global__ void __launch_bounds__(1024)
preprocess_col_kernel(const ColParams P) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) unsigned char smem[];
  const int S = P.S, Ho = S / 2, Wo = S / 2;
  uint32_t* s_src = reinterpret_cast<uint32_t*>(smem);                       // 32 x 32 packed pixels
  uint32_t* s_res = s_src + SRC * SRC;                                       // S x S packed pixels (augmenting images)
  uint2* s_ring = reinterpret_cast<uint2*>(s_res + (size_t)S * S);           // RING x S normalised bf16 pixels (r, g, b, 0)
  uint2* s_tab = s_ring + (size_t)RING * S;                                  // [4][S]: resize y / x taps, crop y / x taps
  uint2 *t_ry = s_tab, *t_rx = s_tab + S, *t_cy = s_tab + 2 * S, *t_cx = s_tab + 3 * S;
  __shared__ float s_red[32];
  __shared__ float s_mean;
  __shared__ float s_sum[32];
  const int img = blockIdx.x;
  const int tid = threadIdx.x;
  constexpr int NT = 1024;
  {
    const uint8_t* g = P.images + (size_t)img * SRC * SRC * 3;
    for (int i = tid; i < SRC * SRC; i += NT)
      s_src[i] = (uint32_t)g[3 * i] | ((uint32_t)g[3 * i + 1] << 8) | ((uint32_t)g[3 * i + 2] << 16);
  }
  if (tid < 32) s_sum[tid] = 0.f;
  AugP a{};
  bool enabled = false;
  if (P.params != nullptr) {
    const float* pr = P.params + (size_t)img * NP;
    a.top = pr[0]; a.left = pr[1]; a.h = pr[2]; a.w = pr[3]; a.flip = pr[4];
    a.fb = pr[5]; a.fc = pr[6]; a.fs = pr[7]; a.fh = pr[8];
    a.order = min(max((int)pr[9], 0), 23);
    a.angle = pr[10];
    enabled = pr[11] > 0.f;
  }
  const float scale = (float)SRC / (float)S;
  for (int i = tid; i < 2 * S; i += NT) {           // resize taps (the same along both axes)
    int i0, i1;
    float f;
    src_tap(i < S ? i : i - S, scale, i0, i1, f);
    s_tab[i] = tap_pack(i0, i1, f);
  }
  if (enabled) {
    for (int i = tid; i < 2 * S; i += NT) {         // crop taps: rows of the box, columns of the box (flip folded in)
      const bool isx = i >= S;
      int o = isx ? i - S : i;
      if (isx && a.flip > 0.f) o = S - 1 - o;
      const float start = isx ? a.left : a.top, len = isx ? a.w : a.h;
      float pos = start + ((float)o + 0.5f) * (len / (float)S) - 0.5f;
      const float lo = start, hi = start + len - 1.f;
      pos = fminf(fmaxf(pos, lo), hi);
      const float p0 = floorf(pos);
      s_tab[2 * S + i] = tap_pack((int)p0, (int)fminf(p0 + 1.f, hi), pos - p0);
    }
  }
  uint32_t perm = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) perm |= (uint32_t)c_perm[a.order][i] << (2 * i);
  __syncthreads();
  const int npx = S * S;
  // (y, x) of pixel index tid, advanced by NT pixels per iteration without dividing
  const int dy = NT / S, dx = NT - dy * S;
  float mean = 0.f;
  if (enabled) {
    // phase 0: resized image -> shared memory
    {
      int y = tid / S, x = tid - y * S;
      for (int p = tid; p < npx; p += NT) {
        s_res[p] = pack_rgb(sample_tab(s_src, SRC, t_ry[y], t_rx[x]));
        y += dy; x += dx;
        if (x >= S) { x -= S; ++y; }
      }
    }
    __syncthreads();
    // phase 1: grey mean of the image as it enters the contrast op
    int cpos = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (c_perm[a.order][i] == 1) cpos = i;
    float part = 0.f;
    {
      int y = tid / S, x = tid - y * S;
      for (int p = tid; p < npx; p += NT) {
        part += gray_of(color_chain_r(sample_tab(s_res, S, t_cy[y], t_cx[x]), a, perm, cpos, 0.f));
        y += dy; x += dx;
        if (x >= S) { x -= S; ++y; }
      }
    }
    part = warp_sum(part);
    if ((tid & 31) == 0) s_red[tid >> 5] = part;
    __syncthreads();
    if (tid < 32) {
      float t = s_red[tid];
      t = warp_sum(t);
      if (tid == 0) s_mean = floorf(t / (float)npx + 0.5f);
    }
    __syncthreads();
    mean = s_mean;
  }
  // rotation: inverse map, nearest, zero fill
  const float ang = -a.angle * 0.017453292519943295f;
  const float ca = cosf(ang), sa = sinf(ang);
  const float cx = (float)S * 0.5f;
  const float ra = ca, rb = sa, rd = -sa, re = ca;
  const float rc = cx - (ra * cx + rb * cx);
  const float rf = cx - (rd * cx + re * cx);
  auto final_px = [&](int y, int x) -> uint2 {          // normalised output pixel (y, x) of this image
    if (!enabled) return norm_px(sample_tab(s_src, SRC, t_ry[y], t_rx[x]));
    const float xin = ra * ((float)x + 0.5f) + rb * ((float)y + 0.5f) + rc;
    const float yin = rd * ((float)x + 0.5f) + re * ((float)y + 0.5f) + rf;
    const int u = (int)floorf(xin), v = (int)floorf(yin);
    float3 val = make_float3(0.f, 0.f, 0.f);
    if (u >= 0 && u < S && v >= 0 && v < S) val = color_chain_r(sample_tab(s_res, S, t_cy[v], t_cx[u]), a, perm, 4, mean);
    return norm_px(val);
  };
  // ---- bands: 2*BAND new image rows -> ring, then BAND im2col rows of Wo pixels out of the ring
  const int half = tid >> 9;               // warp-uniform: threads 0..511 write k = 0..15 of a row, 512..1023 k = 16..31
  const int slot = tid & 511;
  float cs[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) cs[k] = 0.f;
  // image row y lives in ring slot (y + 1) % RING; row -1 (zero padding above the image) is slot 0
  for (int i = tid; i < S; i += NT) s_ring[i] = make_uint2(0u, 0u);
  const int nbands = Ho / BAND;
  for (int b = 0; b < nbands; ++b) {
    const int y_first = 2 * BAND * b;                     // new rows y_first .. y_first + 2*BAND - 1
    for (int i = tid; i < 2 * BAND * S; i += NT) {
      const int r = i / S, x = i - r * S;                 // r < 8: a shift when S is a power of two, else one small division
      const int y = y_first + r;
      s_ring[((y + 1) % RING) * S + x] = final_px(y, x);
    }
    __syncthreads();
    const int nrows = BAND * Wo;
    if (slot < nrows) {
      const int hr = slot / Wo, wo = slot - hr * Wo;
      const int ho = BAND * b + hr;
      // taps of this half: k = 3 * t + c.  half 0: taps 0..4 whole, tap 5 channel 0; half 1: tap 5 channels 1, 2, taps 6..8
      uint32_t h[16];
      auto px_at = [&](int t) -> uint2 {
        const int iy = 2 * ho + t / 3 - 1, ix = 2 * wo + t % 3 - 1;
        if (ix < 0) return make_uint2(0u, 0u);            // left padding (ix <= S - 1 always; iy >= -1 is ring slot 0)
        return s_ring[((iy + 1) % RING) * S + ix];
      };
      if (half == 0) {
#pragma unroll
        for (int t = 0; t < 5; ++t) {
          const uint2 v = px_at(t);
          h[3 * t + 0] = v.x & 0xffffu; h[3 * t + 1] = v.x >> 16; h[3 * t + 2] = v.y & 0xffffu;
        }
        h[15] = px_at(5).x & 0xffffu;
      } else {
        const uint2 v5 = px_at(5);
        h[0] = v5.x >> 16; h[1] = v5.y & 0xffffu;
#pragma unroll
        for (int t = 6; t < 9; ++t) {
          const uint2 v = px_at(t);
          h[3 * t - 16] = v.x & 0xffffu; h[3 * t - 15] = v.x >> 16; h[3 * t - 14] = v.y & 0xffffu;
        }
#pragma unroll
        for (int k = 11; k < 16; ++k) h[k] = 0u;
      }
      const size_t row = ((size_t)img * Ho + ho) * Wo + wo;
      uint4* dst = reinterpret_cast<uint4*>(P.col + row * 32 + half * 16);
      dst[0] = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
      dst[1] = make_uint4(h[8] | (h[9] << 16), h[10] | (h[11] << 16), h[12] | (h[13] << 16), h[14] | (h[15] << 16));
      if (P.asum != nullptr) {
#pragma unroll
        for (int k = 0; k < 16; ++k) cs[k] += __uint_as_float(h[k] << 16);
      }
    }
    __syncthreads();
  }
  if (P.asum != nullptr) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float t = warp_sum(cs[k]);
      if ((tid & 31) == 0 && half * 16 + k < 27) atomicAdd(&s_sum[half * 16 + k], t);
    }
    __syncthreads();
    if (tid < 27) atomicAdd(P.asum + tid, s_sum[tid]);
  }
}

// ---- inference on photographs: arbitrary H x W uint8 RGB -> normalised bf16 [S, S, 4] ------------------------------
// What the reference's predict helper and Gradio app do with `transforms.Resize((224, 224))` on the ORIGINAL image
// (cifar10_128batch.py:42-47, :157, :165): PIL's two-pass bilinear resample -- triangle filter whose support grows with
// the down-scaling factor (anti-aliasing), horizontal pass rounded to 8 bits, then the vertical pass -- followed by
// ToTensor and Normalize.  One thread per output pixel walks its (vertical x horizontal) window of the source.
struct ResizeParams {
  const uint8_t* src;       // [batch][H][W][3]
  __nv_bfloat16* out;       // [batch][S][S][4]
  int H, W, S;
};

__device__ __forceinline__ void pil_window(int o, float scale, int in_size, int& lo, int& hi, float& center, float& fscale) {
  fscale = fmaxf(scale, 1.f);
  const float support = fscale;                       // bilinear filter support 1.0, stretched when shrinking
  center = ((float)o + 0.5f) * scale;
  lo = max(0, (int)(center - support + 0.5f));
  hi = min(in_size, (int)(center + support + 0.5f));
}
__device__ __forceinline__ float tri(float t) { return fmaxf(0.f, 1.f - fabsf(t)); }

__global__ void __launch_bounds__(256)
resize_normalize_kernel(const ResizeParams P) {
  pdl_trigger();
  pdl_wait();
  const int S = P.S;
  const int img = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= S * S) return;
  const int oy = p / S, ox = p - oy * S;
  const float sx = (float)P.W / (float)S, sy = (float)P.H / (float)S;
  int x0, x1, y0, y1;
  float cxn, cyn, fx, fy;
  pil_window(ox, sx, P.W, x0, x1, cxn, fx);
  pil_window(oy, sy, P.H, y0, y1, cyn, fy);
  float wxs = 0.f, wys = 0.f;
  for (int x = x0; x < x1; ++x) wxs += tri(((float)x - cxn + 0.5f) / fx);
  for (int y = y0; y < y1; ++y) wys += tri(((float)y - cyn + 0.5f) / fy);
  const uint8_t* base = P.src + (size_t)img * P.H * P.W * 3;
  float acc[3] = {0.f, 0.f, 0.f};
  for (int y = y0; y < y1; ++y) {
    const float wy = tri(((float)y - cyn + 0.5f) / fy) / wys;
    const uint8_t* row = base + (size_t)y * P.W * 3;
    float h[3] = {0.f, 0.f, 0.f};
    for (int x = x0; x < x1; ++x) {
      const float wx = tri(((float)x - cxn + 0.5f) / fx) / wxs;
      h[0] += wx * (float)row[3 * x];
      h[1] += wx * (float)row[3 * x + 1];
      h[2] += wx * (float)row[3 * x + 2];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) acc[c] += wy * round_u8(h[c]);      // the horizontal pass is stored as 8-bit pixels
  }
  const float3 v = make_float3(round_u8(acc[0]), round_u8(acc[1]), round_u8(acc[2]));
  *reinterpret_cast<uint2*>(P.out + ((size_t)img * S * S + p) * 4) = norm_px(v);
}

size_t col_smem_bytes(int S) { return (size_t)SRC * SRC * 4 + (size_t)S * S * 4 + (size_t)RING * S * 8 + (size_t)4 * S * 8; }

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_preprocess(const uint8_t* images, const float* params, void* out_bf16, int batch,
                                 int size, int out_channels, void* stream) {
  B200DP_REQUIRE(images && out_bf16 && batch > 0, "preprocess: bad arguments");
  B200DP_REQUIRE(out_channels == 3 || out_channels == 4, "preprocess: out_channels must be 3 or 4");
  B200DP_REQUIRE(size >= 32 && size <= 256, "preprocess: size %d unsupported", size);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* out = static_cast<__nv_bfloat16*>(out_bf16);
  if (params == nullptr) {
    const int rows = 16;
    dim3 grid((size + rows - 1) / rows, batch);
    if (out_channels == 4) B200DP_CUDA(launch_k(preprocess_plain_kernel<4>, grid, dim3(256), 0, s, images, out, size, rows));
    else B200DP_CUDA(launch_k(preprocess_plain_kernel<3>, grid, dim3(256), 0, s, images, out, size, rows));
    B200DP_CHECK_LAUNCH();
    return B200DP_OK;
  }
  const size_t smem = SRC * SRC * 3 + (size_t)size * size * 3;
  // per-device attribute; cheap enough to set on every call
  B200DP_CUDA(cudaFuncSetAttribute(preprocess_aug_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  B200DP_CUDA(cudaFuncSetAttribute(preprocess_aug_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  if (out_channels == 4) B200DP_CUDA(launch_k(preprocess_aug_kernel<4>, dim3(batch), dim3(1024), smem, s, images, params, out, size));
  else B200DP_CUDA(launch_k(preprocess_aug_kernel<3>, dim3(batch), dim3(1024), smem, s, images, params, out, size));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

// Fused preprocess + stem im2col (see preprocess_col_kernel): `col` = bf16 [batch * (size/2)^2][32], `asum` (may be
// NULL) fp32 [32] += column sums.  Supported for even sizes whose working set fits in shared memory (size <= 224 and
// (size/2) % 4 == 0); returns B200DP_EINVAL otherwise -- callers fall back to b200dp_preprocess + the im2col pass.
extern "C" int b200dp_preprocess_col_supported(int size) {
  return size >= 32 && size % 8 == 0 && (size / 2) * 4 <= 512 && col_smem_bytes(size) + 512 <= 227 * 1024 ? 1 : 0;
}

extern "C" int b200dp_preprocess_col(const uint8_t* images, const float* params, void* col_bf16, float* asum, int batch,
                                     int size, void* stream) {
  B200DP_REQUIRE(images && col_bf16 && batch > 0, "preprocess_col: bad arguments");
  B200DP_REQUIRE(b200dp_preprocess_col_supported(size), "preprocess_col: size %d unsupported", size);
  ColParams P;
  P.images = images; P.params = params; P.col = static_cast<__nv_bfloat16*>(col_bf16); P.asum = asum; P.S = size;
  const size_t smem = col_smem_bytes(size);
  B200DP_CUDA(cudaFuncSetAttribute(preprocess_col_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 512));
  B200DP_CUDA(launch_k(preprocess_col_kernel, dim3(batch), dim3(1024), smem, static_cast<cudaStream_t>(stream), P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

// Resize (PIL-style anti-aliased bilinear) + ToTensor + Normalize of `batch` RGB images of one size, for inference on
// full-resolution photographs: src uint8 [batch][H][W][3] -> out bf16 [batch][size][size][4] (4th channel 0), the
// layout b200dp_engine_forward_preprocessed consumes.
extern "C" int b200dp_resize_normalize(const uint8_t* src, int batch, int H, int W, void* out_bf16, int size, void* stream) {
  B200DP_REQUIRE(src && out_bf16 && batch > 0 && H > 0 && W > 0, "resize_normalize: bad arguments");
  B200DP_REQUIRE(size >= 32 && size <= 1024, "resize_normalize: size %d unsupported", size);
  ResizeParams P;
  P.src = src; P.out = static_cast<__nv_bfloat16*>(out_bf16); P.H = H; P.W = W; P.S = size;
  B200DP_CUDA(launch_k(resize_normalize_kernel, dim3((size * size + 255) / 256, batch), dim3(256), 0,
                       static_cast<cudaStream_t>(stream), P));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}
