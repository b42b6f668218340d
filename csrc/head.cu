// Loss / metrics / inference tail (SURVEY W9 tail, W10; K10, K11, K15).
//
// Reference behaviour replaced: nn.CrossEntropyLoss (log_softmax + nll_loss), torch.max(outputs, 1),
// (preds == labels).sum().item() and loss.item() -- five ATen kernels and two host syncs per step
// (cifar10_mpi_mobilenet_224.py:178-184) -- and softmax + topk at inference
// (cifar10_128batch.py:170-172).  One warp per sample; loss sum / correct count / sample count are
// accumulated on the device as doubles and read once per epoch.
#include "bn_inline.cuh"
#include "common.cuh"

namespace b200dp {
namespace {

__global__ void __launch_bounds__(128)
softmax_xent_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                    float* __restrict__ loss_out, float* __restrict__ dlogits, double* __restrict__ metrics,
                    int batch, int classes, float inv_mean) {
  pdl_trigger();
  pdl_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  __shared__ float s_loss[4];
  __shared__ int s_corr[4];
  float my_loss = 0.f;
  int my_corr = 0;
  if (warp < batch) {
    const float* row = logits + (size_t)warp * classes;
    float mx = -INFINITY;
    int arg = 0;
    for (int c = lane; c < classes; c += 32) {
      const float v = row[c];
      if (v > mx) { mx = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (ov > mx || (ov == mx && oa < arg)) { mx = ov; arg = oa; }   // first max, like torch.max
    }
    float se = 0.f;
    for (int c = lane; c < classes; c += 32) se += __expf(row[c] - mx);
    se = warp_sum(se);
    const int label = (int)labels[warp];
    const float lse = mx + __logf(se);
    if (dlogits != nullptr) {
      for (int c = lane; c < classes; c += 32) {
        const float p = __expf(row[c] - lse);
        dlogits[(size_t)warp * classes + c] = (p - (c == label ? 1.f : 0.f)) * inv_mean;
      }
    }
    if (lane == 0) {
      my_loss = lse - row[label];
      my_corr = (arg == label) ? 1 : 0;
      if (loss_out != nullptr) loss_out[warp] = my_loss;
    }
  }
  if (lane == 0) {
    s_loss[threadIdx.x >> 5] = my_loss;
    s_corr[threadIdx.x >> 5] = my_corr;
  }
  __syncthreads();
  if (threadIdx.x == 0 && metrics != nullptr) {
    double l = 0.0;
    int c = 0, n = 0;
    for (int w = 0; w < 4; ++w) {
      l += s_loss[w];
      c += s_corr[w];
      if (blockIdx.x * 4 + w < batch) ++n;
    }
    atomicAdd(&metrics[0], l);
    atomicAdd(&metrics[1], (double)c);
    atomicAdd(&metrics[2], (double)n);
  }
}

__global__ void __launch_bounds__(128)
softmax_topk_kernel(const float* __restrict__ logits, float* __restrict__ probs, float* __restrict__ topv,
                    int32_t* __restrict__ topi, int batch, int classes, int k) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= batch) return;
  const float* row = logits + (size_t)warp * classes;
  float mx = -INFINITY;
  for (int c = lane; c < classes; c += 32) mx = fmaxf(mx, row[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float se = 0.f;
  for (int c = lane; c < classes; c += 32) se += __expf(row[c] - mx);
  se = warp_sum(se);
  const float inv = 1.f / se;
  if (probs != nullptr)
    for (int c = lane; c < classes; c += 32) probs[(size_t)warp * classes + c] = __expf(row[c] - mx) * inv;
  // k rounds of warp arg-max with already-chosen entries masked out
  int chosen_prev = -1;
  float chosen_val = INFINITY;
  for (int j = 0; j < k; ++j) {
    float best = -INFINITY;
    int barg = 0x7fffffff;
    for (int c = lane; c < classes; c += 32) {
      const float v = row[c];
      const bool taken = (v > chosen_val) || (v == chosen_val && c <= chosen_prev);
      if (!taken && (v > best || (v == best && c < barg))) { best = v; barg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oa = __shfl_xor_sync(0xffffffffu, barg, o);
      if (ov > best || (ov == best && oa < barg)) { best = ov; barg = oa; }
    }
    if (lane == 0) {
      topv[(size_t)warp * k + j] = __expf(best - mx) * inv;
      topi[(size_t)warp * k + j] = barg;
    }
    chosen_val = best;
    chosen_prev = barg;
  }
}


// ------------------------------------------------------------------------------------------------------
// Fused classifier head for the native engine.
//   pool:    pooled[b][c] = drop[b][c] * mean_pos relu6(scale[c]*raw[b][pos][c] + shift[c])
//            (BatchNorm + ReLU6 of features.18 applied on load; dropout mask from a counter-based hash)
//   fc:      logits[b][j] = pooled[b] . W[j] + bias[j]
//   bwd_fc:  dpool[b][c] = drop[b][c]/HW * sum_j dlogits[b][j] W[j][c];  dW[j][c] += sum_b dlogits[b][j] pooled[b][c]
//   bwd_mask: dz[b][pos][c] = dpool[b][c] * [0 < scale*raw+shift < 6]  (+ per-channel sum dz, sum dz*raw)
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

__global__ void __launch_bounds__(256)
head_pool_kernel(const __nv_bfloat16* __restrict__ raw, const float* __restrict__ coef, float* __restrict__ pooled,
                 float* __restrict__ drop, int B, int HW, int C, float p_drop, uint32_t seed, const int* step,
                 const BnFwdSrc bsrc) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (c >= C) return;
  float sc, sh;
  if (bsrc.sums != nullptr) bn_fwd_channel(bsrc, c, b == 0, sc, sh);
  else { sc = coef[c]; sh = coef[C + c]; }
  const __nv_bfloat16* src = raw + (size_t)b * HW * C + c;
  float acc = 0.f;
  for (int p = 0; p < HW; ++p) acc += relu6f(fmaf(__bfloat162float(src[(size_t)p * C]), sc, sh));
  acc /= (float)HW;
  float keep = 1.f;
  if (drop != nullptr) {
    if (p_drop > 0.f) {
      const uint32_t st = step != nullptr ? (uint32_t)(*step) : 0u;
      const uint32_t h = hash32(hash32(seed ^ (st * 0x9e3779b9u)) ^ (uint32_t)(b * C + c));
      keep = ((float)(h >> 8) * (1.f / 16777216.f)) >= p_drop ? 1.f / (1.f - p_drop) : 0.f;
    }
    drop[(size_t)b * C + c] = keep;
  }
  pooled[(size_t)b * C + c] = acc * keep;
}

__global__ void __launch_bounds__(512)
head_fc_kernel(const float* __restrict__ pooled, const float* __restrict__ w, const float* __restrict__ bias,
               float* __restrict__ logits, int C, int classes) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = warp; j < classes; j += blockDim.x >> 5) {
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(pooled[(size_t)b * C + c], w[(size_t)j * C + c], acc);
    acc = warp_sum(acc);
    if (lane == 0) logits[(size_t)b * classes + j] = acc + bias[j];
  }
}

constexpr int HEAD_MAX_CLASSES = 16;

// grid (C / 128, B / HEAD_BROWS): a block takes 128 channels x HEAD_BROWS images (the earlier one-block-per-128-channels
// version walked the whole batch serially, one dependent load after another: 84 us for 164k MACs)
constexpr int HEAD_BROWS = 8;

__global__ void __launch_bounds__(128)
head_bwd_fc_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled, const float* __restrict__ drop,
                   const float* __restrict__ w, float* __restrict__ dpool, float* __restrict__ dw,
                   float* __restrict__ dbias, int B, int HW, int C, int classes) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_dl[HEAD_BROWS * HEAD_MAX_CLASSES];
  const int b0 = blockIdx.y * HEAD_BROWS;
  const int nb = min(HEAD_BROWS, B - b0);
  for (int i = threadIdx.x; i < nb * classes; i += blockDim.x) s_dl[i] = dlogits[(size_t)b0 * classes + i];
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    float wc[HEAD_MAX_CLASSES], gw[HEAD_MAX_CLASSES];
#pragma unroll
    for (int j = 0; j < HEAD_MAX_CLASSES; ++j) { wc[j] = j < classes ? w[(size_t)j * C + c] : 0.f; gw[j] = 0.f; }
    const float inv = 1.f / (float)HW;
    float pv[HEAD_BROWS], dr[HEAD_BROWS];
#pragma unroll
    for (int bb = 0; bb < HEAD_BROWS; ++bb) {
      const bool ok = bb < nb;
      pv[bb] = ok ? pooled[(size_t)(b0 + bb) * C + c] : 0.f;
      dr[bb] = ok ? drop[(size_t)(b0 + bb) * C + c] : 0.f;
    }
#pragma unroll
    for (int bb = 0; bb < HEAD_BROWS; ++bb) {
      if (bb < nb) {
        float dp = 0.f;
#pragma unroll
        for (int j = 0; j < HEAD_MAX_CLASSES; ++j) {
          if (j < classes) {
            const float d = s_dl[bb * classes + j];
            dp = fmaf(d, wc[j], dp);
            gw[j] = fmaf(d, pv[bb], gw[j]);
          }
        }
        dpool[(size_t)(b0 + bb) * C + c] = dp * dr[bb] * inv;
      }
    }
#pragma unroll
    for (int j = 0; j < HEAD_MAX_CLASSES; ++j)
      if (j < classes) atomicAdd(dw + (size_t)j * C + c, gw[j]);
  }
  if (blockIdx.x == 0 && threadIdx.x < classes) {
    float sum = 0.f;
    for (int bb = 0; bb < nb; ++bb) sum += s_dl[bb * classes + threadIdx.x];
    atomicAdd(dbias + threadIdx.x, sum);
  }
}

__global__ void __launch_bounds__(256)
head_bwd_mask_kernel(const __nv_bfloat16* __restrict__ raw, const float* __restrict__ coef,
                     const float* __restrict__ dpool, __nv_bfloat16* __restrict__ dz, float* __restrict__ sums,
                     int stride, int B, int HW, int C) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float s_red[];   // [2][C]
  const int CG = C >> 3;
  const int ipb = blockDim.x / CG;
  const int cg = threadIdx.x % CG, ti = threadIdx.x / CG;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();
  if (ti < ipb) {
    float sc[8], sh[8], s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { sc[j] = coef[cg * 8 + j]; sh[j] = coef[C + cg * 8 + j]; s0[j] = s1[j] = 0.f; }
    const long long rows = (long long)B * HW;
    for (long long r = (long long)blockIdx.x * ipb + ti; r < rows; r += (long long)gridDim.x * ipb) {
      const int b = (int)(r / HW);
      const uint4 v = ldg_stream(raw + r * C + cg * 8);
      const float x[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                          bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
      const float4 d0 = *reinterpret_cast<const float4*>(dpool + (size_t)b * C + cg * 8);
      const float4 d1 = *reinterpret_cast<const float4*>(dpool + (size_t)b * C + cg * 8 + 4);
      const float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(x[j], sc[j], sh[j]);
        o[j] = (z > 0.f && z < 6.f) ? d[j] : 0.f;
      }
      const uint4 pk = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                                  pack_bf16x2(o[6], o[7]));
      stg_stream(dz + r * C + cg * 8, pk);
      const float q[8] = {bf16_lo(pk.x), bf16_hi(pk.x), bf16_lo(pk.y), bf16_hi(pk.y),
                          bf16_lo(pk.z), bf16_hi(pk.z), bf16_lo(pk.w), bf16_hi(pk.w)};
#pragma unroll
      for (int j = 0; j < 8; ++j) { s0[j] += q[j]; s1[j] = fmaf(q[j], x[j], s1[j]); }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&s_red[cg * 8 + j], s0[j]);
      atomicAdd(&s_red[C + cg * 8 + j], s1[j]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(sums + i, s_red[i]);
    atomicAdd(sums + stride + i, s_red[C + i]);
  }
}

}  // namespace

int head_pool_launch(const void* raw, const float* coef, float* pooled, float* drop, int B, int HW, int C,
                     float p_drop, uint32_t seed, const int* step, cudaStream_t s, const BnFwdSrc* src) {
  dim3 grid((C + 255) / 256, B);
  B200DP_CUDA(launch_k(head_pool_kernel, grid, dim3(256), 0, s, static_cast<const __nv_bfloat16*>(raw), coef, pooled, drop, B,
                       HW, C, p_drop, seed, step, src ? *src : BnFwdSrc{}));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int head_fc_launch(const float* pooled, const float* w, const float* bias, float* logits, int B, int C, int classes,
                   cudaStream_t s) {
  B200DP_CUDA(launch_k(head_fc_kernel, dim3(B), dim3(32 * (classes < 16 ? classes : 16)), 0, s, pooled, w, bias, logits, C, classes));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int head_bwd_launch(const float* dlogits, const float* pooled, const float* drop, const float* w, float* dpool,
                    float* dw, float* dbias, const void* raw, const float* coef, void* dz, float* sums, int stride,
                    int B, int HW, int C, int classes, cudaStream_t s) {
  B200DP_REQUIRE(classes <= HEAD_MAX_CLASSES, "head: at most %d classes", HEAD_MAX_CLASSES);
  B200DP_CUDA(launch_k(head_bwd_fc_kernel, dim3((C + 127) / 128, (B + HEAD_BROWS - 1) / HEAD_BROWS), dim3(128), 0, s, dlogits,
                       pooled, drop, w, dpool, dw, dbias, B, HW, C, classes));
  B200DP_CHECK_LAUNCH();
  const int CG = C / 8;
  B200DP_REQUIRE(C % 8 == 0 && CG <= 256, "head: C=%d unsupported", C);
  const int threads = (256 / CG) * CG;
  const int ipb = threads / CG;
  const long long rows = (long long)B * HW;
  long long blocks = (rows + ipb - 1) / ipb;
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  B200DP_CUDA(launch_k(head_bwd_mask_kernel, dim3((unsigned)blocks), dim3(threads), 2 * C * sizeof(float), s,
                       static_cast<const __nv_bfloat16*>(raw), coef, (const float*)dpool, static_cast<__nv_bfloat16*>(dz), sums, stride,
                       B, HW, C));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_softmax_xent(const float* logits, const int64_t* labels, float* loss_out,
                                   float* dlogits_out, double* metrics, int batch, int classes,
                                   float mean_over, void* stream) {
  B200DP_REQUIRE(logits && labels && batch > 0 && classes > 0, "softmax_xent: bad arguments");
  const float inv = mean_over > 0.f ? 1.f / mean_over : 1.f;
  B200DP_CUDA(launch_k(softmax_xent_kernel, dim3((batch + 3) / 4), dim3(128), 0, static_cast<cudaStream_t>(stream), logits, labels,
                       loss_out, dlogits_out, metrics, batch, classes, inv));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

extern "C" int b200dp_softmax_topk(const float* logits, float* probs_out, float* topv_out, int32_t* topi_out,
                                   int batch, int classes, int k, void* stream) {
  B200DP_REQUIRE(logits && topv_out && topi_out && batch > 0, "softmax_topk: bad arguments");
  B200DP_REQUIRE(k >= 1 && k <= classes, "softmax_topk: k=%d outside [1, %d]", k, classes);
  softmax_topk_kernel<<<(batch + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, probs_out, topv_out, topi_out, batch, classes, k);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}
