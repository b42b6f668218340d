// Loss / metrics / inference tail (SURVEY W9 tail, W10; K10, K11, K15).
//
// Reference behaviour replaced: nn.CrossEntropyLoss (log_softmax + nll_loss), torch.max(outputs, 1),
// (preds == labels).sum().item() and loss.item() -- five ATen kernels and two host syncs per step
// (cifar10_mpi_mobilenet_224.py:178-184) -- and softmax + topk at inference
// (cifar10_128batch.py:170-172).  One warp per sample; loss sum / correct count / sample count are
// accumulated on the device as doubles and read once per epoch.
#include "common.cuh"

namespace b200dp {
namespace {

__global__ void __launch_bounds__(128)
softmax_xent_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                    float* __restrict__ loss_out, float* __restrict__ dlogits, double* __restrict__ metrics,
                    int batch, int classes, float inv_mean) {
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

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_softmax_xent(const float* logits, const int64_t* labels, float* loss_out,
                                   float* dlogits_out, double* metrics, int batch, int classes,
                                   float mean_over, void* stream) {
  B200DP_REQUIRE(logits && labels && batch > 0 && classes > 0, "softmax_xent: bad arguments");
  const float inv = mean_over > 0.f ? 1.f / mean_over : 1.f;
  softmax_xent_kernel<<<(batch + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, labels, loss_out, dlogits_out, metrics, batch, classes, inv);
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
