// BatchNorm finalisation folded into the FIRST CONSUMER of the coefficients.
//
// Producers leave per-channel sums in global memory.  Turning them into (scale, shift) or (ca, cb, cc) is a handful
// of flops per channel; as separate kernels that was 104 launches per training step, ~3 us each even inside a CUDA
// graph.  Instead the first kernel that needs the coefficients computes them in its prologue, every CTA for the
// channels it uses (a few dozen to a few hundred rsqrt's), and ONE designated thread per channel also performs the
// side effects: the coefficient array for later readers, saved mean / inv-std, running statistics, dgamma / dbeta.
// (Folding the finalisation into the PRODUCER's tail -- "last CTA done" ticket -- was tried first and is slower: it
// puts a fence -> atomic -> reload -> store chain on the critical path of every producer; see DESIGN.md.)
//
// A descriptor with sums == nullptr means "read the coefficient array as before" (evaluation mode, C API).
// Reference behaviour replaced: the statistics half of cuDNN batch-norm forward / backward.
#pragma once

#include "common.cuh"

namespace b200dp {

struct BnFwdSrc {
  const float* sums = nullptr;   // [2][stride]: sum y, sum y^2 of the producing conv
  int stride = 0, C = 0;
  float count = 1.f, eps = 0.f, momentum = 0.f;
  const float* gamma = nullptr;
  const float* beta = nullptr;
  float* rmean = nullptr;        // side effects (designated writer only)
  float* rvar = nullptr;
  float* coef = nullptr;         // [2][C] scale, shift
  float* saved = nullptr;        // [2][C] mean, inv-std
};

struct BnBwdSrc {
  const float* sums = nullptr;   // [2][stride]: sum dz, sum dz*x
  int stride = 0, C = 0;
  float count = 1.f;
  const float* saved = nullptr;
  const float* gamma = nullptr;
  float* coef = nullptr;         // [3][C] ca, cb, cc (side effect)
  float* dgamma = nullptr;       // accumulated (side effect)
  float* dbeta = nullptr;
};

#ifdef __CUDACC__

__device__ __forceinline__ void bn_fwd_channel(const BnFwdSrc& b, int c, bool writer, float& sc, float& sh) {
  const float mean = b.sums[c] / b.count;
  const float var = fmaxf(b.sums[b.stride + c] / b.count - mean * mean, 0.f);
  const float inv = rsqrtf(var + b.eps);
  sc = b.gamma[c] * inv;
  sh = b.beta[c] - mean * sc;
  if (writer) {
    b.coef[c] = sc;
    b.coef[b.C + c] = sh;
    b.saved[c] = mean;
    b.saved[b.C + c] = inv;
    if (b.rmean != nullptr) {   // torch: running = (1-m) running + m batch, variance unbiased
      b.rmean[c] = (1.f - b.momentum) * b.rmean[c] + b.momentum * mean;
      const float unb = b.count > 1.f ? var * b.count / (b.count - 1.f) : var;
      b.rvar[c] = (1.f - b.momentum) * b.rvar[c] + b.momentum * unb;
    }
  }
}

// dx = ca*dz + cb*x + cc
__device__ __forceinline__ void bn_bwd_channel(const BnBwdSrc& b, int c, bool writer, float& ca, float& cb, float& cc) {
  const float mean = b.saved[c], r = b.saved[b.C + c], g = b.gamma[c];
  const float db = b.sums[c];
  const float dg = r * (b.sums[b.stride + c] - mean * db);
  ca = g * r;
  cb = -ca * r * dg / b.count;
  cc = -ca * db / b.count - cb * mean;
  if (writer) {
    b.coef[c] = ca;
    b.coef[b.C + c] = cb;
    b.coef[2 * b.C + c] = cc;
    b.dgamma[c] += dg;
    b.dbeta[c] += db;
  }
}

#endif  // __CUDACC__

}  // namespace b200dp
