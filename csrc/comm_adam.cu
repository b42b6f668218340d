// Fused gradient all-reduce + average + Adam over peer memory (NVLink 5 / NVSwitch), sm_100a.
//
// Reference behaviour replaced (cifar10_mpi_mobilenet_224.py:176-180): DDP's bucketed NCCL
// all-reduce launched from autograd hooks, then torch.optim.Adam's multi-tensor kernels, then
// zero_grad.  Here that is ONE kernel per step:
//
//   flag barrier (st.release.sys / ld.acquire.sys on peer-mapped flag pads)
//   -> 16-byte loads of every rank's gradient vector in rank order 0..W-1 (same order on every
//      rank => bit-identical sums, so replicas stay bit-identical without a broadcast)
//   -> * 1/W -> Adam on fp32 master weights + moments (all in registers)
//   -> optional bf16 working-weight refresh
//   -> flag barrier -> zero own gradients.
//
// one-shot (algo 0): every rank reduces and updates everything; traffic in = (W-1) * bytes.
// two-shot (algo 1): rank r reduces + updates slice r only, then stores the new weights into
//                    every peer (reduce-scatter + all-gather); traffic = 2 (W-1)/W * bytes.
// No host involvement per step: learning rate, Adam step count and barrier epochs live in
// device memory, so the launch is CUDA-graph capturable.
#include <cmath>

#include "common.cuh"

namespace b200dp {

namespace {

constexpr int kThreads = 512;
constexpr int kDefaultBlocks = 148;
constexpr unsigned long long kDefaultSpin = 400ull * 1000 * 1000;  // polls; seconds, not minutes

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer data is read with system-scope strong loads: coherent with the owner's L2, never served
// from a stale local L1 line.
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_bf4(const __nv_bfloat16* p) {
  uint32_t a, b;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
  return make_float4(bf16_lo(a), bf16_hi(a), bf16_lo(b), bf16_hi(b));
}
__device__ __forceinline__ void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_peer_u2(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}

// All threads of the block call this.  Block b of this rank meets block b of every peer.
__device__ __forceinline__ void flag_barrier(const b200dp_comm_adam& P, uint32_t target,
                                             unsigned long long limit) {
  __syncthreads();
  if (threadIdx.x < (unsigned)P.world) {
    __threadfence_system();
    uint32_t* remote = P.flags[threadIdx.x] + (size_t)blockIdx.x * B200DP_MAX_WORLD + P.rank;
    st_release_sys(remote, target);
    const uint32_t* mine = P.flags[P.rank] + (size_t)blockIdx.x * B200DP_MAX_WORLD + threadIdx.x;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - target) < 0) {
      if (++spins > limit) {  // a peer is dead or stalled: report, do not hang the GPU
        atomicExch(P.error_flag, 1 + (int)threadIdx.x);
        break;
      }
      if ((spins & 63) == 63) __nanosleep(64);
    }
  }
  __syncthreads();
}

template <int GD>
__device__ __forceinline__ float4 load_grad(const void* base, long long i4) {
  if constexpr (GD == 0) {
    return ld_peer_f4(reinterpret_cast<const float*>(base) + i4 * 4);
  } else {
    return ld_peer_bf4(reinterpret_cast<const __nv_bfloat16*>(base) + i4 * 4);
  }
}

template <int GD>
__device__ __forceinline__ void zero_grad(void* base, long long i4) {
  if constexpr (GD == 0) {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + i4 * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(base) + i4 * 4) = make_uint2(0u, 0u);
  }
}

__device__ __forceinline__ void adam1(float g, float& p, float& m, float& v, float b1, float b2, float ob1,
                                      float ob2, float eps, float wd, float step_size, float isb2) {
  g = fmaf(wd, p, g);
  m = fmaf(b1, m, ob1 * g);
  v = fmaf(b2, v, ob2 * g * g);
  const float denom = fmaf(sqrtf(v), isb2, eps);
  p = p - step_size * __fdividef(m, denom);
}

// WT = compile-time world size (1,2,4,8) or 0 = runtime loop.
template <int GD, int ALGO, int WT>
__global__ void __launch_bounds__(kThreads)
comm_adam_kernel(const __grid_constant__ b200dp_comm_adam P) {
  __shared__ float s_scal[2];
  const int W = (WT > 0) ? WT : P.world;
  const int rank = P.rank;
  const uint32_t epoch = P.epoch_state[blockIdx.x];
  const int step = P.step_state[blockIdx.x];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;

  if (threadIdx.x == 0) {
    const double t = (double)(step + 1);
    const double bc1 = 1.0 - pow(P.beta1, t);
    const double bc2 = 1.0 - pow(P.beta2, t);
    s_scal[0] = (float)((double)(*P.lr) / bc1);
    s_scal[1] = (float)(1.0 / sqrt(bc2));
  }
  if (W > 1) {
    flag_barrier(P, epoch + 1, limit);  // every peer has finished its backward pass
  } else {
    __syncthreads();
  }
  const float step_size = s_scal[0], isb2 = s_scal[1];
  const float b1 = (float)P.beta1, b2 = (float)P.beta2, eps = P.eps, wd = P.weight_decay;
  const float ob1 = (float)(1.0 - P.beta1), ob2 = (float)(1.0 - P.beta2);
  const float scale = 1.0f / (float)W;

  const long long n4 = P.numel >> 2;
  long long lo = 0, hi = n4;
  long long per = n4;
  if (ALGO == 1) {
    per = (n4 + W - 1) / W;
    lo = per * rank;
    hi = lo + per < n4 ? lo + per : n4;
  }
  const long long stride = (long long)gridDim.x * kThreads;
  float* __restrict__ myp = P.params[rank];
  float* __restrict__ mm = P.m;
  float* __restrict__ vv = P.v;

  for (long long i = lo + (long long)blockIdx.x * kThreads + threadIdx.x; i < hi; i += stride) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if constexpr (WT > 0) {
      float4 g[WT];
#pragma unroll
      for (int r = 0; r < WT; ++r) g[r] = load_grad<GD>(P.grads[r], i);   // all loads in flight first
#pragma unroll
      for (int r = 0; r < WT; ++r) {                                       // fixed order 0..W-1
        acc.x += g[r].x; acc.y += g[r].y; acc.z += g[r].z; acc.w += g[r].w;
      }
    } else {
      for (int r = 0; r < W; ++r) {
        const float4 g = load_grad<GD>(P.grads[r], i);
        acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
      }
    }
    acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
    if (!P.apply_adam) {
      *reinterpret_cast<float4*>(P.out_avg + i * 4) = acc;
      continue;
    }
    float4 p = *reinterpret_cast<const float4*>(myp + i * 4);
    float4 m = *reinterpret_cast<const float4*>(mm + i * 4);
    float4 v = *reinterpret_cast<const float4*>(vv + i * 4);
    adam1(acc.x, p.x, m.x, v.x, b1, b2, ob1, ob2, eps, wd, step_size, isb2);
    adam1(acc.y, p.y, m.y, v.y, b1, b2, ob1, ob2, eps, wd, step_size, isb2);
    adam1(acc.z, p.z, m.z, v.z, b1, b2, ob1, ob2, eps, wd, step_size, isb2);
    adam1(acc.w, p.w, m.w, v.w, b1, b2, ob1, ob2, eps, wd, step_size, isb2);
    *reinterpret_cast<float4*>(mm + i * 4) = m;
    *reinterpret_cast<float4*>(vv + i * 4) = v;
    const uint32_t w01 = pack_bf16x2(p.x, p.y), w23 = pack_bf16x2(p.z, p.w);
    if (ALGO == 0) {
      *reinterpret_cast<float4*>(myp + i * 4) = p;
      if (P.work_bf16[rank] != nullptr)
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(P.work_bf16[rank]) + i * 4) =
            make_uint2(w01, w23);
    } else {
      for (int r = 0; r < W; ++r) {  // all-gather: push the updated slice into every replica
        st_peer_f4(P.params[r] + i * 4, p);
        if (P.work_bf16[r] != nullptr)
          st_peer_u2(reinterpret_cast<__nv_bfloat16*>(P.work_bf16[r]) + i * 4, w01, w23);
      }
    }
  }

  const bool second = (W > 1) && (P.zero_grads || ALGO == 1);
  if (second) flag_barrier(P, epoch + 2, limit);  // peers are done reading my gradients / writing my weights
  if (P.zero_grads) {
    void* mine = const_cast<void*>(P.grads[rank]);
    if (ALGO == 0 || W == 1) {
      for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n4; i += stride)
        zero_grad<GD>(mine, i);
    } else {
      for (int q = 0; q < W; ++q) {  // the part of slice q that block b of rank q has read
        const long long qlo = per * q;
        const long long qhi = qlo + per < n4 ? qlo + per : n4;
        for (long long i = qlo + (long long)blockIdx.x * kThreads + threadIdx.x; i < qhi; i += stride)
          zero_grad<GD>(mine, i);
      }
    }
  }
  if (threadIdx.x == 0) {
    P.epoch_state[blockIdx.x] = epoch + 2;
    if (P.apply_adam) P.step_state[blockIdx.x] = step + 1;
  }
}

__global__ void __launch_bounds__(kThreads)
comm_barrier_kernel(const __grid_constant__ b200dp_comm_adam P) {
  const uint32_t epoch = P.epoch_state[blockIdx.x];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (P.world > 1) flag_barrier(P, epoch + 2, limit);
  if (threadIdx.x == 0) P.epoch_state[blockIdx.x] = epoch + 2;
}

struct BcastArgs {
  float* dst[B200DP_MAX_WORLD];
  const float* src;
  long long n4;
  int root;
};

__global__ void __launch_bounds__(kThreads)
comm_broadcast_kernel(const __grid_constant__ b200dp_comm_adam P, const __grid_constant__ BcastArgs A) {
  const uint32_t epoch = P.epoch_state[blockIdx.x];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (P.world > 1) flag_barrier(P, epoch + 1, limit);  // nobody is still using dst
  if (P.rank == A.root) {
    const long long stride = (long long)gridDim.x * kThreads;
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < A.n4; i += stride) {
      const float4 v = *reinterpret_cast<const float4*>(A.src + i * 4);
      for (int r = 0; r < P.world; ++r)
        if (A.dst[r] + i * 4 != A.src + i * 4) st_peer_f4(A.dst[r] + i * 4, v);
    }
  }
  if (P.world > 1) flag_barrier(P, epoch + 2, limit);  // root's stores have landed everywhere
  if (threadIdx.x == 0) P.epoch_state[blockIdx.x] = epoch + 2;
}

int grid_for(const b200dp_comm_adam& a) {
  long long n4 = a.numel >> 2;
  if (a.algo == 1 && a.world > 1) n4 = (n4 + a.world - 1) / a.world;
  long long blocks = (n4 + (long long)kThreads * 2 - 1) / ((long long)kThreads * 2);
  int cap = a.max_blocks > 0 ? a.max_blocks : kDefaultBlocks;
  if (cap > B200DP_COMM_MAX_BLOCKS) cap = B200DP_COMM_MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

int validate(const b200dp_comm_adam* a) {
  B200DP_REQUIRE(a != nullptr, "comm_adam: null args");
  B200DP_REQUIRE(a->world >= 1 && a->world <= B200DP_MAX_WORLD, "comm_adam: world %d out of range", a->world);
  B200DP_REQUIRE(a->rank >= 0 && a->rank < a->world, "comm_adam: rank %d outside world %d", a->rank, a->world);
  B200DP_REQUIRE(a->numel > 0 && (a->numel & 3) == 0, "comm_adam: numel %lld must be a positive multiple of 4",
                 a->numel);
  B200DP_REQUIRE(a->epoch_state && a->step_state && a->error_flag, "comm_adam: state pointers missing");
  for (int r = 0; r < a->world; ++r) {
    B200DP_REQUIRE(a->grads[r] != nullptr, "comm_adam: grads[%d] is null", r);
    if (a->world > 1) B200DP_REQUIRE(a->flags[r] != nullptr, "comm_adam: flags[%d] is null", r);
  }
  return B200DP_OK;
}

template <int GD, int ALGO>
int launch_w(const b200dp_comm_adam& a, int grid, cudaStream_t s) {
  switch (a.world) {
    case 1: comm_adam_kernel<GD, ALGO, 1><<<grid, kThreads, 0, s>>>(a); break;
    case 2: comm_adam_kernel<GD, ALGO, 2><<<grid, kThreads, 0, s>>>(a); break;
    case 4: comm_adam_kernel<GD, ALGO, 4><<<grid, kThreads, 0, s>>>(a); break;
    case 8: comm_adam_kernel<GD, ALGO, 8><<<grid, kThreads, 0, s>>>(a); break;
    default: comm_adam_kernel<GD, ALGO, 0><<<grid, kThreads, 0, s>>>(a); break;
  }
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_comm_adam_grid(const b200dp_comm_adam* a) {
  if (validate(a) != B200DP_OK) return B200DP_EINVAL;
  return grid_for(*a);
}

extern "C" int b200dp_comm_adam_step(const b200dp_comm_adam* a, void* stream) {
  if (int rc = validate(a)) return rc;
  if (a->apply_adam) {
    B200DP_REQUIRE(a->params[a->rank] && a->m && a->v && a->lr, "comm_adam: Adam state pointers missing");
    if (a->algo == 1)
      for (int r = 0; r < a->world; ++r)
        B200DP_REQUIRE(a->params[r] != nullptr, "comm_adam: two-shot needs params[%d]", r);
  } else {
    B200DP_REQUIRE(a->out_avg != nullptr, "comm_adam: out_avg missing");
    B200DP_REQUIRE(a->algo == 0, "comm_adam: plain all-reduce is one-shot only");
  }
  B200DP_REQUIRE(a->grad_dtype == 0 || a->grad_dtype == 1, "comm_adam: grad_dtype must be 0 or 1");
  const int grid = grid_for(*a);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a->grad_dtype == 0) {
    return a->algo == 0 ? launch_w<0, 0>(*a, grid, s) : launch_w<0, 1>(*a, grid, s);
  }
  return a->algo == 0 ? launch_w<1, 0>(*a, grid, s) : launch_w<1, 1>(*a, grid, s);
}

extern "C" int b200dp_comm_barrier(const b200dp_comm_adam* a, void* stream) {
  if (int rc = validate(a)) return rc;
  comm_barrier_kernel<<<grid_for(*a), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(*a);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

extern "C" int b200dp_comm_broadcast(const b200dp_comm_adam* a, int root, float* const* dst_by_rank,
                                     const float* src, long long numel, void* stream) {
  if (int rc = validate(a)) return rc;
  B200DP_REQUIRE(root >= 0 && root < a->world, "broadcast: root %d outside world", root);
  B200DP_REQUIRE(numel > 0 && (numel & 3) == 0, "broadcast: numel must be a positive multiple of 4");
  BcastArgs b{};
  for (int r = 0; r < a->world; ++r) {
    B200DP_REQUIRE(dst_by_rank[r] != nullptr, "broadcast: dst[%d] is null", r);
    b.dst[r] = dst_by_rank[r];
  }
  b.src = src;
  b.n4 = numel >> 2;
  b.root = root;
  comm_broadcast_kernel<<<grid_for(*a), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(*a, b);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}
