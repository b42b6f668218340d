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
// NVLS     (algo 2): two-shot through the NVSwitch multicast mapping of the arena (vmm.cu):
//                    `multimem.ld_reduce.add.v4.f32` returns the sum over all ranks computed in the
//                    switch, `multimem.st.v4.f32` writes the updated slice into every replica.
// A launch covers the element range [range_lo, range_hi): the executor (engine.cu) launches one
// per gradient bucket from a branch of the step graph, so buckets reduce while backward is
// still producing the next ones (DDP's Reducer, ref :142-145, :179).  Such launches use a small
// fixed grid: block b only ever waits for block b of its peers, never for another block of its
// own GPU, so a grid that is only partly resident next to the compute kernels still completes.
// A barrier that times out (dead or stalled peer) makes the block return WITHOUT touching
// weights, moments or gradients, and raises both a device flag (sticky: later launches return at
// once) and, when given, a host-mapped mirror the trainer polls without synchronising.
// No host involvement per step: learning rate, Adam step count and barrier epochs live in
// device memory, so the launch is CUDA-graph capturable.
#include <cmath>
#include <cstdlib>

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

__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ void raise_error(const b200dp_comm_adam& P, int code) {
  atomicExch(P.error_flag, code);
  for (int q = 0; q < P.world; ++q)          // tell everybody: a peer that arrives late must not update anything either
    if (P.abort_words[q] != nullptr) st_release_sys(P.abort_words[q], (uint32_t)code);
  if (P.error_host != nullptr) {
    *reinterpret_cast<volatile int*>(P.error_host) = code;
    __threadfence_system();
  }
}

// All threads of the block call this.  Block b of this rank meets block b of every peer.
// Returns false (for the whole block) when a peer did not arrive within `limit` polls.
__device__ __forceinline__ bool flag_barrier(const b200dp_comm_adam& P, uint32_t target,
                                             unsigned long long limit) {
  int failed = 0;
  __syncthreads();
  if (threadIdx.x < (unsigned)P.world) {
    __threadfence_system();
    uint32_t* remote = P.flags[threadIdx.x] + (size_t)blockIdx.x * B200DP_MAX_WORLD + P.rank;
    st_release_sys(remote, target);
    const uint32_t* mine = P.flags[P.rank] + (size_t)blockIdx.x * B200DP_MAX_WORLD + threadIdx.x;
    unsigned long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - target) < 0) {
      if (++spins > limit) {  // a peer is dead or stalled: report, do not hang the GPU
        raise_error(P, 1 + (int)threadIdx.x);
        failed = 1;
        break;
      }
      if ((spins & 63) == 63) __nanosleep(64);
    }
  }
  return __syncthreads_or(failed) == 0;
}

// sticky failure: once any launch on this rank has raised the flag, or any PEER has given up on a barrier (its abort
// word in this rank's arena), launches do nothing.  Checked at the start and again after the first barrier.
__device__ __forceinline__ bool already_failed(const b200dp_comm_adam& P) {
  int bad = 0;
  if (threadIdx.x == 0) {
    bad = *reinterpret_cast<volatile int*>(P.error_flag);
    if (bad == 0 && P.world > 1 && P.abort_words[P.rank] != nullptr) {
      const uint32_t a = ld_acquire_sys(P.abort_words[P.rank]);
      if (a != 0) {
        bad = (int)a;
        atomicExch(P.error_flag, bad);
        if (P.error_host != nullptr) {
          *reinterpret_cast<volatile int*>(P.error_host) = bad;
          __threadfence_system();
        }
      }
    }
  }
  return __syncthreads_or(bad) != 0;
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
  const long long n4_all = P.numel >> 2;
  const long long r_lo = P.range_hi > 0 ? (P.range_lo >> 2) : 0;
  const long long r_hi = P.range_hi > 0 ? (P.range_hi >> 2) : n4_all;
  const int step = P.step_state[0];     // advanced by the last block of the step's LAST launch (bump_step)
  const bool barriers = W > 1 && !P.skip_barriers;
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (already_failed(P)) return;

  if (threadIdx.x == 0) {
    const double t = (double)(step + 1);
    const double bc1 = 1.0 - pow(P.beta1, t);
    const double bc2 = 1.0 - pow(P.beta2, t);
    s_scal[0] = (float)((double)(*P.lr) / bc1);
    s_scal[1] = (float)(1.0 / sqrt(bc2));
  }
  if (barriers) {
    if (!flag_barrier(P, epoch + 1, limit)) return;  // every peer has finished producing this range
    if (already_failed(P)) return;                   // ... unless one of them gave up waiting for us in the meantime
  } else {
    __syncthreads();      // (skip_barriers: a one-CTA barrier launch ahead of this one did the waiting)
  }
  const float step_size = s_scal[0], isb2 = s_scal[1];
  const float b1 = (float)P.beta1, b2 = (float)P.beta2, eps = P.eps, wd = P.weight_decay;
  const float ob1 = (float)(1.0 - P.beta1), ob2 = (float)(1.0 - P.beta2);
  const float scale = 1.0f / (float)W;

  const long long n4 = r_hi - r_lo;
  long long lo = r_lo, hi = r_hi;
  long long per = n4;
  if (ALGO >= 1) {
    per = (n4 + W - 1) / W;
    lo = r_lo + per * rank;
    hi = lo + per < r_hi ? lo + per : r_hi;
  }
  const long long stride = (long long)gridDim.x * blockDim.x;
  float* __restrict__ myp = P.params[rank];
  float* __restrict__ mm = P.m;
  float* __restrict__ vv = P.v;

  // one 16-byte element: (already summed) gradient -> average -> Adam -> store
  auto update = [&](long long i, float4 acc) {
    if constexpr (ALGO == 2) {
    } else if constexpr (WT > 0) {
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
      return;
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
    } else if (ALGO == 2) {
      mc_st_f4(P.mc_params + i * 4, p);   // one store, the switch replicates it into every rank's weights
    } else {
      for (int r = 0; r < W; ++r) {  // all-gather: push the updated slice into every replica
        st_peer_f4(P.params[r] + i * 4, p);
        if (P.work_bf16[r] != nullptr)
          st_peer_u2(reinterpret_cast<__nv_bfloat16*>(P.work_bf16[r]) + i * 4, w01, w23);
      }
    }
  };
  const long long first = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if constexpr (ALGO == 2) {
    // four independent switch reductions in flight per thread before any of the (dependent) Adam chains starts
    for (long long i = first; i < hi; i += 4 * stride) {
      float4 acc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < hi) acc[u] = mc_ld_reduce_f4(reinterpret_cast<const float*>(P.mc_grads) + (i + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i + u * stride < hi) update(i + u * stride, acc[u]);
    }
  } else {
    for (long long i = first; i < hi; i += stride) update(i, make_float4(0.f, 0.f, 0.f, 0.f));
  }

  if (W > 1) __threadfence_system();   // my stores into the peers' weights are performed before anything that follows
  const bool second = barriers && (P.zero_grads || ALGO >= 1);
  if (second && !flag_barrier(P, epoch + 2, limit)) return;  // peers are done reading my gradients / writing my weights
  if (P.zero_grads) {
    void* mine = const_cast<void*>(P.grads[rank]);
    if (ALGO == 0 || W == 1) {
      for (long long i = r_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < r_hi; i += stride)
        zero_grad<GD>(mine, i);
    } else {
      for (int q = 0; q < W; ++q) {  // the part of slice q that block b of rank q has read
        const long long qlo = r_lo + per * q;
        const long long qhi = qlo + per < r_hi ? qlo + per : r_hi;
        for (long long i = qlo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < qhi; i += stride)
          zero_grad<GD>(mine, i);
      }
    }
  }
  if (P.zero_hi > P.zero_lo) {      // gradients of earlier buckets of this step: every peer is done with them by now
    float* mine = reinterpret_cast<float*>(const_cast<void*>(P.grads[rank]));
    for (long long i = (P.zero_lo >> 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (P.zero_hi >> 2); i += stride)
      zero_grad<GD>(mine, i);
  }
  if (threadIdx.x == 0) {
    if (barriers) P.epoch_state[blockIdx.x] = epoch + 2;
    if (P.apply_adam && P.bump_step) {
      // the step count advances when the LAST block of the step's last launch is done: every block has read it by then
      __threadfence();
      if (atomicAdd(&P.step_state[1], 1) == (int)gridDim.x - 1) {
        P.step_state[1] = 0;
        P.step_state[0] = step + 1;
      }
    }
  }
}

// Barrier across ranks.  With one CTA this is how the early in-graph buckets wait: the waiting costs one CTA slot,
// not the grid of the data launch that follows.
__global__ void __launch_bounds__(kThreads)
comm_barrier_kernel(const __grid_constant__ b200dp_comm_adam P) {
  const uint32_t epoch = P.epoch_state[blockIdx.x];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (already_failed(P)) return;
  if (P.world > 1 && !flag_barrier(P, epoch + 2, limit)) return;
  if (threadIdx.x == 0) P.epoch_state[blockIdx.x] = epoch + 2;
}

struct BcastArgs {
  float* dst[B200DP_MAX_WORLD];
  float* mc_dst;              // multicast alias of dst (NVLS): one store reaches every rank
  const float* src;
  long long n4;
  int root;
};

__global__ void __launch_bounds__(kThreads)
comm_broadcast_kernel(const __grid_constant__ b200dp_comm_adam P, const __grid_constant__ BcastArgs A) {
  const uint32_t epoch = P.epoch_state[blockIdx.x];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (already_failed(P)) return;
  if (P.world > 1 && !flag_barrier(P, epoch + 1, limit)) return;  // nobody is still using dst
  if (P.rank == A.root) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < A.n4; i += stride) {
      const float4 v = *reinterpret_cast<const float4*>(A.src + i * 4);
      if (A.mc_dst != nullptr) {
        mc_st_f4(A.mc_dst + i * 4, v);
      } else {
        for (int r = 0; r < P.world; ++r)
          if (A.dst[r] + i * 4 != A.src + i * 4) st_peer_f4(A.dst[r] + i * 4, v);
      }
    }
  }
  if (P.world > 1 && !flag_barrier(P, epoch + 2, limit)) return;  // root's stores have landed everywhere
  if (threadIdx.x == 0) P.epoch_state[blockIdx.x] = epoch + 2;
}

// All-reduce of k <= 32 doubles (sum or max): every rank stores its values into row `rank` of every peer's slot
// table, barrier, every rank combines the rows of its own table in rank order (bit-identical results), barrier.
// One block; shares flag row 0 and epoch slot 0 with the kernels above (same stream, same order on every rank).
__global__ void __launch_bounds__(64)
comm_reduce_f64_kernel(const __grid_constant__ b200dp_comm_adam P, const double* in, double* out, int k, int op) {
  const uint32_t epoch = P.epoch_state[0];
  const unsigned long long limit = P.spin_limit ? P.spin_limit : kDefaultSpin;
  if (already_failed(P)) return;
  const int W = P.world;
  if (W == 1) {
    if ((int)threadIdx.x < k) out[threadIdx.x] = in[threadIdx.x];
    return;
  }
  if ((int)threadIdx.x < k) {
    const double v = in[threadIdx.x];
    for (int r = 0; r < W; ++r) {
      double* dst = P.slots[r] + (size_t)P.rank * B200DP_REDUCE_MAX + threadIdx.x;
      asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(dst), "d"(v) : "memory");
    }
  }
  if (!flag_barrier(P, epoch + 1, limit)) return;
  if ((int)threadIdx.x < k) {
    double acc = 0.0;
    for (int r = 0; r < W; ++r) {
      double v;
      const double* src = P.slots[P.rank] + (size_t)r * B200DP_REDUCE_MAX + threadIdx.x;
      asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(src) : "memory");
      acc = r == 0 ? v : (op == 0 ? acc + v : fmax(acc, v));
    }
    out[threadIdx.x] = acc;
  }
  if (!flag_barrier(P, epoch + 2, limit)) return;       // nobody overwrites a table that is still being read
  if (threadIdx.x == 0) P.epoch_state[0] = epoch + 2;
}

// threads per CTA: 512 by default; bucket launches from the step graph (exact_blocks > 0) may use fewer so that a
// comm CTA fits into the registers / thread slots the persistent compute CTAs leave free (B200DP_COMM_THREADS)
int threads_for(const b200dp_comm_adam& a) {
  static int small = [] {
    const char* v = getenv("B200DP_COMM_THREADS");
    const int t = v ? atoi(v) : 0;
    return (t >= 32 && t <= kThreads && t % 32 == 0) ? t : 0;
  }();
  return (a.exact_blocks > 0 && small > 0) ? small : kThreads;
}

int grid_for(const b200dp_comm_adam& a);
int validate(const b200dp_comm_adam* a);
}  // namespace
int comm_barrier_launch(const b200dp_comm_adam& a, int blocks, cudaStream_t s);
namespace {

int grid_for(const b200dp_comm_adam& a) {
  if (a.exact_blocks > 0) return a.exact_blocks < B200DP_COMM_MAX_BLOCKS ? a.exact_blocks : B200DP_COMM_MAX_BLOCKS;
  // sized by the whole range, not by the 1/W slice a rank reduces: the gradient clearing at the end covers the whole
  // range on every rank, and with one 16-byte element per thread the reduce + Adam loop is a single latency chain
  long long n4 = (a.range_hi > 0 ? a.range_hi - a.range_lo : a.numel) >> 2;
  if (a.zero_hi > a.zero_lo) n4 += (a.zero_hi - a.zero_lo) >> 2;
  long long blocks = (n4 + (long long)kThreads * 8 - 1) / ((long long)kThreads * 8);
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
  B200DP_REQUIRE(a->range_hi == 0 || (a->range_lo >= 0 && a->range_lo < a->range_hi && a->range_hi <= a->numel &&
                                      (a->range_lo & 3) == 0 && (a->range_hi & 3) == 0),
                 "comm_adam: range [%lld, %lld) must be a non-empty multiple-of-4 sub-range of [0, %lld)", a->range_lo,
                 a->range_hi, a->numel);
  B200DP_REQUIRE(a->epoch_state && a->step_state && a->error_flag, "comm_adam: state pointers missing");
  B200DP_REQUIRE(a->zero_hi == 0 || (a->zero_lo >= 0 && a->zero_lo < a->zero_hi && a->zero_hi <= a->numel &&
                                     (a->zero_lo & 3) == 0 && (a->zero_hi & 3) == 0 && a->grad_dtype == 0),
                 "comm_adam: bad extra zero range [%lld, %lld)", a->zero_lo, a->zero_hi);
  for (int r = 0; r < a->world; ++r) {
    B200DP_REQUIRE(a->grads[r] != nullptr, "comm_adam: grads[%d] is null", r);
    if (a->world > 1) B200DP_REQUIRE(a->flags[r] != nullptr, "comm_adam: flags[%d] is null", r);
  }
  return B200DP_OK;
}

template <int GD, int ALGO>
int launch_w(const b200dp_comm_adam& a, int grid, cudaStream_t s) {
  const int threads = threads_for(a);
  switch (a.world) {
    case 1: comm_adam_kernel<GD, ALGO, 1><<<grid, threads, 0, s>>>(a); break;
    case 2: comm_adam_kernel<GD, ALGO, 2><<<grid, threads, 0, s>>>(a); break;
    case 4: comm_adam_kernel<GD, ALGO, 4><<<grid, threads, 0, s>>>(a); break;
    case 8: comm_adam_kernel<GD, ALGO, 8><<<grid, threads, 0, s>>>(a); break;
    default: comm_adam_kernel<GD, ALGO, 0><<<grid, threads, 0, s>>>(a); break;
  }
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace

int comm_adam_launch(const b200dp_comm_adam& a, cudaStream_t s) {
  if (int rc = validate(&a)) return rc;
  if (a.apply_adam) {
    B200DP_REQUIRE(a.params[a.rank] && a.m && a.v && a.lr, "comm_adam: Adam state pointers missing");
    if (a.algo == 1)
      for (int r = 0; r < a.world; ++r)
        B200DP_REQUIRE(a.params[r] != nullptr, "comm_adam: two-shot needs params[%d]", r);
  } else {
    B200DP_REQUIRE(a.out_avg != nullptr, "comm_adam: out_avg missing");
    B200DP_REQUIRE(a.algo == 0, "comm_adam: plain all-reduce is one-shot only");
  }
  B200DP_REQUIRE(a.grad_dtype == 0 || a.grad_dtype == 1, "comm_adam: grad_dtype must be 0 or 1");
  B200DP_REQUIRE(a.algo >= 0 && a.algo <= 2, "comm_adam: algo must be 0, 1 or 2");
  int algo = a.algo;
  if (algo == 2) {
    B200DP_REQUIRE(a.grad_dtype == 0, "comm_adam: the NVLS path reduces fp32 gradients");
    if (a.world == 1) algo = 0;
    else B200DP_REQUIRE(a.mc_grads && a.mc_params, "comm_adam: NVLS needs the multicast addresses");
  }
  const int grid = grid_for(a);
  int rc;
  if (a.grad_dtype == 0) {
    rc = algo == 0 ? launch_w<0, 0>(a, grid, s) : algo == 1 ? launch_w<0, 1>(a, grid, s) : launch_w<0, 2>(a, grid, s);
  } else {
    rc = algo == 0 ? launch_w<1, 0>(a, grid, s) : launch_w<1, 1>(a, grid, s);
  }
  return rc;
}

int comm_barrier_launch(const b200dp_comm_adam& a, int blocks, cudaStream_t s) {
  if (int rc = validate(&a)) return rc;
  comm_barrier_kernel<<<blocks, kThreads, 0, s>>>(a);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_comm_adam_grid(const b200dp_comm_adam* a) {
  if (validate(a) != B200DP_OK) return B200DP_EINVAL;
  return grid_for(*a);
}

extern "C" int b200dp_comm_adam_step(const b200dp_comm_adam* a, void* stream) {
  B200DP_REQUIRE(a != nullptr, "comm_adam: null args");
  return comm_adam_launch(*a, static_cast<cudaStream_t>(stream));
}

extern "C" int b200dp_comm_barrier(const b200dp_comm_adam* a, void* stream) {
  if (int rc = validate(a)) return rc;
  return comm_barrier_launch(*a, grid_for(*a), static_cast<cudaStream_t>(stream));
}

extern "C" int b200dp_comm_broadcast(const b200dp_comm_adam* a, int root, float* const* dst_by_rank, float* mc_dst,
                                     const float* src, long long numel, void* stream) {
  if (int rc = validate(a)) return rc;
  B200DP_REQUIRE(root >= 0 && root < a->world, "broadcast: root %d outside world", root);
  B200DP_REQUIRE(numel > 0 && (numel & 3) == 0, "broadcast: numel must be a positive multiple of 4");
  BcastArgs b{};
  for (int r = 0; r < a->world; ++r) {
    B200DP_REQUIRE(dst_by_rank[r] != nullptr, "broadcast: dst[%d] is null", r);
    b.dst[r] = dst_by_rank[r];
  }
  b.mc_dst = a->world > 1 ? mc_dst : nullptr;
  b.src = src;
  b.n4 = numel >> 2;
  b.root = root;
  b200dp_comm_adam g = *a;          // grid from the broadcast size, not from the optimiser's vector
  g.numel = numel; g.range_lo = g.range_hi = 0; g.algo = 0; g.exact_blocks = 0;
  comm_broadcast_kernel<<<grid_for(g), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(*a, b);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

extern "C" int b200dp_comm_reduce_f64(const b200dp_comm_adam* a, const double* in, double* out, int k, int op,
                                      void* stream) {
  if (int rc = validate(a)) return rc;
  B200DP_REQUIRE(in && out && k >= 1 && k <= B200DP_REDUCE_MAX, "reduce_f64: k=%d outside [1, %d]", k, B200DP_REDUCE_MAX);
  B200DP_REQUIRE(op == 0 || op == 1, "reduce_f64: op must be 0 (sum) or 1 (max)");
  if (a->world > 1)
    for (int r = 0; r < a->world; ++r) B200DP_REQUIRE(a->slots[r] != nullptr, "reduce_f64: slots[%d] is null", r);
  comm_reduce_f64_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(*a, in, out, k, op);
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}
