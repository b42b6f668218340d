/* b200dp native library -- C ABI.
 *
 * Everything the Python layer calls goes through these `extern "C"` entry points (ctypes;
 * no torch headers, so the library builds in seconds with plain nvcc for sm_100a).  All
 * pointers are device pointers unless said otherwise; `stream` is a `cudaStream_t` passed
 * as `void*`.  Every function returns 0 on success or a negative B200DP_E* code and records a
 * message retrievable with b200dp_last_error().  None of them synchronises the device unless
 * documented.
 *
 * The reference project has no native code at all; each entry point cites the
 * reference *behaviour* it implements (file:line in /root/reference).
 */
#ifndef B200DP_H_
#define B200DP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200DP_OK 0
#define B200DP_EINVAL (-1)
#define B200DP_ECUDA (-2)
#define B200DP_ENOMEM (-3)
#define B200DP_ESTATE (-4)

#define B200DP_MAX_WORLD 16
#define B200DP_COMM_MAX_BLOCKS 256
#define B200DP_REDUCE_MAX 32
#define B200DP_AUG_NPARAM 12

const char* b200dp_last_error(void);
int b200dp_version(void);
/* Number of kernels this library has launched in this process (bench.py's "gpu_launches"). */
unsigned long long b200dp_launch_count(void);
void b200dp_reset_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Device memory arenas shared between ranks of one node.
 * A rank allocates an arena, exports a 64-byte IPC handle, and opens its peers' handles; the
 * kernels below then load/store peer memory directly over NVLink / NVSwitch.
 * (Replaces the NCCL communicator DDP builds at cifar10_mpi_mobilenet_224.py:142-145.)
 * ------------------------------------------------------------------------------------------- */
int b200dp_arena_alloc(size_t bytes, void** out_ptr);           /* cudaMalloc + zero fill (synchronises) */
int b200dp_arena_free(void* ptr);
int b200dp_ipc_export(void* ptr, unsigned char handle_out[64]);
int b200dp_ipc_open(const unsigned char handle[64], void** out_ptr);
int b200dp_ipc_close(void* ptr);
int b200dp_enable_peer_access(int peer_device);                 /* same-process multi-GPU (tests) */

/* ---------------------------------------------------------------------------------------------
 * Fused gradient all-reduce + 1/W average + Adam (K-C5 + K12 + K13).
 *   reference: loss.backward() -> DDP bucketed ncclAllReduce, then optimizer.step(), then
 *   optimizer.zero_grad()  (cifar10_mpi_mobilenet_224.py:176-180)
 *
 * One kernel per step.  Every rank's gradient vector lives in a peer-mapped arena; after a
 * flag barrier each rank reads all W copies in rank order 0..W-1 (bit-identical sums on every
 * rank), scales, runs Adam on its fp32 master weights / moments, optionally refreshes a bf16
 * working copy, then (after a second flag barrier) zeroes its own gradient vector.
 * ------------------------------------------------------------------------------------------- */
typedef struct b200dp_comm_adam {
  int world;                                /* number of ranks (1..B200DP_MAX_WORLD) */
  int rank;
  long long numel;                          /* elements; must be a multiple of 4 */
  int grad_dtype;                           /* 0 = fp32 gradients, 1 = bf16 gradients */
  int algo;                                 /* 0 = one-shot (every rank reduces everything),
                                               1 = two-shot (rank r reduces + updates slice r, then
                                                   stores the new weights into every peer),
                                               2 = NVLS: two-shot with multimem.ld_reduce / multimem.st on
                                                   the multicast mapping (mc_grads, mc_params) */
  const void* grads[B200DP_MAX_WORLD];      /* grads[r] = rank r's gradient vector as mapped HERE */
  uint32_t* flags[B200DP_MAX_WORLD];        /* flags[r] = rank r's flag pad as mapped HERE
                                               (B200DP_COMM_MAX_BLOCKS * B200DP_MAX_WORLD uint32) */
  float* params[B200DP_MAX_WORLD];          /* two-shot only: every rank's master weights as mapped
                                               HERE; one-shot uses params[rank] only */
  void* work_bf16[B200DP_MAX_WORLD];        /* optional bf16 working copies (same indexing); NULL = none */
  float* m;                                 /* Adam first moment  (local) */
  float* v;                                 /* Adam second moment (local) */
  const float* lr;                          /* device scalar: learning rate for this step */
  int* step_state;                          /* device int[2]: [0] = completed Adam steps, [1] = block ticket (0 between launches) */
  uint32_t* epoch_state;                    /* device uint32[B200DP_COMM_MAX_BLOCKS]: barrier epochs */
  int* error_flag;                          /* device int: set non-zero if a barrier spin timed out */
  double beta1, beta2;                      /* doubles: (1 - beta) and the bias corrections match torch.optim.Adam exactly */
  float eps, weight_decay;
  int zero_grads;                           /* zero this rank's gradient vector at the end */
  int apply_adam;                           /* 0 = plain all-reduce: the averaged gradient is written
                                               to out_avg instead (bus-bandwidth sweep, tests) */
  float* out_avg;                           /* apply_adam == 0: fp32 output vector (local) */
  unsigned long long spin_limit;            /* polls before a barrier gives up (0 = default) */
  int max_blocks;                           /* 0 = default grid */
  /* ---- bucketed launches (engine.cu launches one per gradient bucket from a branch of the step graph) ---- */
  long long range_lo, range_hi;             /* element sub-range handled by this launch, multiples of 4;
                                               range_hi == 0 means the whole vector */
  int bump_step;                            /* advance the Adam step count when this launch is done (the LAST launch
                                               of a step sets it) */
  int skip_barriers;                        /* the caller has already waited for the peers with a one-CTA
                                               b200dp_comm_barrier ahead of this launch (in-graph buckets): do not
                                               wait again (zero_grads must then be 0 unless world == 1) */
  int exact_blocks;                         /* > 0: use exactly this grid */
  long long zero_lo, zero_hi;               /* extra range of this rank's fp32 gradients to clear at the end (after the second
                                               barrier): the ranges of the step's earlier skip_barriers launches */
  /* ---- NVLS ---- */
  const void* mc_grads;                     /* multicast address of the gradient vector (algo 2) */
  float* mc_params;                         /* multicast address of the master weights (algo 2) */
  /* ---- failure reporting / small reductions ---- */
  int* error_host;                          /* optional host-mapped int mirroring error_flag: the trainer polls it
                                               without synchronising the device */
  double* slots[B200DP_MAX_WORLD];          /* slots[r] = rank r's table double[B200DP_MAX_WORLD][B200DP_REDUCE_MAX]
                                               as mapped HERE (b200dp_comm_reduce_f64) */
  uint32_t* abort_words[B200DP_MAX_WORLD];  /* optional: abort_words[r] = one word in rank r's arena as mapped HERE.  A
                                               rank whose barrier times out sets it on EVERY rank; a rank that finds its
                                               own word set after a barrier returns without updating anything, so a
                                               late peer cannot store weights into a rank that has already given up */
} b200dp_comm_adam;

/* Launch the fused step on `stream`.  All ranks must call it the same number of times with the
 * same numel/algo.  With world == 1 no flags are touched. */
int b200dp_comm_adam_step(const b200dp_comm_adam* args, void* stream);
/* Grid the call above will use for these arguments (every rank must agree; exposed for tests). */
int b200dp_comm_adam_grid(const b200dp_comm_adam* args);

/* Device-side barrier across ranks with the same flag protocol. */
int b200dp_comm_barrier(const b200dp_comm_adam* args, void* stream);

/* Rank `root` stores `numel` fp32 values into every peer's `dst` (K-C3, K-C4:
 * DDP's parameter broadcast at construction and buffer broadcast, ref :142, :177). */
int b200dp_comm_broadcast(const b200dp_comm_adam* args, int root, float* const* dst_by_rank, float* mc_dst,
                          const float* src, long long numel, void* stream);
/* `mc_dst` (may be NULL): multicast alias of the destination; the root then issues one multimem.st per 16 bytes
 * and the NVSwitch replicates it, instead of W-1 peer stores. */

/* All-reduce (op 0 = sum, 1 = max) of k <= B200DP_REDUCE_MAX doubles over the peer slot tables: the epoch-level
 * metric sums and the max-over-ranks timers (K-C6, K-C7: dist.all_reduce of [loss_sum, count],
 * ref :188-195, :216-223).  `in` / `out` are local device arrays; identical results on every rank. */
int b200dp_comm_reduce_f64(const b200dp_comm_adam* args, const double* in, double* out, int k, int op,
                           void* stream);

/* ---------------------------------------------------------------------------------------------
 * Symmetric memory with an NVSwitch multicast alias (csrc/vmm.cu; SURVEY §5.8 "memory model").
 * CUDA VMM allocations exported as POSIX file descriptors (the Python layer passes them between
 * ranks over AF_UNIX sockets) and one multicast object bound over every rank's allocation.
 * (Replaces what NCCL does internally for the communicator of ref :142-145.)
 * ------------------------------------------------------------------------------------------- */
int b200dp_vmm_query(int* flags_out, size_t* granularity_out);     /* bit 0: VMM + fd export, bit 1: multicast */
int b200dp_vmm_alloc(size_t bytes, void** ptr_out, int* fd_out, unsigned long long* handle_out);
int b200dp_vmm_import(int fd, size_t bytes, void** ptr_out);
int b200dp_vmm_free(void* ptr);
int b200dp_mc_create(int world, size_t bytes, unsigned long long* mc_out, int* fd_out);    /* rank 0 */
int b200dp_mc_import(int fd, size_t bytes, unsigned long long* mc_out);                    /* other ranks */
int b200dp_mc_add_device(unsigned long long mc);        /* every rank, before any bind */
int b200dp_mc_bind(unsigned long long mc, unsigned long long mem_handle, size_t bytes);
int b200dp_mc_map(unsigned long long mc, size_t bytes, void** ptr_out);
int b200dp_mc_release(unsigned long long mc);
int b200dp_close_fd(int fd);

/* ---------------------------------------------------------------------------------------------
 * Input pipeline (P1-P7, K14).
 *   reference: transforms.Compose([...]) on PIL images in DataLoader workers, then a 77 MB fp32
 *   H2D copy per step (cifar10_mpi_mobilenet_224.py:72-89, :174)
 * uint8 [B,32,32,3] -> normalised bf16 NHWC [B,size,size,out_channels] (out_channels 3 or 4; the
 * 4th channel is zero).  params == NULL selects the test transform; otherwise one row of
 * B200DP_AUG_NPARAM floats per sample (data/transforms.py documents the columns).
 * ------------------------------------------------------------------------------------------- */
int b200dp_preprocess(const uint8_t* images, const float* params, void* out_bf16, int batch,
                      int size, int out_channels, void* stream);

/* The same pipeline fused with the stem convolution's im2col: col = bf16 [batch * (size/2)^2][32] rows of the 27 taps
 * (ky, kx, channel) of the 3x3 stride-2 stem + 5 zeros, never materialising the image; asum (may be NULL) fp32 [32] +=
 * column sums of the matrix.  b200dp_preprocess_col_supported(size) says whether the working set fits (size <= 224). */
int b200dp_preprocess_col(const uint8_t* images, const float* params, void* col_bf16, float* asum, int batch, int size,
                          void* stream);
int b200dp_preprocess_col_supported(int size);
/* Inference on full-resolution photographs: PIL-style anti-aliased bilinear Resize((size, size)) + ToTensor + Normalize of
 * uint8 [batch][H][W][3] -> bf16 [batch][size][size][4] (what cifar10_128batch.py:157-165 and the Gradio app of
 * GROUP03.pdf p.26-27 do on the CPU with torchvision.transforms). */
int b200dp_resize_normalize(const uint8_t* src, int batch, int H, int W, void* out_bf16, int size, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Classifier head (K7-K11, K15).
 *   reference: adaptive_avg_pool2d -> Dropout(0.2) -> Linear(1280,10) -> CrossEntropyLoss ->
 *   argmax / correct count / loss.item()  (torchvision mobilenetv2.py:166-171;
 *   cifar10_mpi_mobilenet_224.py:178-184)
 * ------------------------------------------------------------------------------------------- */
/* logits fp32 [B,C], labels int64 [B] -> per-sample loss (optional), dlogits (optional, already
 * divided by `mean_over`), and metrics[3] += {sum of losses, correct count, B} as doubles. */
int b200dp_softmax_xent(const float* logits, const int64_t* labels, float* loss_out,
                        float* dlogits_out, double* metrics, int batch, int classes,
                        float mean_over, void* stream);
/* probs = softmax(logits) and top-k (values, indices) per row -- predict_cifar10_image,
 * cifar10_128batch.py:170-172. */
int b200dp_softmax_topk(const float* logits, float* probs_out, float* topv_out, int32_t* topi_out,
                        int batch, int classes, int k, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Convolution building blocks of the native engine (stand-alone entry points; the executor in
 * csrc/engine.cu drives the same kernels).  All activations are NHWC bf16, channels % 16 == 0.
 *   reference: torchvision MobileNetV2 ConvBNReLU / InvertedResidual forward + autograd backward,
 *   i.e. cuDNN conv + cuDNN batch-norm + ATen hardtanh (cifar10_mpi_mobilenet_224.py:177, :179)
 * ------------------------------------------------------------------------------------------- */
/* Pointwise conv as a tcgen05 GEMM, out[M,N] = Aeff[M,K] * w[N,K]^T (csrc/pw_gemm.cu).
 * amode 0: a_tma is the bf16 A matrix; 1: Aeff = relu6?(acoef[0]*a0 + acoef[1]); 2: Aeff = acoef[0]*a0 + acoef[1]*a1 + acoef[2].
 * emode 0: store; 1: store + stats[0] += sum(out), stats[1] += sum(out^2); 2: out masked by ReLU6'(ecoef[0]*aux+ecoef[1]),
 * stats[0] += sum(out), stats[1] += sum(out*aux).  `stats` is [2][stats_stride] fp32 and is ACCUMULATED into. */
int b200dp_pw_gemm(int amode, int emode, int M, int N, int K, const void* a_tma, int lda, const void* a0,
                   const void* a1, const float* acoef, int a_relu6, const void* w, void* out, int ldo,
                   const void* aux, const float* ecoef, float* stats, int stats_stride, int max_ctas,
                   void* stream);
/* Pointwise conv weight gradient dw[N][ldw] += G^T A on tcgen05 with MN-major operands (csrc/pw_wgrad.cu).
 * mode 0: `pure` = A [M,K]; G = coef[0]*t0 + coef[1]*t1 + coef[2]; `asum` = [K] column sums of A over all rows (the
 * affine map is factored out of the reduction).  mode 1: `pure` = G [M,N]; A = relu6(coef[0]*t0 + coef[1]); asum unused. */
int b200dp_pw_wgrad(int mode, int M, int N, int K, const void* pure, int ld_pure, const void* t0, const void* t1,
                    int ld_t, const float* coef, const float* asum, float* dw, int ldw, int max_ctas, void* stream);
/* Same over dense tensors with row packing: `pack` consecutive pixels are one row of pack*C channels, only the diagonal
 * blocks of the product are accumulated (pack = 0: chosen by the library; pays when the wide side has <= 64 channels). */
int b200dp_pw_wgrad_packed(int mode, int M, int N, int K, const void* pure, const void* t0, const void* t1,
                           const float* coef, const float* asum, float* dw, int max_ctas, int pack, void* stream);
/* Pointwise GEMM with row packing: `w_blockdiag` = blockdiag_pack(W) [pack*N][pack*K]; dense tensors, true sizes. */
int b200dp_pw_gemm_packed(int amode, int emode, int M, int N, int K, const void* a_tma, const void* a0, const void* a1,
                          const float* acoef, int a_relu6, const void* w_blockdiag, void* out, const void* aux,
                          const float* ecoef, float* stats, int stats_stride, int max_ctas, int pack, void* stream);
int b200dp_pw_gemm_pick_pack(int M, int N, int K);
/* Depthwise 3x3 (pad 1, stride 1|2): y = dwconv(relu6(scale*x + shift)), w fp32 [9][C];
 * stats (may be NULL) [2][stats_stride] += sum(y), sum(y^2): a scratch accumulator that must be zero on entry. */
int b200dp_dw_fwd(const void* x, const float* scale, const float* shift, const float* w, void* y, float* stats,
                  int stats_stride, int B, int H, int W, int C, int stride, void* stream);
/* Depthwise backward: data gradient (masked by ReLU6' of the input activation), weight gradient and the sums the
 * producer's BatchNorm backward needs, in one pass (csrc/dwconv.cu). */
int b200dp_dw_bwd(const void* e_raw, const float* escale, const float* eshift, const void* dzd, const void* d_raw,
                  const float* dcoef, const float* w, void* dze, float* stats, int stats_stride, float* dw, int B,
                  int H, int W, int C, int stride, void* stream);
/* Stem 3x3 stride-2 conv 3->32 on a 4-channel bf16 image, and its weight gradient (csrc/stem.cu). */
int b200dp_stem_fwd(const void* x, const float* w, void* y, float* stats, int stats_stride, int B, int H, int W,
                    void* stream);
int b200dp_stem_wgrad(const void* x, const void* dz, const void* yraw, const float* coef, float* dw, int B, int H,
                      int W, void* stream);
/* BatchNorm finalisation: sums -> (scale, shift), (mean, invstd), running statistics (rmean/rvar may be NULL);
 * and sums of (dz, dz*x) -> (ca, cb, cc) with dx = ca*dz + cb*x + cc, plus dgamma / dbeta (accumulated). */
int b200dp_bn_fwd_finalize(const float* sums, int stride, const float* gamma, const float* beta, float count,
                           float eps, float momentum, float* rmean, float* rvar, float* coef, float* saved, int C,
                           void* stream);
int b200dp_bn_bwd_finalize(const float* sums, int stride, const float* saved, const float* gamma, float count,
                           float* coef3, float* dgamma, float* dbeta, int C, void* stream);
/* Passes over the narrow block outputs: out = c0*a + c1 (+ b), or (three) c0*a + c1*b + c2; and the gradient
 * join dout = g1 (+ g2) with the two per-channel sums of the projection BatchNorm backward. */
int b200dp_affine(const void* a, const void* b, const float* coef, int three, void* out, long long rows, int C,
                  void* stream);
int b200dp_grad_join_stats(const void* g1, const void* g2, const void* x, void* dout, float* sums, int stride,
                           long long rows, int C, void* stream);
/* Block output of an inverted residual: out = coef[0]*raw + coef[1] (+ res); colsum (may be NULL) [C] += column sums of
 * the rounded result (the next expand conv's factored weight gradient needs them). */
int b200dp_block_out(const void* raw, const void* res, const float* coef, void* out, float* colsum, long long rows,
                     int C, void* stream);

/* Depthwise conv + the BatchNorm coefficients of its output, with NO device allocation by anyone.
 * This entry point has no counterpart in the reference.
 * It is b200dp_dw_fwd followed by b200dp_bn_fwd_finalize, except that the per-channel sums -- scratch that
 * the two-call form makes the caller allocate and zero -- live in a static 8 KB __device__ array inside this
 * library (one copy on every GPU that loads it, kept until exit).  It exists for callers that may not
 * allocate device memory at the point of the call: plug-ins inside another runtime's enqueue hook, serving
 * loops with a fixed up-front memory budget, latency-critical paths that must never enter the allocator, or
 * processes whose free memory is all held by a framework allocator.
 *   - C must be <= B200DP_DW_NOALLOC_MAX_C (else B200DP_EINVAL; use the two-call form);
 *   - the call clears the part of the scratch it uses, runs the conv, writes coef_out [2][C] = (scale, shift)
 *     and saved_out [2][C] = (mean, inv-std); running statistics are not touched;
 *   - all calls on one GPU share the scratch: calls that may overlap on the device race -- serialise them;
 *   - the FIRST call on each device looks up the scratch's device address (cudaGetSymbolAddress); make that
 *     first call before any stream capture begins.  The look-up is not synchronised between host threads: two
 *     threads making the first call at once both store the same address, which is harmless but is a data race in
 *     the letter of the C++ memory model -- make the first call from one thread. */
#define B200DP_DW_NOALLOC_MAX_C 1024
int b200dp_dw_fwd_bn_noalloc(const void* x, const float* in_scale, const float* in_shift, const float* w,
                             const float* gamma, const float* beta, float eps, void* y, float* coef_out,
                             float* saved_out, int B, int H, int W, int C, int stride, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Whole-model executor (csrc/engine.cu): MobileNetV2 + 10-class head, one training or evaluation step as one
 * CUDA-graph launch over the kernels above.
 *   reference: the hot loop zero_grad -> forward -> loss -> backward -> step of
 *   cifar10_mpi_mobilenet_224.py:173-185 and the eval loop :199-213.
 * `offsets` is the table of models/arena.py (per conv: weight, gamma, beta, running mean, running var element
 * offsets; then classifier weight, bias); params / grads / buffers are flat fp32 arenas owned by the caller.
 * ------------------------------------------------------------------------------------------- */
int b200dp_engine_create(void** out, int batch, int img_size, int num_classes, const long long* offsets, int n_offsets,
                         long long param_total, long long buffer_total, float* params, float* grads, float* buffers,
                         float dropout, float bn_eps, float bn_momentum, unsigned seed, int use_graphs);
int b200dp_engine_destroy(void* engine);
size_t b200dp_engine_workspace_bytes(void* engine);
void* b200dp_engine_buffer(void* engine, int which);
void* b200dp_engine_tap(void* engine, int what, int index);
/* forward + loss + backward (gradients accumulate into `grads`); with b200dp_engine_set_comm also the bucketed
 * gradient all-reduce + Adam, overlapped with backward (DDP's Reducer + optimizer.step, ref :142-145, :179-180). */
int b200dp_engine_train_step(void* engine, const uint8_t* images, const int64_t* labels, const float* aug, int n,
                             void* stream);
int b200dp_engine_eval_step(void* engine, const uint8_t* images, const int64_t* labels, int n, void* stream);
/* Forward of already pre-processed input: `x` is normalised bf16 NHWC [n, S, S, 4] (inference on full-resolution
 * images that were resized outside, ref cifar10_128batch.py:157-172); logits land in buffer 3. */
int b200dp_engine_forward_preprocessed(void* engine, const void* x, int n, void* stream);
int b200dp_engine_invalidate(void* engine);
int b200dp_engine_set_comm(void* engine, const b200dp_comm_adam* args, const int* bucket_first, int n_buckets,
                           int comm_blocks);
int b200dp_engine_profile_step(void* engine, const uint8_t* images, const int64_t* labels, const float* aug, int n,
                               char* names, int names_cap, float* ms, int ms_cap, int* count, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
