// Native executor: builds the MobileNetV2 layer plan, owns the activation / statistics workspace, and runs a
// whole training or evaluation step as one sequence of hand-written kernels -- captured into a CUDA graph
// per batch size, so a step costs one graph launch on the host.
//
// Reference behaviour replaced: the Python hot loop zero_grad -> forward -> loss -> backward of
// cifar10_mpi_mobilenet_224.py:173-184 with ~700 ATen/cuDNN kernel launches and two host syncs per step.
// The optimiser / gradient all-reduce is NOT in here: gradients are left in the caller's flat fp32 arena
// (which lives in peer-mapped memory) and the fused all-reduce + Adam kernel (comm_adam.cu) consumes them.
//
// Data flow per inverted-residual block (training):
//   x (finished, bf16) --expand GEMM [TMA A, stats epilogue]--> e_raw --dw conv [BN+ReLU6 on load, stats]--> d_raw
//     --project GEMM [BN+ReLU6 prologue, stats epilogue]--> p_raw --affine (+x)--> next x
// and the mirror image backwards; every expanded tensor is written once and read once per direction.
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "common.cuh"
#include "kernels.h"

extern "C" int b200dp_preprocess(const uint8_t* images, const float* params, void* out_bf16, int batch, int size,
                                 int out_channels, void* stream);
namespace b200dp {
int comm_adam_launch(const b200dp_comm_adam& a, cudaStream_t s);
int comm_barrier_launch(const b200dp_comm_adam& a, int blocks, cudaStream_t s);
}
extern "C" int b200dp_preprocess_col(const uint8_t* images, const float* params, void* col_bf16, float* asum, int batch,
                                     int size, void* stream);
extern "C" int b200dp_preprocess_col_supported(int size);
extern "C" int b200dp_softmax_xent(const float* logits, const int64_t* labels, float* loss_out, float* dlogits_out,
                                   double* metrics, int batch, int classes, float mean_over, void* stream);

namespace b200dp {
namespace {

inline int pad16(int c) { return (c + 15) / 16 * 16; }

struct Conv {
  int kind;            // 0 stem, 1 dw, 2 pw
  int role;            // 0 stem, 1 expand, 2 dw, 3 project, 4 last
  int cin, cout, cinp, coutp, stride, hin, hout, block;
  long long off_w, off_g, off_b, off_rm, off_rv;    // element offsets: params arena (w, gamma, beta), buffer arena
  int pack_f = 1, pack_d = 1;                        // row packing of the forward / data-gradient GEMM (pw_gemm.h)
  long long off_pf = 0, off_pd = 0;                  // block-diagonal weight copies inside w_pack
  // workspace
  __nv_bfloat16* raw = nullptr;                      // raw conv output [B, hout, hout, coutp]
  float *sums_f = nullptr, *sums_b = nullptr, *coef_f = nullptr, *saved = nullptr, *coef_b = nullptr;
};

struct Block {
  int index, cin, cout, cinp, coutp, stride, expand, hin, hout;
  bool use_res;
  int ce, cd, cp;      // conv indices (ce = -1 when expand == 1)
  __nv_bfloat16* xout = nullptr;   // finished block output [B, hout, hout, coutp]
  float* xsum = nullptr;           // per-channel column sums of xout (for the consumer's factored weight gradient)
  float* mask1 = nullptr;          // [2][coutp]: scale 0, shift 3 -> ReLU6'(0 * aux + 3) = 1 for every element
};

struct WPrepEntry { long long off; int n, k; int rf, rd; long long off_pf, off_pd; };

struct Plans {            // per batch size
  int n = 0;
  std::vector<PwGemmPlan> fwd, fwd_eval, dgrad;    // indexed by conv (index 0 = the stem run as an im2col GEMM)
  cudaGraphExec_t train_graph = nullptr, eval_graph = nullptr, eval_pre_graph = nullptr;
  int train_aug = -1;      // augmentation flag the train graph was captured with
  unsigned long long train_kernels = 0, eval_kernels = 0;   // kernel nodes in each graph (for the launch counter)
};

}  // namespace

struct Engine {
  int B, S, classes;
  float p_drop, eps, momentum;
  uint32_t seed;
  std::vector<Conv> convs;
  std::vector<Block> blocks;
  long long off_fcw, off_fcb, param_total, buffer_total;
  float *params, *grads, *buffers;
  // workspace
  char* ws = nullptr;
  size_t ws_bytes = 0;
  uint8_t* in_images = nullptr;
  int64_t* in_labels = nullptr;
  float* in_aug = nullptr;
  __nv_bfloat16 *x0 = nullptr, *w_bf16 = nullptr, *wt_bf16 = nullptr, *w_pack = nullptr;
  long long pack_total = 0;
  // stem as a GEMM (B200DP_STEM=simt: the CUDA-core stem kernels): im2col matrix, bf16 weights (plain + block
  // diagonal), column sums of the im2col matrix, [32][32] gradient scratch
  bool stem_gemm = true;
  bool fuse_col = true;     // preprocess writes the im2col matrix directly (B200DP_FUSE_COL=0: image, then an im2col pass)
  int stem_pack = 1;
  __nv_bfloat16 *x0col = nullptr, *stem_w = nullptr;
  float *stem_asum = nullptr, *stem_dw = nullptr;
  // dze / dp are double-buffered: the weight-gradient kernels that read them run on side branches of the step graph
  // and may still be busy when the main chain produces the next block's tensors
  __nv_bfloat16 *dzd = nullptr, *dze[2] = {nullptr, nullptr}, *dp[2] = {nullptr, nullptr}, *gx[2] = {nullptr, nullptr},
                *dout[2] = {nullptr, nullptr};
  float *sums_all = nullptr;
  size_t sums_bytes = 0;
  float *pooled = nullptr, *drop = nullptr, *logits = nullptr, *dlogits = nullptr, *dpool = nullptr;
  double *metrics_train = nullptr, *metrics_eval = nullptr;
  int* step_dev = nullptr;
  // B200DP_JOIN_IN_DGRAD=1: where a block's output gradient is the consumer's data gradient alone (no residual branch
  // joins it), the two sums of its projection BatchNorm backward come out of that data-gradient GEMM's epilogue
  // (E_MASKSTATS with an always-open mask: Block::mask1) and the grad_join_stats launch disappears
  bool join_in_dgrad = false;
  WPrepEntry* wprep_dev = nullptr;
  int n_pw = 0;
  bool use_graphs = true;
  bool bn_fuse = true;      // BatchNorm finalisation in the first consumer's prologue (B200DP_BN_FUSE=0: own kernels)
  bool pw_pack = true;      // row packing of narrow pointwise layers (B200DP_PW_PACK=0: off)
  bool dw_tiled = true;     // stride-1 depthwise convs on the TMA-tiled kernels (B200DP_DW_IMPL=simt: sliding-window kernels)
  cudaStream_t cap_stream = nullptr;   // graphs are captured here (the caller's stream may be the legacy default stream)
  // fork / join inside a step (B200DP_FORK=0: one linear chain).  side[0]: project-conv weight gradients,
  // side[1]: expand-conv / last-conv / stem weight gradients, side[2]: gradient all-reduce + Adam per bucket
  // Measured (1x B200, batch 128): forking the weight gradients is SLOWER at every threshold -- 5.09 ms/step with none,
  // 5.10 / 5.14 / 5.18 with the blocks up to 7 / 14 / 28 rows tall forked, 5.19 with all.  Every kernel here is sized to
  // fill the machine on its own, so a concurrent branch only takes SMs, TMEM columns and L2 away from the critical
  // path.  The weight-gradient fork is therefore opt-in (B200DP_FORK_WGRAD=1, B200DP_FORK_MAXH=<rows>); the forward
  // weight-preparation fork and the communication branch, which do not compete with anything, are always on.
  bool fork = true;
  bool fork_wgrad = false;
  int fork_max_h = 14;       // only blocks whose input is at most this tall fork their weight gradients
  int fork_sms = 40;         // SMs given to the weight-gradient branches of a forked block; the main chain sizes for the rest
  cudaStream_t side[3] = {nullptr, nullptr, nullptr};
  std::vector<cudaEvent_t> ev_pool;
  size_t ev_next = 0;
  // Gradient all-reduce + Adam inside the step (set_comm): one launch per bucket, on side[2], as soon as backward
  // has produced the bucket.  Buckets are address ranges of the flat arena; bucket_first[k] = index of the first
  // block of bucket k (descending: backward reaches them in this order), the last bucket ends with the stem.
  bool comm_on = false;
  b200dp_comm_adam comm{};
  int comm_blocks = 148;
  std::vector<int> bucket_first;
  bool eval_coef_valid = false;
  std::map<int, Plans> plans;
  // profiling
  bool profiling = false;
  std::vector<std::pair<std::string, cudaEvent_t>> prof_events;
};

namespace {

#define ENG_TRY(call)            \
  do {                           \
    int rc__ = (call);           \
    if (rc__ != B200DP_OK) return rc__; \
  } while (0)

void build_spec(Engine& e, const long long* off) {
  static const int setting[7][4] = {{1, 16, 1, 1}, {6, 24, 2, 2}, {6, 32, 3, 2}, {6, 64, 4, 2},
                                    {6, 96, 3, 1}, {6, 160, 3, 2}, {6, 320, 1, 1}};
  auto add = [&](int kind, int role, int cin, int cout, int stride, int hin, int block) {
    Conv c{};
    c.kind = kind; c.role = role; c.cin = cin; c.cout = cout; c.cinp = pad16(cin); c.coutp = pad16(cout);
    c.stride = stride; c.hin = hin; c.hout = (hin - 1) / stride + 1; c.block = block;
    const int i = (int)e.convs.size();
    c.off_w = off[i * 5 + 0]; c.off_g = off[i * 5 + 1]; c.off_b = off[i * 5 + 2];
    c.off_rm = off[i * 5 + 3]; c.off_rv = off[i * 5 + 4];
    e.convs.push_back(c);
    return i;
  };
  int h = e.S;
  add(0, 0, 3, 32, 2, h, 0);
  h = e.convs[0].hout;
  int cin = 32, feat = 1;
  for (auto& st : setting) {
    for (int i = 0; i < st[2]; ++i) {
      const int stride = i == 0 ? st[3] : 1;
      const int hidden = cin * st[0];
      Block b{};
      b.index = feat; b.cin = cin; b.cout = st[1]; b.cinp = pad16(cin); b.coutp = pad16(st[1]);
      b.stride = stride; b.expand = st[0]; b.hin = h; b.hout = (h - 1) / stride + 1;
      b.use_res = stride == 1 && cin == st[1];
      b.ce = st[0] != 1 ? add(2, 1, cin, hidden, 1, h, feat) : -1;
      b.cd = add(1, 2, hidden, hidden, stride, h, feat);
      b.cp = add(2, 3, hidden, st[1], 1, b.hout, feat);
      e.blocks.push_back(b);
      cin = st[1];
      h = b.hout;
      ++feat;
    }
  }
  add(2, 4, cin, 1280, 1, h, feat);
  const int nc = (int)e.convs.size();
  e.off_fcw = off[nc * 5];
  e.off_fcb = off[nc * 5 + 1];
}

template <typename T>
T* carve(char*& p, size_t count) {
  T* r = reinterpret_cast<T*>(p);
  p += (count * sizeof(T) + 255) / 256 * 256;
  return r;
}

size_t layout_workspace(Engine& e, char* base) {
  char* p = base;
  const size_t B = e.B;
  e.in_images = carve<uint8_t>(p, B * 32 * 32 * 3);
  e.in_labels = carve<int64_t>(p, B);
  e.in_aug = carve<float>(p, B * B200DP_AUG_NPARAM);
  e.x0 = carve<__nv_bfloat16>(p, B * e.S * e.S * 4);
  e.x0col = carve<__nv_bfloat16>(p, e.stem_gemm ? B * (e.S / 2) * (e.S / 2) * 32 : 8);
  e.stem_w = carve<__nv_bfloat16>(p, 1024 + 32 * 32 * 16);
  e.w_bf16 = carve<__nv_bfloat16>(p, e.param_total);
  e.wt_bf16 = carve<__nv_bfloat16>(p, e.param_total);
  e.w_pack = carve<__nv_bfloat16>(p, e.pack_total > 0 ? e.pack_total : 8);
  size_t max_d = 0, max_e = 0, max_x = 0;
  // statistics region first (one memset per step)
  char* sums0 = p;
  for (auto& c : e.convs) {
    c.sums_f = carve<float>(p, 2 * c.coutp);
    c.sums_b = carve<float>(p, 2 * c.coutp);
  }
  for (auto& b : e.blocks) b.xsum = carve<float>(p, b.coutp);
  e.stem_asum = carve<float>(p, 32);
  e.stem_dw = carve<float>(p, 32 * 32);
  e.sums_all = reinterpret_cast<float*>(sums0);
  e.sums_bytes = (size_t)(p - sums0);
  for (auto& c : e.convs) {
    c.coef_f = carve<float>(p, 2 * c.coutp);
    c.saved = carve<float>(p, 2 * c.coutp);
    c.coef_b = carve<float>(p, 3 * c.coutp);
    const size_t elems = B * c.hout * c.hout * c.coutp;
    c.raw = carve<__nv_bfloat16>(p, elems);
    if (c.role == 2) max_d = elems > max_d ? elems : max_d;
    if (c.role == 1 || c.role == 0 || c.role == 4) max_e = elems > max_e ? elems : max_e;
  }
  for (auto& b : e.blocks) {
    const size_t elems = B * b.hout * b.hout * b.coutp;
    b.xout = carve<__nv_bfloat16>(p, elems);
    max_x = elems > max_x ? elems : max_x;
    const size_t ein = B * b.hin * b.hin * b.cinp;
    max_x = ein > max_x ? ein : max_x;
  }
  e.dzd = carve<__nv_bfloat16>(p, max_d);
  for (int i = 0; i < 2; ++i) {
    e.dze[i] = carve<__nv_bfloat16>(p, max_e);
    e.dp[i] = carve<__nv_bfloat16>(p, max_x);
    e.gx[i] = carve<__nv_bfloat16>(p, max_x);
    e.dout[i] = carve<__nv_bfloat16>(p, max_x);
  }
  e.pooled = carve<float>(p, B * 1280);
  e.drop = carve<float>(p, B * 1280);
  e.dpool = carve<float>(p, B * 1280);
  e.logits = carve<float>(p, B * e.classes);
  e.dlogits = carve<float>(p, B * e.classes);
  e.metrics_train = carve<double>(p, 4);
  e.metrics_eval = carve<double>(p, 4);
  e.step_dev = carve<int>(p, 4);
  for (auto& b : e.blocks) b.mask1 = carve<float>(p, 2 * b.coutp);
  e.wprep_dev = carve<WPrepEntry>(p, 64);
  return (size_t)(p - base);
}

// profiling hook: call after each launch
void mark(Engine& e, cudaStream_t s, const char* name, int conv = -1) {
  if (!e.profiling) return;
  cudaEvent_t ev;
  cudaEventCreate(&ev);
  cudaEventRecord(ev, s);
  std::string label = name;
  if (conv >= 0) label += "#" + std::to_string(conv);
  e.prof_events.emplace_back(label, ev);
}

// Descriptors for "finalise this BatchNorm in your prologue" (bn_inline.cuh), handed to the first consumer of the
// coefficients instead of launching a finalisation kernel.
BnFwdSrc src_f(const Engine& e, const Conv& c, float count) {
  BnFwdSrc b;
  b.sums = c.sums_f; b.stride = c.coutp; b.C = c.coutp;
  b.count = count; b.eps = e.eps; b.momentum = e.momentum;
  b.gamma = e.params + c.off_g; b.beta = e.params + c.off_b;
  b.rmean = e.buffers + c.off_rm; b.rvar = e.buffers + c.off_rv;
  b.coef = c.coef_f; b.saved = c.saved;
  return b;
}
BnBwdSrc src_b(const Engine& e, const Conv& c, float count) {
  BnBwdSrc b;
  b.sums = c.sums_b; b.stride = c.coutp; b.C = c.coutp; b.count = count;
  b.saved = c.saved; b.gamma = e.params + c.off_g;
  b.coef = c.coef_b; b.dgamma = e.grads + c.off_g; b.dbeta = e.grads + c.off_b;
  return b;
}

// Which copy of dze / dp a block uses: consecutive blocks alternate, the head (index = number of blocks) writes dze[0].
inline int dze_idx(const Engine& e, int bi) { return ((int)e.blocks.size() - bi) & 1; }
inline int dp_idx(int bi) { return bi & 1; }

int build_plans(Engine& e, Plans& pl, int n) {
  pl.n = n;
  const int nc = (int)e.convs.size();
  pl.fwd.assign(nc, PwGemmPlan{});
  pl.fwd_eval.assign(nc, PwGemmPlan{});
  pl.dgrad.assign(nc, PwGemmPlan{});
  // forward / data-gradient weights of a conv: the plain bf16 copy, or its block-diagonal copy when the layer is packed
  auto wf = [&](const Conv& c) { return c.pack_f > 1 ? e.w_pack + c.off_pf : e.w_bf16 + c.off_w; };
  auto wd = [&](const Conv& c) { return c.pack_d > 1 ? e.w_pack + c.off_pd : e.wt_bf16 + c.off_w; };
  // data-gradient GEMMs of blocks whose weight gradients run on side branches leave fork_sms SMs to those branches
  auto main_ctas = [&](int h) { return (e.fork && e.fork_wgrad && h <= e.fork_max_h) ? num_sms() - e.fork_sms : 0; };
  if (e.stem_gemm) {
    const Conv& st = e.convs[0];
    const int M = n * st.hout * st.hout;
    const int r = M % e.stem_pack == 0 ? e.stem_pack : 1;
    for (int ev = 0; ev < 2; ++ev)
      ENG_TRY(pw_gemm_plan(ev ? &pl.fwd_eval[0] : &pl.fwd[0], A_TMA, ev ? E_PLAIN : E_STATS, M, 32, 32, e.x0col, 32,
                           nullptr, nullptr, nullptr, 0, r > 1 ? e.stem_w + 1024 : e.stem_w, st.raw, 32, nullptr,
                           nullptr, st.sums_f, st.coutp, 0, r));
  }
  for (auto& b : e.blocks) {
    const Conv& cp = e.convs[b.cp];
    const Conv& cd = e.convs[b.cd];
    const int Mout = n * b.hout * b.hout;
    const int Min = n * b.hin * b.hin;
    const __nv_bfloat16* xin = b.index == 1 ? nullptr : e.blocks[b.index - 2].xout;
    if (b.ce >= 0) {
      const Conv& ce = e.convs[b.ce];
      for (int ev = 0; ev < 2; ++ev)
        ENG_TRY(pw_gemm_plan(ev ? &pl.fwd_eval[b.ce] : &pl.fwd[b.ce], A_TMA, ev ? E_PLAIN : E_STATS, Min, ce.coutp,
                             ce.cinp, xin, ce.cinp, nullptr, nullptr, nullptr, 0, wf(ce), ce.raw, ce.coutp,
                             nullptr, nullptr, ce.sums_f, ce.coutp, 0, ce.pack_f));
      // dgrad: dx[M, cin] = BNgrad(dze, e_raw)[M, hidden] * W[hidden, cin]   (weights transposed: [cin][hidden])
      // the gradient this GEMM produces belongs to the previous block's output; without a residual branch it IS that
      // gradient, and the sums of that block's projection BatchNorm backward can be taken here
      const bool stats_here = e.join_in_dgrad && !b.use_res && b.index >= 2;
      const Conv* cpp = b.index >= 2 ? &e.convs[e.blocks[b.index - 2].cp] : nullptr;
      ENG_TRY(pw_gemm_plan(&pl.dgrad[b.ce], A_BNGRAD, stats_here ? E_MASKSTATS : E_PLAIN, Min, ce.cinp, ce.coutp, nullptr, ce.coutp,
                           e.dze[dze_idx(e, b.index - 1)], ce.raw, ce.coef_b, 0, wd(ce), e.gx[0], ce.cinp,
                           stats_here ? cpp->raw : nullptr, stats_here ? e.blocks[b.index - 2].mask1 : nullptr,
                           stats_here ? cpp->sums_b : nullptr, stats_here ? cpp->coutp : 0, main_ctas(b.hin), ce.pack_d));
      if (e.fork && e.fork_wgrad && e.bn_fuse) pl.dgrad[b.ce].gsrc = src_b(e, ce, (float)Min);
    }
    for (int ev = 0; ev < 2; ++ev)
      ENG_TRY(pw_gemm_plan(ev ? &pl.fwd_eval[b.cp] : &pl.fwd[b.cp], A_BNRELU, ev ? E_PLAIN : E_STATS, Mout, cp.coutp,
                           cp.cinp, nullptr, cp.cinp, cd.raw, nullptr, cd.coef_f, 1, wf(cp), cp.raw,
                           cp.coutp, nullptr, nullptr, cp.sums_f, cp.coutp, 0, cp.pack_f));
    if (e.bn_fuse) pl.fwd[b.cp].asrc = src_f(e, cd, (float)Mout);     // training plan only: eval reads running-stat coefficients
    // project dgrad: dzd[M, hidden] = mask(dp[M, cout] * W[cout, hidden])   (weights transposed: [hidden][cout])
    ENG_TRY(pw_gemm_plan(&pl.dgrad[b.cp], A_TMA, E_MASKSTATS, Mout, cp.cinp, cp.coutp, e.dp[dp_idx(b.index - 1)], cp.coutp, nullptr, nullptr,
                         nullptr, 0, wd(cp), e.dzd, cp.cinp, cd.raw, cd.coef_f, cd.sums_b, cd.coutp, main_ctas(b.hout), cp.pack_d));
  }
  {
    const int li = nc - 1;
    const Conv& cl = e.convs[li];
    const Block& lb = e.blocks.back();
    const int M = n * cl.hin * cl.hin;
    for (int ev = 0; ev < 2; ++ev)
      ENG_TRY(pw_gemm_plan(ev ? &pl.fwd_eval[li] : &pl.fwd[li], A_TMA, ev ? E_PLAIN : E_STATS, M, cl.coutp, cl.cinp,
                           lb.xout, cl.cinp, nullptr, nullptr, nullptr, 0, e.w_bf16 + cl.off_w, cl.raw, cl.coutp, nullptr,
                           nullptr, cl.sums_f, cl.coutp, 0));
    const Conv& cpl = e.convs[lb.cp];
    const bool jd = e.join_in_dgrad;
    ENG_TRY(pw_gemm_plan(&pl.dgrad[li], A_BNGRAD, jd ? E_MASKSTATS : E_PLAIN, M, cl.cinp, cl.coutp, nullptr, cl.coutp, e.dze[0], cl.raw,
                         cl.coef_b, 0, e.wt_bf16 + cl.off_w, e.gx[0], cl.cinp, jd ? cpl.raw : nullptr, jd ? lb.mask1 : nullptr,
                         jd ? cpl.sums_b : nullptr, jd ? cpl.coutp : 0, 0));
    if (e.fork && e.fork_wgrad && e.bn_fuse) pl.dgrad[li].gsrc = src_b(e, cl, (float)M);
  }
  return B200DP_OK;
}

// A dgrad plan writes to a fixed output buffer chosen at plan time; the executor alternates gx[0]/gx[1], so the
// output tensor map is re-encoded when the target changes (cheap; host only).
int retarget_out(PwGemmPlan& pl, void* out) {
  return make_tmap_2d_bf16(&pl.tmO, out, (uint64_t)pl.p.M, (uint64_t)pl.p.N, (uint64_t)pl.p.ldo, 64, 32);
}

int bn_fin_f(Engine& e, Conv& c, float count, cudaStream_t s) {
  return bn_fwd_finalize_launch(c.sums_f, c.coutp, e.params + c.off_g, e.params + c.off_b, count, e.eps, e.momentum,
                                e.buffers + c.off_rm, e.buffers + c.off_rv, c.coef_f, c.saved, c.coutp, s);
}
int bn_fin_b(Engine& e, Conv& c, float count, cudaStream_t s) {
  return bn_bwd_finalize_launch(c.sums_b, c.coutp, c.saved, e.params + c.off_g, count, c.coef_b, e.grads + c.off_g,
                                e.grads + c.off_b, c.coutp, s);
}

// Fork / join helpers.  They work the same on real streams and under stream capture (where they become graph edges).
cudaEvent_t next_event(Engine& e) {
  if (e.ev_next == e.ev_pool.size()) {
    cudaEvent_t ev;
    cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    e.ev_pool.push_back(ev);
  }
  return e.ev_pool[e.ev_next++];
}
// everything enqueued on `from` so far happens before whatever is enqueued on `to` from now on
int order_after(Engine& e, cudaStream_t from, cudaStream_t to) {
  if (from == to) return B200DP_OK;
  cudaEvent_t ev = next_event(e);
  B200DP_CUDA(cudaEventRecord(ev, from));
  B200DP_CUDA(cudaStreamWaitEvent(to, ev, 0));
  return B200DP_OK;
}

int forward(Engine& e, Plans& pl, int n, bool train, bool has_aug, cudaStream_t s, bool pre = false) {
  // the bf16 weight copies do not depend on the input, the preprocessing does not depend on the weights: two branches
  e.ev_next = 0;
  cudaStream_t sw = (e.fork && !e.profiling) ? e.side[0] : s;
  ENG_TRY(order_after(e, s, sw));
  ENG_TRY(weight_prep_launch(e.params, e.w_bf16, e.wt_bf16, e.w_pack, e.wprep_dev, e.n_pw, sw));
  mark(e, s, "weight_prep");
  if (e.stem_gemm) ENG_TRY(stem_weight_prep_launch(e.params + e.convs[0].off_w, e.stem_w, e.stem_pack, sw));
  if (train) {
    B200DP_CUDA(cudaMemsetAsync(e.sums_all, 0, e.sums_bytes, s));
  } else if (!e.eval_coef_valid) {
    for (auto& c : e.convs)
      ENG_TRY(bn_eval_coef_launch(e.params + c.off_g, e.params + c.off_b, e.buffers + c.off_rm, e.buffers + c.off_rv,
                                  e.eps, c.coef_f, c.coutp, s));
    mark(e, s, "bn_eval_coef");
  }
  // input pipeline: straight into the stem's im2col matrix when the stem is a GEMM (no 224x224 image in between)
  const bool fused_col = !pre && e.stem_gemm && e.fuse_col;
  if (fused_col)
    ENG_TRY(b200dp_preprocess_col(e.in_images, (train && has_aug) ? e.in_aug : nullptr, e.x0col, train ? e.stem_asum : nullptr,
                                  n, e.S, s));
  else if (!pre)
    ENG_TRY(b200dp_preprocess(e.in_images, (train && has_aug) ? e.in_aug : nullptr, e.x0, n, e.S, 4, s));
  mark(e, s, "preprocess");
  Conv& st = e.convs[0];
  if (e.stem_gemm) {
    if (!fused_col) ENG_TRY(stem_im2col_launch(e.x0, e.x0col, train ? e.stem_asum : nullptr, n, e.S, e.S, s));
    mark(e, s, "stem_im2col", 0);
    ENG_TRY(order_after(e, sw, s));
    ENG_TRY(pw_gemm_launch(train ? pl.fwd[0] : pl.fwd_eval[0], s));
  } else {
    ENG_TRY(order_after(e, sw, s));
    ENG_TRY(stem_fwd_launch(e.x0, e.params + st.off_w, st.raw, train ? st.sums_f : nullptr, st.coutp, n, e.S, e.S, s));
  }
  mark(e, s, "stem_fwd", 0);
  const bool fuse = train && e.bn_fuse;
  if (train && !fuse) ENG_TRY(bn_fin_f(e, st, (float)n * st.hout * st.hout, s));
  for (auto& b : e.blocks) {
    Conv& cd = e.convs[b.cd];
    Conv& cp = e.convs[b.cp];
    Conv* cin_raw = &st;
    if (b.ce >= 0) {
      Conv& ce = e.convs[b.ce];
      ENG_TRY(pw_gemm_launch(train ? pl.fwd[b.ce] : pl.fwd_eval[b.ce], s));
      mark(e, s, "pw_fwd_expand", b.ce);
      if (train && !fuse) ENG_TRY(bn_fin_f(e, ce, (float)n * ce.hout * ce.hout, s));
      cin_raw = &ce;
    }
    const BnFwdSrc in_src = src_f(e, *cin_raw, (float)n * cin_raw->hout * cin_raw->hout);
    if (e.dw_tiled && dw_tile_fwd_supported(n, cd.hin, cd.hin, cd.coutp, cd.stride))
      ENG_TRY(dw_tile_fwd_launch(cin_raw->raw, cin_raw->coef_f, cin_raw->coef_f + cin_raw->coutp, e.params + cd.off_w,
                                 cd.raw, train ? cd.sums_f : nullptr, cd.coutp, n, cd.hin, cd.hin, cd.coutp, cd.stride, s,
                                 fuse ? &in_src : nullptr));
    else
      ENG_TRY(dw_fwd_launch(cin_raw->raw, cin_raw->coef_f, cin_raw->coef_f + cin_raw->coutp, e.params + cd.off_w, cd.raw,
                            train ? cd.sums_f : nullptr, cd.coutp, n, cd.hin, cd.hin, cd.coutp, cd.stride, s,
                            fuse ? &in_src : nullptr));
    mark(e, s, "dw_fwd", b.cd);
    if (train && !fuse) ENG_TRY(bn_fin_f(e, cd, (float)n * cd.hout * cd.hout, s));
    ENG_TRY(pw_gemm_launch(train ? pl.fwd[b.cp] : pl.fwd_eval[b.cp], s));       // fused: its plan carries src_f(cd)
    mark(e, s, "pw_fwd_project", b.cp);
    if (train && !fuse) ENG_TRY(bn_fin_f(e, cp, (float)n * cp.hout * cp.hout, s));
    const void* res = b.use_res ? e.blocks[b.index - 2].xout : nullptr;
    const BnFwdSrc cp_src = src_f(e, cp, (float)n * cp.hout * cp.hout);
    ENG_TRY(block_out_launch(cp.raw, res, cp.coef_f, b.xout, train ? b.xsum : nullptr, (long long)n * b.hout * b.hout,
                             cp.coutp, s, fuse ? &cp_src : nullptr));
    mark(e, s, "block_out", b.cp);
  }
  const int li = (int)e.convs.size() - 1;
  Conv& cl = e.convs[li];
  ENG_TRY(pw_gemm_launch(train ? pl.fwd[li] : pl.fwd_eval[li], s));
  mark(e, s, "pw_fwd_last", li);
  if (train && !fuse) ENG_TRY(bn_fin_f(e, cl, (float)n * cl.hout * cl.hout, s));
  const int HW = cl.hout * cl.hout;
  const BnFwdSrc cl_src = src_f(e, cl, (float)n * HW);
  ENG_TRY(head_pool_launch(cl.raw, cl.coef_f, e.pooled, train ? e.drop : nullptr, n, HW, cl.coutp, train ? e.p_drop : 0.f,
                           e.seed, e.step_dev, s, fuse ? &cl_src : nullptr));
  ENG_TRY(head_fc_launch(e.pooled, e.params + e.off_fcw, e.params + e.off_fcb, e.logits, n, cl.coutp, e.classes, s));
  ENG_TRY(b200dp_softmax_xent(e.logits, e.in_labels, nullptr, train ? e.dlogits : nullptr,
                              train ? e.metrics_train : e.metrics_eval, n, e.classes, (float)n, s));
  mark(e, s, "head_fwd");
  return B200DP_OK;
}

__global__ void bump_step_kernel(int* step) {
  pdl_trigger();
  pdl_wait();
  *step += 1;
}

// Backward pass.  The main chain is the data-gradient path (grad_join -> affine -> project dgrad -> depthwise backward
// -> expand dgrad); the pointwise weight gradients only READ what that chain produces, so each is forked onto a side
// branch right after its inputs exist and joined again when its input buffer is about to be overwritten (two blocks
// later: dze / dp are double-buffered) or at the end of the step.
int backward(Engine& e, Plans& pl, int n, cudaStream_t s) {
  const bool forked = e.fork && e.fork_wgrad && !e.profiling;
  cudaStream_t sp = forked ? e.side[0] : s;      // project-conv weight gradients
  cudaStream_t se = forked ? e.side[1] : s;      // expand / last / stem weight gradients
  cudaStream_t const sp_side = sp, se_side = se;
  bool sp_used = false, se_used = false;
  cudaEvent_t dze_free[2] = {nullptr, nullptr}, dp_free[2] = {nullptr, nullptr};   // last side-branch reader done
  auto wait_free = [&](cudaEvent_t& ev) -> int {
    if (ev != nullptr) B200DP_CUDA(cudaStreamWaitEvent(s, ev, 0));
    ev = nullptr;
    return B200DP_OK;
  };
  auto mark_free = [&](cudaEvent_t& ev, cudaStream_t side) -> int {
    if (side == s) return B200DP_OK;
    ev = next_event(e);
    B200DP_CUDA(cudaEventRecord(ev, side));
    return B200DP_OK;
  };
  // ---- communication branch: bucket k covers the arena from the first conv of block bucket_first[k] up to the
  // start of bucket k-1 (bucket 0 runs to the end of the arena: last conv + classifier)
  cudaStream_t sc = (e.fork && !e.profiling) ? e.side[2] : s;
  size_t next_bucket = 0;
  long long bucket_hi = e.param_total;
  bool sc_used = false;
  auto launch_bucket = [&](long long lo, bool last) -> int {
    if (sc != s) {
      ENG_TRY(order_after(e, s, sc));
      if (sp_used) ENG_TRY(order_after(e, sp_side, sc));
      if (se_used) ENG_TRY(order_after(e, se_side, sc));
      sc_used = true;
    }
    // Early buckets: waiting and moving data are separate launches -- a ONE-CTA barrier kernel absorbs the skew
    // between ranks while the compute kernels keep every other SM, then a wide, short launch reduces + updates the
    // bucket without waiting for anybody.  The LAST bucket is one launch with both barriers inside: its second
    // barrier tells every rank that all its weights are complete and all its gradients have been read, so it also
    // clears the gradients of the earlier buckets and advances the step count.
    b200dp_comm_adam a = e.comm;
    const bool multi = a.world > 1;
    a.range_lo = lo;
    a.range_hi = bucket_hi;
    a.exact_blocks = 0;
    a.max_blocks = e.comm_blocks;
    if (!last) {
      if (multi) ENG_TRY(comm_barrier_launch(a, 1, sc));
      a.skip_barriers = multi ? 1 : 0;
      a.zero_grads = multi ? 0 : 1;
      a.bump_step = 0;
    } else {
      a.skip_barriers = 0;
      a.zero_grads = 1;
      a.bump_step = 1;
      if (multi && bucket_hi < e.param_total) { a.zero_lo = bucket_hi; a.zero_hi = e.param_total; }
    }
    ENG_TRY(comm_adam_launch(a, sc));
    mark(e, s, "comm_adam");
    bucket_hi = lo;
    return B200DP_OK;
  };
  const int li = (int)e.convs.size() - 1;
  Conv& cl = e.convs[li];
  const int HW = cl.hout * cl.hout;
  ENG_TRY(head_bwd_launch(e.dlogits, e.pooled, e.drop, e.params + e.off_fcw, e.dpool, e.grads + e.off_fcw,
                          e.grads + e.off_fcb, cl.raw, cl.coef_f, e.dze[0], cl.sums_b, cl.coutp, n, HW, cl.coutp,
                          e.classes, s));
  mark(e, s, "head_bwd");
  const bool fuse = e.bn_fuse;
  if (!fuse) ENG_TRY(bn_fin_b(e, cl, (float)n * HW, s));
  Block& lb = e.blocks.back();
  const int Ml = n * HW;
  const BnBwdSrc cl_src = src_b(e, cl, (float)n * HW);
  ENG_TRY(order_after(e, s, se));
  se_used |= se != s;
  ENG_TRY(pw_wgrad_launch(0, Ml, cl.coutp, cl.cinp, lb.xout, cl.cinp, e.dze[0], cl.raw, cl.coutp, cl.coef_b, lb.xsum,
                          e.grads + cl.off_w, cl.cinp, 0, se, 1, fuse ? &cl_src : nullptr));
  mark(e, s, "pw_wgrad_last", li);
  ENG_TRY(mark_free(dze_free[0], se));
  int gxi = 0, doi = 0;
  ENG_TRY(retarget_out(pl.dgrad[li], e.gx[gxi]));
  ENG_TRY(pw_gemm_launch(pl.dgrad[li], s));
  mark(e, s, "pw_dgrad_last", li);
  const __nv_bfloat16* g1 = e.gx[gxi];         // gradient arriving from the consumer's data-gradient
  const __nv_bfloat16* dnext = nullptr;        // dout of the next block (valid when that block has a residual)
  bool next_res = false;
  for (int bi = (int)e.blocks.size() - 1; bi >= 0; --bi) {
    Block& b = e.blocks[bi];
    Conv& cd = e.convs[b.cd];
    Conv& cp = e.convs[b.cp];
    // the big early layers fill the machine on their own: a concurrent weight gradient only takes SMs and L2 away
    // from the critical path there (measured), so they stay on the main chain
    sp = b.hout <= e.fork_max_h ? sp_side : s;
    se = b.hin <= e.fork_max_h ? se_side : s;
    // SM budgets of a forked block: the weight-gradient launches size their grids for fork_sms SMs, everything on
    // the main chain for the rest (RAII: back to the whole device at the end of the iteration)
    struct Budget { int saved; Budget() : saved(sm_budget()) {} ~Budget() { sm_budget() = saved; } } budget_guard;
    const int total_sms = num_sms();
    const int side_sms = (sp != s || se != s) ? e.fork_sms : 0;
    auto budget_main = [&]() { sm_budget() = side_sms ? total_sms - side_sms : 0; };
    auto budget_side = [&]() { sm_budget() = side_sms; };
    budget_main();
    __nv_bfloat16* dp = e.dp[dp_idx(bi)];
    __nv_bfloat16* dze = e.dze[dze_idx(e, bi)];
    const long long rows_out = (long long)n * b.hout * b.hout;
    // ---- gradient w.r.t. this block's output: consumer dgrad (+ residual path of the next block)
    const __nv_bfloat16* dcur = g1;
    if (next_res) {
      ENG_TRY(grad_join_stats_launch(g1, dnext, cp.raw, e.dout[doi], cp.sums_b, cp.coutp, rows_out, cp.coutp, s));
      dcur = e.dout[doi];
      doi ^= 1;
    } else if (!(e.join_in_dgrad && (bi == (int)e.blocks.size() - 1 || e.blocks[bi + 1].ce >= 0))) {
      ENG_TRY(grad_join_stats_launch(g1, nullptr, cp.raw, nullptr, cp.sums_b, cp.coutp, rows_out, cp.coutp, s));
    }   // else: the data-gradient GEMM that produced g1 has already accumulated the two sums (E_MASKSTATS epilogue)
    mark(e, s, "grad_join", b.cp);
    if (!fuse) ENG_TRY(bn_fin_b(e, cp, (float)rows_out, s));
    const BnBwdSrc cp_src = src_b(e, cp, (float)rows_out);
    ENG_TRY(wait_free(dp_free[dp_idx(bi)]));
    ENG_TRY(affine_launch(dcur, cp.raw, cp.coef_b, 1, dp, rows_out, cp.coutp, s, fuse ? &cp_src : nullptr));
    mark(e, s, "bn_bwd_apply", b.cp);
    // ---- projection conv: weight gradient (side branch), data gradient (+ReLU6 mask, statistics for the dw BatchNorm)
    ENG_TRY(order_after(e, s, sp));
    sp_used |= sp != s;
    if (sp != s) budget_side();
    ENG_TRY(pw_wgrad_launch(1, (int)rows_out, cp.coutp, cp.cinp, dp, cp.coutp, cd.raw, nullptr, cp.cinp, cd.coef_f,
                            nullptr, e.grads + cp.off_w, cp.cinp, 0, sp,
                            e.pw_pack ? pw_wgrad_pick_pack(1, (int)rows_out, cp.coutp, cp.cinp) : 1));
    mark(e, s, "pw_wgrad_project", b.cp);
    budget_main();
    ENG_TRY(mark_free(dp_free[dp_idx(bi)], sp));
    ENG_TRY(pw_gemm_launch(pl.dgrad[b.cp], s));
    mark(e, s, "pw_dgrad_project", b.cp);
    if (!fuse) ENG_TRY(bn_fin_b(e, cd, (float)rows_out, s));
    // ---- depthwise conv backward (data + weight gradient, mask, statistics for the producer's BatchNorm)
    Conv& cin_raw = b.ce >= 0 ? e.convs[b.ce] : e.convs[0];
    const BnBwdSrc cd_src = src_b(e, cd, (float)rows_out);
    ENG_TRY(wait_free(dze_free[dze_idx(e, bi)]));
    if (e.dw_tiled && dw_tile_supported(n, cd.hin, cd.hin, cd.coutp, cd.stride))
      ENG_TRY(dw_tile_bwd_launch(cin_raw.raw, cin_raw.coef_f, cin_raw.coef_f + cin_raw.coutp, e.dzd, cd.raw, cd.coef_b,
                                 e.params + cd.off_w, dze, cin_raw.sums_b, cin_raw.coutp, e.grads + cd.off_w, n, cd.hin,
                                 cd.hin, cd.coutp, s, fuse ? &cd_src : nullptr));
    else if (e.dw_tiled && cd.stride == 2 && dw_tile_bwd_supported(n, cd.hin, cd.hin, cd.coutp, 2))
      ENG_TRY(dw_tile_bwd2_launch(cin_raw.raw, cin_raw.coef_f, cin_raw.coef_f + cin_raw.coutp, e.dzd, cd.raw, cd.coef_b,
                                  e.params + cd.off_w, dze, cin_raw.sums_b, cin_raw.coutp, e.grads + cd.off_w, n, cd.hin,
                                  cd.hin, cd.coutp, s, fuse ? &cd_src : nullptr));
    else
      ENG_TRY(dw_bwd_launch(cin_raw.raw, cin_raw.coef_f, cin_raw.coef_f + cin_raw.coutp, e.dzd, cd.raw, cd.coef_b,
                            e.params + cd.off_w, dze, cin_raw.sums_b, cin_raw.coutp, e.grads + cd.off_w, n, cd.hin,
                            cd.hin, cd.coutp, cd.stride, s, fuse ? &cd_src : nullptr));
    mark(e, s, "dw_bwd", b.cd);
    const long long rows_in = (long long)n * b.hin * b.hin;
    const bool in_fused = fuse && (b.ce >= 0 || e.stem_gemm);      // the CUDA-core stem weight gradient reads the array
    if (!in_fused) ENG_TRY(bn_fin_b(e, cin_raw, (float)rows_in, s));
    const BnBwdSrc in_src = src_b(e, cin_raw, (float)rows_in);
    ENG_TRY(order_after(e, s, se));
    se_used |= se != s;
    if (se != s) budget_side();
    if (b.ce >= 0) {
      Conv& ce = e.convs[b.ce];
      const __nv_bfloat16* xin = e.blocks[bi - 1].xout;
      ENG_TRY(pw_wgrad_launch(0, (int)rows_in, ce.coutp, ce.cinp, xin, ce.cinp, dze, ce.raw, ce.coutp, ce.coef_b,
                              e.blocks[bi - 1].xsum, e.grads + ce.off_w, ce.cinp, 0, se,
                              e.pw_pack ? pw_wgrad_pick_pack(0, (int)rows_in, ce.coutp, ce.cinp) : 1,
                              in_fused ? &in_src : nullptr));
      mark(e, s, "pw_wgrad_expand", b.ce);
      budget_main();
      ENG_TRY(mark_free(dze_free[dze_idx(e, bi)], se));
      gxi ^= 1;
      ENG_TRY(retarget_out(pl.dgrad[b.ce], e.gx[gxi]));
      ENG_TRY(pw_gemm_launch(pl.dgrad[b.ce], s));
      mark(e, s, "pw_dgrad_expand", b.ce);
      g1 = e.gx[gxi];
      dnext = dcur;
      next_res = b.use_res;
    } else {
      Conv& st = e.convs[0];
      if (e.stem_gemm) {
        const int M = n * st.hout * st.hout;
        ENG_TRY(pw_wgrad_launch(0, M, 32, 32, e.x0col, 32, dze, st.raw, 32, st.coef_b, e.stem_asum, e.stem_dw, 32, 0, se,
                                pw_wgrad_pick_pack(0, M, 32, 32), in_fused ? &in_src : nullptr));
        ENG_TRY(stem_dw_fold_launch(e.stem_dw, e.grads + st.off_w, se));
      } else {
        ENG_TRY(stem_wgrad_launch(e.x0, dze, st.raw, st.coef_b, e.grads + st.off_w, n, e.S, e.S, se));
      }
      mark(e, s, "stem_wgrad", 0);
    }
    // every gradient of blocks >= bi exists now: hand the finished bucket to the communication branch
    if (e.comm_on) {
      if (bi == 0) {
        ENG_TRY(launch_bucket(0, true));
      } else if (next_bucket < e.bucket_first.size() && e.bucket_first[next_bucket] == bi) {
        const Conv& first = e.convs[b.ce >= 0 ? b.ce : b.cd];
        ENG_TRY(launch_bucket(first.off_w, false));
        ++next_bucket;
      }
    }
  }
  if (sc_used) ENG_TRY(order_after(e, sc, s));
  // join the side branches: the step (and the capture) ends on the main stream
  if (sp_used) ENG_TRY(order_after(e, sp_side, s));
  if (se_used) ENG_TRY(order_after(e, se_side, s));
  B200DP_CUDA(launch_k(bump_step_kernel, dim3(1), dim3(1), 0, s, e.step_dev));
  B200DP_CHECK_LAUNCH();
  return B200DP_OK;
}

int stage_inputs(Engine& e, const uint8_t* images, const int64_t* labels, const float* aug, int n, cudaStream_t s) {
  if (images != e.in_images)
    B200DP_CUDA(cudaMemcpyAsync(e.in_images, images, (size_t)n * 3072, cudaMemcpyDeviceToDevice, s));
  if (labels != e.in_labels)
    B200DP_CUDA(cudaMemcpyAsync(e.in_labels, labels, (size_t)n * 8, cudaMemcpyDeviceToDevice, s));
  if (aug != nullptr && aug != e.in_aug)
    B200DP_CUDA(cudaMemcpyAsync(e.in_aug, aug, (size_t)n * B200DP_AUG_NPARAM * 4, cudaMemcpyDeviceToDevice, s));
  return B200DP_OK;
}

int get_plans(Engine& e, int n, Plans** out) {
  auto it = e.plans.find(n);
  if (it == e.plans.end()) {
    Plans pl;
    ENG_TRY(build_plans(e, pl, n));
    it = e.plans.emplace(n, std::move(pl)).first;
  }
  *out = &it->second;
  return B200DP_OK;
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" {

int b200dp_engine_create(void** out, int batch, int img_size, int num_classes, const long long* offsets, int n_offsets,
                         long long param_total, long long buffer_total, float* params, float* grads, float* buffers,
                         float dropout, float bn_eps, float bn_momentum, unsigned seed, int use_graphs) {
  B200DP_REQUIRE(out && offsets && params && grads && buffers, "engine_create: null argument");
  B200DP_REQUIRE(batch >= 1 && batch <= 1024 && img_size % 32 == 0 && img_size >= 32 && img_size <= 256,
                 "engine_create: batch=%d img_size=%d unsupported", batch, img_size);
  B200DP_REQUIRE(n_offsets == 52 * 5 + 2, "engine_create: offsets table has %d entries, expected %d", n_offsets, 52 * 5 + 2);
  Engine* e = new Engine();
  e->B = batch; e->S = img_size; e->classes = num_classes;
  e->p_drop = dropout; e->eps = bn_eps; e->momentum = bn_momentum; e->seed = seed;
  e->params = params; e->grads = grads; e->buffers = buffers;
  e->param_total = param_total; e->buffer_total = buffer_total;
  e->use_graphs = use_graphs != 0;
  {
    const char* impl = getenv("B200DP_DW_IMPL");
    e->dw_tiled = !(impl != nullptr && std::string(impl) == "simt");
  }
  build_spec(*e, offsets);
  e->pw_pack = !(getenv("B200DP_PW_PACK") != nullptr && atoi(getenv("B200DP_PW_PACK")) == 0);
  e->bn_fuse = !(getenv("B200DP_BN_FUSE") != nullptr && atoi(getenv("B200DP_BN_FUSE")) == 0);
  e->fork = !(getenv("B200DP_FORK") != nullptr && atoi(getenv("B200DP_FORK")) == 0);
  e->join_in_dgrad = getenv("B200DP_JOIN_IN_DGRAD") != nullptr && atoi(getenv("B200DP_JOIN_IN_DGRAD")) != 0;
  e->fork_wgrad = getenv("B200DP_FORK_WGRAD") != nullptr && atoi(getenv("B200DP_FORK_WGRAD")) != 0;
  if (getenv("B200DP_FORK_MAXH") != nullptr) e->fork_max_h = atoi(getenv("B200DP_FORK_MAXH"));
  if (getenv("B200DP_FORK_SMS") != nullptr) e->fork_sms = atoi(getenv("B200DP_FORK_SMS"));
  {
    const char* st = getenv("B200DP_STEM");
    e->stem_gemm = !(st != nullptr && std::string(st) == "simt");
    e->stem_pack = e->pw_pack ? 2 : 1;        // K = N = 32: two pixels per 128-byte row
    e->fuse_col = b200dp_preprocess_col_supported(img_size) != 0 &&
                  !(getenv("B200DP_FUSE_COL") != nullptr && atoi(getenv("B200DP_FUSE_COL")) == 0);
  }
  if (e->pw_pack) {
    for (auto& c : e->convs) {
      if (c.kind != 2) continue;
      const int rows = c.hout * c.hout;      // per image; every packed layer has rows % 4 == 0, so any batch divides
      c.pack_f = pw_gemm_pick_pack(rows, c.coutp, c.cinp);
      c.pack_d = pw_gemm_pick_pack(rows, c.cinp, c.coutp);
      if (c.pack_f > 1) { c.off_pf = e->pack_total; e->pack_total += (long long)c.pack_f * c.pack_f * c.coutp * c.cinp; }
      if (c.pack_d > 1) { c.off_pd = e->pack_total; e->pack_total += (long long)c.pack_d * c.pack_d * c.coutp * c.cinp; }
    }
  }
  e->ws_bytes = layout_workspace(*e, nullptr) + 1024;
  cudaError_t err = cudaMalloc(&e->ws, e->ws_bytes);
  if (err != cudaSuccess) {
    set_error("engine_create: cudaMalloc(%zu MB) failed: %s", e->ws_bytes >> 20, cudaGetErrorString(err));
    delete e;
    return B200DP_ENOMEM;
  }
  B200DP_CUDA(cudaMemset(e->ws, 0, e->ws_bytes));
  layout_workspace(*e, e->ws);
  std::vector<WPrepEntry> table;
  for (auto& c : e->convs)
    if (c.kind == 2) table.push_back(WPrepEntry{c.off_w, c.coutp, c.cinp, c.pack_f, c.pack_d, c.off_pf, c.off_pd});
  e->n_pw = (int)table.size();
  B200DP_CUDA(cudaMemcpy(e->wprep_dev, table.data(), table.size() * sizeof(WPrepEntry), cudaMemcpyHostToDevice));
  for (auto& b : e->blocks) {
    std::vector<float> one(2 * b.coutp, 0.f);
    for (int i = b.coutp; i < 2 * b.coutp; ++i) one[i] = 3.f;
    B200DP_CUDA(cudaMemcpy(b.mask1, one.data(), one.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  B200DP_CUDA(cudaStreamCreateWithFlags(&e->cap_stream, cudaStreamNonBlocking));
  {
    int prio_lo = 0, prio_hi = 0;     // numerically lowest = highest priority
    B200DP_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int i = 0; i < 3; ++i)       // the communication branch outranks compute: its few CTAs take the first free slots
      B200DP_CUDA(cudaStreamCreateWithPriority(&e->side[i], cudaStreamNonBlocking, i == 2 ? prio_hi : prio_lo));
  }
  B200DP_CUDA(cudaDeviceSynchronize());
  *out = e;
  return B200DP_OK;
}

int b200dp_engine_destroy(void* h) {
  Engine* e = static_cast<Engine*>(h);
  if (!e) return B200DP_OK;
  cudaDeviceSynchronize();
  for (auto& kv : e->plans) {
    if (kv.second.train_graph) cudaGraphExecDestroy(kv.second.train_graph);
    if (kv.second.eval_graph) cudaGraphExecDestroy(kv.second.eval_graph);
    if (kv.second.eval_pre_graph) cudaGraphExecDestroy(kv.second.eval_pre_graph);
  }
  if (e->cap_stream) cudaStreamDestroy(e->cap_stream);
  for (auto& st : e->side) if (st) cudaStreamDestroy(st);
  for (auto ev : e->ev_pool) cudaEventDestroy(ev);
  if (e->ws) cudaFree(e->ws);
  delete e;
  return B200DP_OK;
}

size_t b200dp_engine_workspace_bytes(void* h) { return static_cast<Engine*>(h)->ws_bytes; }

// Pointers into the engine's workspace: static input staging buffers, logits, metric accumulators
// (double[3]: loss sum, correct, count).  which: 0 images u8, 1 labels i64, 2 aug params f32, 3 logits f32,
// 4 train metrics f64, 5 eval metrics f64, 6 pre-processed input bf16 [B,S,S,4]
void* b200dp_engine_buffer(void* h, int which) {
  Engine* e = static_cast<Engine*>(h);
  switch (which) {
    case 0: return e->in_images;
    case 1: return e->in_labels;
    case 2: return e->in_aug;
    case 3: return e->logits;
    case 4: return e->metrics_train;
    case 5: return e->metrics_eval;
    case 6: return e->x0;
    default: return nullptr;
  }
}

// Activation / gradient taps for tests: raw conv output (what=0), forward coef (1), saved mean/invstd (2),
// block output (3, index = block 0..16).
void* b200dp_engine_tap(void* h, int what, int index) {
  Engine* e = static_cast<Engine*>(h);
  if (what == 3) return (index >= 0 && index < (int)e->blocks.size()) ? e->blocks[index].xout : nullptr;
  if (index < 0 || index >= (int)e->convs.size()) return nullptr;
  Conv& c = e->convs[index];
  switch (what) {
    case 0: return c.raw;
    case 1: return c.coef_f;
    case 2: return c.saved;
    default: return nullptr;
  }
}

// One training step: forward + loss + backward.  Gradients are ACCUMULATED into the grads arena given at create
// time (it must be zero on entry -- the fused optimiser kernel zeroes it after consuming it).  `aug` may be null
// (no augmentation).  Metrics accumulate into the train metric buffer.
int b200dp_engine_train_step(void* h, const uint8_t* images, const int64_t* labels, const float* aug, int n,
                             void* stream) {
  Engine* e = static_cast<Engine*>(h);
  B200DP_REQUIRE(e && n >= 1 && n <= e->B, "train_step: batch %d outside [1, %d]", n, e ? e->B : 0);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  ENG_TRY(stage_inputs(*e, images, labels, aug, n, s));
  Plans* pl;
  ENG_TRY(get_plans(*e, n, &pl));
  e->eval_coef_valid = false;
  const bool has_aug = aug != nullptr;
  if (!e->use_graphs || e->profiling) {
    ENG_TRY(forward(*e, *pl, n, true, has_aug, s));
    return backward(*e, *pl, n, s);
  }
  cudaGraphExec_t& g = pl->train_graph;
  if (g == nullptr || pl->train_aug != (int)has_aug) {
    if (g) { cudaGraphExecDestroy(g); g = nullptr; }
    cudaGraph_t graph;
    cudaStream_t cs = e->cap_stream;
    const unsigned long long before = g_launches.load();
    B200DP_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    int rc = forward(*e, *pl, n, true, has_aug, cs);
    if (rc == B200DP_OK) rc = backward(*e, *pl, n, cs);
    cudaError_t ce = cudaStreamEndCapture(cs, &graph);
    pl->train_kernels = g_launches.load() - before;     // counted while capturing, executed on every replay
    g_launches.store(before);
    if (rc != B200DP_OK) { if (ce == cudaSuccess) cudaGraphDestroy(graph); return rc; }
    B200DP_CUDA(ce);
    B200DP_CUDA(cudaGraphInstantiate(&g, graph, 0));
    cudaGraphDestroy(graph);
    pl->train_aug = (int)has_aug;
  }
  B200DP_CUDA(cudaGraphLaunch(g, s));
  g_launches.fetch_add(pl->train_kernels, std::memory_order_relaxed);
  return B200DP_OK;
}

// Evaluation step (running statistics, no dropout): logits stay in buffer 3, metrics accumulate into buffer 5.
int b200dp_engine_eval_step(void* h, const uint8_t* images, const int64_t* labels, int n, void* stream) {
  Engine* e = static_cast<Engine*>(h);
  B200DP_REQUIRE(e && n >= 1 && n <= e->B, "eval_step: batch %d outside [1, %d]", n, e ? e->B : 0);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  ENG_TRY(stage_inputs(*e, images, labels, nullptr, n, s));
  Plans* pl;
  ENG_TRY(get_plans(*e, n, &pl));
  if (!e->use_graphs || e->profiling || !e->eval_coef_valid) {
    ENG_TRY(forward(*e, *pl, n, false, false, s));
    e->eval_coef_valid = true;
    return B200DP_OK;
  }
  cudaGraphExec_t& g = pl->eval_graph;
  if (g == nullptr) {
    cudaGraph_t graph;
    cudaStream_t cs = e->cap_stream;
    const unsigned long long before = g_launches.load();
    B200DP_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    int rc = forward(*e, *pl, n, false, false, cs);
    cudaError_t ce = cudaStreamEndCapture(cs, &graph);
    pl->eval_kernels = g_launches.load() - before;
    g_launches.store(before);
    if (rc != B200DP_OK) { if (ce == cudaSuccess) cudaGraphDestroy(graph); return rc; }
    B200DP_CUDA(ce);
    B200DP_CUDA(cudaGraphInstantiate(&g, graph, 0));
    cudaGraphDestroy(graph);
  }
  B200DP_CUDA(cudaGraphLaunch(g, s));
  g_launches.fetch_add(pl->eval_kernels, std::memory_order_relaxed);
  return B200DP_OK;
}

// Evaluation forward of input that is already pre-processed: `x` = normalised bf16 NHWC [n, S, S, 4] (4th channel 0),
// e.g. a full-resolution photograph resized to S x S outside (inference.py).  Logits land in buffer 3.
int b200dp_engine_forward_preprocessed(void* h, const void* x, int n, void* stream) {
  Engine* e = static_cast<Engine*>(h);
  B200DP_REQUIRE(e && x && n >= 1 && n <= e->B, "forward_preprocessed: batch %d outside [1, %d]", n, e ? e->B : 0);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (x != (const void*)e->x0)
    B200DP_CUDA(cudaMemcpyAsync(e->x0, x, (size_t)n * e->S * e->S * 4 * 2, cudaMemcpyDeviceToDevice, s));
  Plans* pl;
  ENG_TRY(get_plans(*e, n, &pl));
  if (!e->use_graphs || e->profiling || !e->eval_coef_valid) {
    ENG_TRY(forward(*e, *pl, n, false, false, s, true));
    e->eval_coef_valid = true;
    return B200DP_OK;
  }
  cudaGraphExec_t& g = pl->eval_pre_graph;
  if (g == nullptr) {
    cudaGraph_t graph;
    cudaStream_t cs = e->cap_stream;
    const unsigned long long before = g_launches.load();
    B200DP_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    int rc = forward(*e, *pl, n, false, false, cs, true);
    cudaError_t ce = cudaStreamEndCapture(cs, &graph);
    pl->eval_kernels = g_launches.load() - before;
    g_launches.store(before);
    if (rc != B200DP_OK) { if (ce == cudaSuccess) cudaGraphDestroy(graph); return rc; }
    B200DP_CUDA(ce);
    B200DP_CUDA(cudaGraphInstantiate(&g, graph, 0));
    cudaGraphDestroy(graph);
  }
  B200DP_CUDA(cudaGraphLaunch(g, s));
  g_launches.fetch_add(pl->eval_kernels, std::memory_order_relaxed);
  return B200DP_OK;
}

// Run the gradient all-reduce + Adam INSIDE the training step (bucketed, overlapped with backward): `args` is the
// fully populated argument block of b200dp_comm_adam_step (copied; range / bump_step / exact_blocks are set per
// bucket here), `bucket_first` the block indices (0..16, strictly descending) at which a new bucket starts, and
// `comm_blocks` the fixed grid of every bucket launch.  args == NULL switches it off again.  Existing step graphs
// are dropped.  Every rank must configure the same buckets.
int b200dp_engine_set_comm(void* h, const b200dp_comm_adam* args, const int* bucket_first, int n_buckets,
                           int comm_blocks) {
  Engine* e = static_cast<Engine*>(h);
  B200DP_REQUIRE(e != nullptr, "set_comm: null engine");
  cudaDeviceSynchronize();
  for (auto& kv : e->plans) {
    if (kv.second.train_graph) { cudaGraphExecDestroy(kv.second.train_graph); kv.second.train_graph = nullptr; }
  }
  e->comm_on = args != nullptr;
  e->bucket_first.clear();
  if (!e->comm_on) return B200DP_OK;
  B200DP_REQUIRE(args->numel == e->param_total, "set_comm: optimiser vector has %lld elements, the arena %lld",
                 args->numel, e->param_total);
  B200DP_REQUIRE(args->grads[args->rank] == (const void*)e->grads && args->params[args->rank] == e->params,
                 "set_comm: the optimiser must work on the engine's gradient and parameter arenas");
  B200DP_REQUIRE(comm_blocks >= 1 && comm_blocks <= B200DP_COMM_MAX_BLOCKS, "set_comm: comm_blocks %d out of range",
                 comm_blocks);
  int prev = (int)e->blocks.size();
  for (int i = 0; i < n_buckets; ++i) {
    B200DP_REQUIRE(bucket_first[i] > 0 && bucket_first[i] < prev, "set_comm: bucket starts must descend within (0, %d)",
                   (int)e->blocks.size());
    prev = bucket_first[i];
    e->bucket_first.push_back(bucket_first[i]);
  }
  e->comm = *args;
  e->comm_blocks = comm_blocks;
  return B200DP_OK;
}

// Tell the engine that parameters / running statistics were changed from outside (load_state_dict, broadcast).
int b200dp_engine_invalidate(void* h) {
  static_cast<Engine*>(h)->eval_coef_valid = false;
  return B200DP_OK;
}

// Per-kernel timing of one un-graphed training step: fills `names` (newline separated) and `ms`.
int b200dp_engine_profile_step(void* h, const uint8_t* images, const int64_t* labels, const float* aug, int n,
                               char* names, int names_cap, float* ms, int ms_cap, int* count, void* stream) {
  Engine* e = static_cast<Engine*>(h);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  e->profiling = true;
  const bool pdl_saved = g_pdl;
  g_pdl = false;                 // per-kernel event timing needs fully serialised launches
  e->prof_events.clear();
  mark(*e, s, "start");
  int rc = b200dp_engine_train_step(h, images, labels, aug, n, stream);
  mark(*e, s, "end");
  e->profiling = false;
  g_pdl = pdl_saved;
  if (rc != B200DP_OK) return rc;
  B200DP_CUDA(cudaStreamSynchronize(s));
  std::string all;
  int k = 0;
  for (size_t i = 1; i < e->prof_events.size(); ++i) {
    float t = 0.f;
    cudaEventElapsedTime(&t, e->prof_events[i - 1].second, e->prof_events[i].second);
    if (k < ms_cap) ms[k] = t;
    all += e->prof_events[i].first + "\n";
    ++k;
  }
  for (auto& pe : e->prof_events) cudaEventDestroy(pe.second);
  e->prof_events.clear();
  snprintf(names, names_cap, "%s", all.c_str());
  *count = k < ms_cap ? k : ms_cap;
  return B200DP_OK;
}

}  // extern "C"
