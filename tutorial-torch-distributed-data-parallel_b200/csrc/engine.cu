// T1 — the native step driver: forward and backward of the toy model as ONE call each.
//
// Reference: `outputs = model(inputs)` and `loss.backward()` in train() (multi-GPU-training-torch.py:121,125) walk
// torchvision's AlexNet (alexnet.py:17-52 + the 10-class head of data_and_toy_model.py:44) layer by layer through
// Python and the autograd engine. Driving OUR ~60 kernels per step the same way costs ~2.0 ms of host time per step
// (measured: one Python autograd node, a handful of tensor allocations and ctypes calls per layer) against 2.3 ms of
// GPU time — every further kernel speed-up would be invisible, and with one process per GPU sharing the host cores a
// rank whose Python thread falls behind stalls the all-reduce of all ranks.
//
// This file is that walk in C++: a plan (shapes, parameter pointers, one step-persistent workspace carved into every
// activation / gate / gradient buffer the step needs) and two functions that enqueue the whole forward or the whole
// backward on the caller's stream by calling the same C-ABI entry points the per-layer modules call, in the same order
// with the same arguments — results are bit-identical to the module path (tests/test_gpu_engine.py). Nothing is
// allocated per step (this also retires the per-step torch.empty of every activation).
//
// Gradient bucketing stays with the caller: backward reports, through a callback, the points at which a group of
// parameters has its final gradient (head + FC2, FC1, conv5..conv2, conv1 — the order gradients become ready,
// SURVEY 3.2); the DDP reducer launches that bucket's all-reduce from inside the callback, so communication still
// overlaps the rest of backward. Weight-gradient kernels of conv5..conv2 run on the plan's side stream (they fill the
// tail waves of the dgrad kernels); the plan joins them before it reports the group.
//
// TF32 mode of the standard geometry only (3x224x224 input, feature map 6x6); everything else uses the module path.
#include "common.cuh"
#include "../../include/b200ddp.h"

#include <new>

namespace b200 {
namespace {

struct ConvGeo { int Cin, H, Cout, K, stride, pad, OH, pool, relu_epilogue, PH; };
// torchvision alexnet.py:22-34 at 224x224: conv(3->64,11,s4,p2)+pool, conv(64->192,5,p2)+pool, conv(192->384,3,p1)+relu,
// conv(384->256,3,p1)+relu, conv(256->256,3,p1)+pool
constexpr ConvGeo kConv[5] = {{3, 224, 64, 11, 4, 2, 55, 1, 0, 27}, {64, 27, 192, 5, 1, 2, 27, 1, 0, 13},
                              {192, 13, 384, 3, 1, 1, 13, 0, 1, 13}, {384, 13, 256, 3, 1, 1, 13, 0, 1, 13},
                              {256, 13, 256, 3, 1, 1, 13, 1, 0, 6}};
constexpr int kFeat = 256 * 6 * 6;

struct Carver {
  char* base;
  size_t off = 0;
  explicit Carver(void* b) : base(reinterpret_cast<char*>(b)) {}
  template <typename T>
  T* take(size_t count) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += (count * sizeof(T) + 255) / 256 * 256;
    return p;
  }
};

struct Plan {
  b200ddp_alexnet_desc d;
  // GEMM copies of the conv weights (conv_weight_prep): forward for all five, dgrad for conv2..5
  float* wf[5]; long long ldf[5];
  float* wd[5]; long long ldd[5];
  // activations (saved for backward)
  float* y[5];            // conv output (pre-pool for pooled layers; post-ReLU for conv3/4)
  float* p[5];            // pooled output (pooled layers only)
  unsigned char* idx[5];
  float* d0; unsigned char* gate0;     // classifier.0 dropout
  float* h1; unsigned char* gate1;
  float* h2; unsigned char* gate2;
  // backward temporaries
  float* dh2; float* dz2; float* dh1; float* dz1; float* dd0; float* dp5;
  float* dzc[5];          // gradient at the conv output (after ReLU / pool backward)
  float* dxc[5];          // gradient at the conv input (dgrad result), layers 1..4
  // scratch
  void* lin_ws; long long lin_ws_bytes;
  void* wg_ws; void* wg_ws_side; long long wg_ws_bytes;
  void* bias_scratch; long long bias_scratch_bytes;
  cudaStream_t side = nullptr;
  cudaEvent_t fork_ev[5], join_ev[5];
  size_t total_bytes = 0;
};

// One layout function serves both the size query (base == null) and plan creation.
size_t layout(Plan& P, void* base) {
  const b200ddp_alexnet_desc& d = P.d;
  const size_t B = (size_t)d.B, H = (size_t)d.hidden;
  Carver c(base);
  for (int l = 0; l < 5; ++l) {
    const ConvGeo& g = kConv[l];
    const long long K = (long long)g.Cin * g.K * g.K;
    P.ldf[l] = (g.Cin % 32 == 0) ? K : (K + 31) / 32 * 32;
    if (l == 0 && d.conv1_order == 2) P.ldf[l] = (g.Cin * ((g.K * 12 + 7) / 8 * 8) + 31) / 32 * 32;      // slab order (conv1.cu)
    P.wf[l] = c.take<float>((size_t)g.Cout * P.ldf[l]);
    P.ldd[l] = (long long)g.Cout * g.K * g.K;
    P.wd[l] = l > 0 ? c.take<float>((size_t)g.Cin * P.ldd[l]) : nullptr;
  }
  for (int l = 0; l < 5; ++l) {
    const ConvGeo& g = kConv[l];
    const size_t out = B * g.Cout * g.OH * g.OH, pooled = B * g.Cout * g.PH * g.PH;
    P.y[l] = c.take<float>(out);
    P.p[l] = g.pool ? c.take<float>(pooled) : nullptr;
    P.idx[l] = g.pool ? c.take<unsigned char>(pooled) : nullptr;
    P.dzc[l] = c.take<float>(out);
    P.dxc[l] = l > 0 ? c.take<float>(B * g.Cin * g.H * g.H) : nullptr;
  }
  P.d0 = c.take<float>(B * kFeat); P.gate0 = c.take<unsigned char>(B * kFeat);
  P.h1 = c.take<float>(B * H); P.gate1 = c.take<unsigned char>(B * H);
  P.h2 = c.take<float>(B * H); P.gate2 = c.take<unsigned char>(B * H);
  P.dh2 = c.take<float>(B * H); P.dz2 = c.take<float>(B * H);
  P.dh1 = c.take<float>(B * H); P.dz1 = c.take<float>(B * H);
  P.dd0 = c.take<float>(B * kFeat); P.dp5 = c.take<float>(B * kFeat);
  long long lw = b200ddp_linear_workspace_bytes(d.B, d.hidden, kFeat);
  long long t = b200ddp_linear_workspace_bytes(d.B, d.hidden, d.hidden);
  if (t > lw) lw = t;
  t = b200ddp_linear_workspace_bytes(d.B, d.num_classes, d.hidden);
  if (t > lw) lw = t;
  P.lin_ws_bytes = lw < 256 ? 256 : lw;
  P.lin_ws = c.take<char>((size_t)P.lin_ws_bytes);
  long long ww = 256, bs = 256;
  // the wgrad split-K plan follows the SM budget in force when a step runs; the partials are largest with every SM
  // available, so size the workspace under that setting
  const int prev_budget = b200ddp_set_sm_budget(0);
  for (int l = 0; l < 5; ++l) {
    const ConvGeo& g = kConv[l];
    const long long need = b200ddp_conv2d_wgrad_workspace_bytes(d.B, g.Cin, g.H, g.H, g.Cout, g.K, g.K, g.stride, g.pad);
    if (need > ww) ww = need;
    const long long b = b200ddp_bias_grad_scratch_bytes(d.B, g.Cout);
    if (b > bs) bs = b;
  }
  b200ddp_set_sm_budget(prev_budget >= num_sms() ? 0 : prev_budget);
  P.wg_ws_bytes = ww;
  P.wg_ws = c.take<char>((size_t)ww);
  P.wg_ws_side = c.take<char>((size_t)ww);
  P.bias_scratch_bytes = bs;
  P.bias_scratch = c.take<char>((size_t)bs);
  return c.off;
}

#define B200_RC(expr)        \
  do {                       \
    int _rc = (expr);        \
    if (_rc) return _rc;     \
  } while (0)

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

long long b200ddp_alexnet_workspace_bytes(int B, int hidden, int num_classes) {
  if (B <= 0 || hidden <= 0 || num_classes <= 0) return -1;
  Plan P{};
  P.d.B = B; P.d.hidden = hidden; P.d.num_classes = num_classes;
  P.d.conv1_order = 2;            // the larger of the two first-layer weight layouts
  return (long long)layout(P, nullptr);
}

int b200ddp_alexnet_plan_create(void** plan_out, const b200ddp_alexnet_desc* desc) {
  B200_CHECK_ARG(plan_out && desc, "alexnet_plan_create: null argument");
  B200_CHECK_ARG(desc->B > 0 && desc->hidden > 0 && desc->num_classes > 0 && desc->num_classes <= 16,
                 "alexnet_plan_create: bad shape (B %d, hidden %d, classes %d)", desc->B, desc->hidden, desc->num_classes);
  B200_CHECK_ARG(desc->workspace && ((uintptr_t)desc->workspace & 255) == 0, "alexnet_plan_create: workspace must be 256-byte aligned");
  B200_CHECK_ARG(desc->conv1_ktab, "alexnet_plan_create: the first layer's offset table is required");
  for (int i = 0; i < 8; ++i) B200_CHECK_ARG(desc->w[i] && desc->b[i], "alexnet_plan_create: parameter %d is null", i);
  B200_DEVICE(desc->device);
  Plan* P = new (std::nothrow) Plan();
  B200_CHECK_ARG(P, "alexnet_plan_create: out of host memory");
  P->d = *desc;
  P->total_bytes = layout(*P, desc->workspace);
  if ((long long)P->total_bytes > desc->workspace_bytes) {
    set_error("alexnet_plan_create: workspace of %lld bytes is too small (need %zu)", desc->workspace_bytes, P->total_bytes);
    delete P;
    return kInvalidArg;
  }
  B200_CUDA(cudaStreamCreateWithFlags(&P->side, cudaStreamNonBlocking));
  for (int l = 0; l < 5; ++l) {
    B200_CUDA(cudaEventCreateWithFlags(&P->fork_ev[l], cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&P->join_ev[l], cudaEventDisableTiming));
  }
  *plan_out = P;
  return 0;
}

int b200ddp_alexnet_plan_destroy(void* plan) {
  if (!plan) return 0;
  Plan* P = reinterpret_cast<Plan*>(plan);
  B200_DEVICE(P->d.device);
  cudaStreamSynchronize(P->side);
  for (int l = 0; l < 5; ++l) {
    cudaEventDestroy(P->fork_ev[l]);
    cudaEventDestroy(P->join_ev[l]);
  }
  cudaStreamDestroy(P->side);
  delete P;
  return 0;
}

// Forward of the whole model on `stream`. x [B][3][224][224]. train != 0: keeps what backward needs (gates, argmax
// codes) and applies dropout with the two seeds (classifier.0 and classifier.1's fused dropout, in module order).
// out: logits [B][classes], or the hidden activations [B][hidden] entering the head when stop_at_hidden != 0
// (evaluate(): the head is then fused with the metrics, eval.cu).
int b200ddp_alexnet_forward(void* plan, const float* x, float* out, int train, unsigned long long seed0,
                            unsigned long long seed1, int stop_at_hidden, void* stream) {
  B200_CHECK_ARG(plan && x && out, "alexnet_forward: null argument");
  Plan& P = *reinterpret_cast<Plan*>(plan);
  const b200ddp_alexnet_desc& d = P.d;
  const int dev = d.device, B = d.B;
  // 1. GEMM copies of the five conv weights (+ the four dgrad copies when a backward will follow): one launch
  b200ddp_weight_prep_job jobs[9];
  int nj = 0;
  for (int l = 0; l < 5; ++l) {
    const ConvGeo& g = kConv[l];
    jobs[nj++] = {d.w[l], P.wf[l], g.Cout, g.Cin, g.K, g.K, P.ldf[l], (g.Cin % 32 == 0) ? 2 : ((l == 0 && d.conv1_order == 2) ? 5 : 0)};
    if (train && l > 0) jobs[nj++] = {d.w[l], P.wd[l], g.Cout, g.Cin, g.K, g.K, P.ldd[l], 3};
  }
  B200_RC(b200ddp_conv2d_weight_prep_multi(jobs, nj, dev, stream));
  // 2. features
  const float* in = x;
  for (int l = 0; l < 5; ++l) {
    const ConvGeo& g = kConv[l];
    const int ko = (g.Cin % 32 == 0) ? 1 : ((l == 0 && d.conv1_order == 2) ? 2 : 0);
    B200_RC(b200ddp_conv2d_fwd(in, P.wf[l], P.ldf[l], d.b[l], P.y[l], ko ? nullptr : d.conv1_ktab, ko, B, g.Cin, g.H, g.H,
                               g.Cout, g.K, g.K, g.stride, g.pad, g.relu_epilogue, dev, stream));
    if (g.pool) {
      B200_RC(b200ddp_relu_maxpool_fwd(P.y[l], P.p[l], train ? P.idx[l] : nullptr, B, g.Cout, g.OH, g.OH, 1, dev, stream));
      in = P.p[l];
    } else {
      in = P.y[l];
    }
  }
  // 3. classifier (the flatten is a view: p5 is [B][256*6*6] already)
  const float pdrop = train ? d.dropout_p : 0.f;
  const float* f = in;
  if (pdrop > 0.f) {
    B200_RC(b200ddp_dropout_fwd(in, P.d0, P.gate0, (long long)B * kFeat, pdrop, seed0, dev, stream));
    f = P.d0;
  }
  B200_RC(b200ddp_linear_fwd(f, d.w[5], d.b[5], P.h1, train ? P.gate1 : nullptr, B, d.hidden, kFeat, 1, pdrop, seed1, 1,
                             P.lin_ws, P.lin_ws_bytes, dev, stream));
  float* h2 = stop_at_hidden ? out : P.h2;
  B200_RC(b200ddp_linear_fwd(P.h1, d.w[6], d.b[6], h2, train ? P.gate2 : nullptr, B, d.hidden, d.hidden, 1, 0.f, 0, 1,
                             P.lin_ws, P.lin_ws_bytes, dev, stream));
  if (stop_at_hidden) return 0;
  return b200ddp_linear_fwd(P.h2, d.w[7], d.b[7], out, nullptr, B, d.num_classes, d.hidden, 0, 0.f, 0, 1, P.lin_ws,
                            P.lin_ws_bytes, dev, stream);
}

// Backward of the forward that ran last with train != 0 (its activations are in the workspace; x = that forward's input).
// dlogits [B][classes];
// dw[i] / db[i], i = conv1..conv5, fc1, fc2, head: where each gradient is WRITTEN (not accumulated; typically the
// parameter's slice of the gradient arena). ready(user, group) is called on the host, in this order, as soon as the
// kernels producing a group's gradients have been enqueued on `stream` (group 0: head + fc2, 1: fc1, 2: conv5..conv2 —
// their side-stream kernels joined into `stream` first —, 3: conv1); null = no reports. use_side_stream = 0 keeps every
// kernel on `stream` (required when the caller will ADD the results to existing gradients on that stream).
int b200ddp_alexnet_backward(void* plan, const float* x, const float* dlogits, float* const* dw, float* const* db,
                             float dropout_p_used,
                             b200ddp_ready_fn ready, void* user, int use_side_stream, void* stream) {
  B200_CHECK_ARG(plan && x && dlogits && dw && db, "alexnet_backward: null argument");
  for (int i = 0; i < 8; ++i) B200_CHECK_ARG(dw[i] && db[i], "alexnet_backward: gradient destination %d is null", i);
  Plan& P = *reinterpret_cast<Plan*>(plan);
  const b200ddp_alexnet_desc& d = P.d;
  const int dev = d.device, B = d.B, H = d.hidden;
  cudaStream_t main_s = as_stream(stream);
  B200_DEVICE(dev);
  // head: no gate
  B200_RC(b200ddp_linear_bwd_dx(dlogits, d.w[7], P.dh2, B, d.num_classes, H, 1, P.lin_ws, P.lin_ws_bytes, dev, stream));
  B200_RC(b200ddp_linear_bwd_dw(dlogits, P.h2, dw[7], db[7], B, d.num_classes, H, 1, dev, stream));
  // fc2 (ReLU gate)
  B200_RC(b200ddp_gate_bias_bwd(P.dh2, P.gate2, 0.f, P.dz2, db[6], B, H, dev, stream));
  B200_RC(b200ddp_linear_bwd_dx(P.dz2, d.w[6], P.dh1, B, H, H, 1, P.lin_ws, P.lin_ws_bytes, dev, stream));
  B200_RC(b200ddp_linear_bwd_dw(P.dz2, P.h1, dw[6], nullptr, B, H, H, 1, dev, stream));
  if (ready) ready(user, 0);
  // fc1 (ReLU + dropout gate); its input is the classifier.0 dropout output when dropout was on
  const bool drop = dropout_p_used > 0.f;
  B200_RC(b200ddp_gate_bias_bwd(P.dh1, P.gate1, dropout_p_used, P.dz1, db[5], B, H, dev, stream));
  B200_RC(b200ddp_linear_bwd_dx(P.dz1, d.w[5], P.dd0, B, H, kFeat, 1, P.lin_ws, P.lin_ws_bytes, dev, stream));
  B200_RC(b200ddp_linear_bwd_dw(P.dz1, drop ? P.d0 : P.p[4], dw[5], nullptr, B, H, kFeat, 1, dev, stream));
  if (ready) ready(user, 1);
  const float* g_in = P.dd0;                 // gradient w.r.t. the flattened pooled features
  if (drop) {
    B200_RC(b200ddp_gate_bias_bwd(P.dd0, P.gate0, dropout_p_used, P.dp5, nullptr, B, kFeat, dev, stream));
    g_in = P.dp5;
  }
  // features, last layer first
  for (int l = 4; l >= 0; --l) {
    const ConvGeo& g = kConv[l];
    if (g.pool)
      B200_RC(b200ddp_relu_maxpool_bwd_bias(g_in, P.idx[l], P.dzc[l], db[l], B, g.Cout, g.OH, g.OH, P.bias_scratch,
                                            P.bias_scratch_bytes, dev, stream));
    else
      B200_RC(b200ddp_relu_bwd_bias(g_in, P.y[l], P.dzc[l], db[l], B, g.Cout, g.OH * g.OH, P.bias_scratch,
                                    P.bias_scratch_bytes, dev, stream));
    const float* xin = l == 0 ? x : (kConv[l - 1].pool ? P.p[l - 1] : P.y[l - 1]);
    const int ko = (g.Cin % 32 == 0) ? 1 : 0;
    if (l > 0) {
      // dgrad first (critical path): the forward kernel on dz with the flipped / transposed weights
      B200_RC(b200ddp_conv2d_fwd(P.dzc[l], P.wd[l], P.ldd[l], nullptr, P.dxc[l], nullptr, 1, B, g.Cout, g.OH, g.OH, g.Cin,
                                 g.K, g.K, 1, g.K - 1 - g.pad, 0, dev, stream));
      void* ws = P.wg_ws;
      void* wstream = stream;
      if (use_side_stream) {                 // wgrad (+ its finalize) beside the next layer's kernels
        B200_CUDA(cudaEventRecord(P.fork_ev[l], main_s));
        B200_CUDA(cudaStreamWaitEvent(P.side, P.fork_ev[l], 0));
        ws = P.wg_ws_side;
        wstream = P.side;
      }
      B200_RC(b200ddp_conv2d_wgrad(xin, P.dzc[l], dw[l], nullptr, ko, B, g.Cin, g.H, g.H, g.Cout, g.K, g.K, g.stride, g.pad,
                                   ws, P.wg_ws_bytes, dev, wstream));
      if (use_side_stream) B200_CUDA(cudaEventRecord(P.join_ev[l], P.side));
      g_in = P.dxc[l];
      if (l == 1) {                          // conv5..conv2 are final once their side-stream kernels have been joined
        if (use_side_stream)
          for (int k = 4; k >= 1; --k) B200_CUDA(cudaStreamWaitEvent(main_s, P.join_ev[k], 0));
        if (ready) ready(user, 2);
      }
    } else {
      B200_RC(b200ddp_conv2d_wgrad(xin, P.dzc[0], dw[0], d.conv1_ktab, d.conv1_order == 2 ? 2 : 0, B, g.Cin, g.H, g.H, g.Cout, g.K, g.K, g.stride,
                                   g.pad, P.wg_ws, P.wg_ws_bytes, dev, stream));
      if (ready) ready(user, 3);
    }
  }
  return 0;
}

}  // extern "C"
