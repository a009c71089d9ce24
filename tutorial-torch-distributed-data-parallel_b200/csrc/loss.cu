// F3 (loss half) — softmax cross-entropy, forward and gradient in ONE launch.
//
// Restates nn.CrossEntropyLoss() (reference multi-GPU-training-torch.py:248,122): mean over the batch
// of -log_softmax(z)[y]; gradient (softmax(z) - onehot(y)) / B. Upstream runs log_softmax + nll_loss
// as 2 forward + 2 backward kernels; here one block does both and keeps the
// gradient for backward. Latency-bound ([B x 10] floats): a single CTA, one thread per row,
// deterministic block reduction for the mean.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

constexpr long long kIgnoreIndex = -100;   // nn.CrossEntropyLoss(ignore_index=-100) default

// Rows whose label equals ignore_index contribute nothing and are left out of the mean (torch semantics);
// any other label outside [0, C) is an error: the row is treated as ignored (zero gradient row, never
// uninitialised memory) and *err_flag is raised for the host to report.
__global__ void __launch_bounds__(1024) softmax_ce_kernel(const float* __restrict__ logits,
                                                          const long long* __restrict__ labels,
                                                          float* __restrict__ loss, float* __restrict__ dlogits,
                                                          int B, int C, int* __restrict__ err_flag) {
  __shared__ float red[32];
  float cnt = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const long long y = labels[b];
    if (y >= 0 && y < C) cnt += 1.f;
    else if (y != kIgnoreIndex && err_flag) atomicExch(err_flag, 1);
  }
  const float nvalid = block_sum<float>(cnt, red);          // exact: B < 2^24
  const float inv = nvalid > 0.f ? 1.0f / nvalid : 0.f;
  float local = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const float* z = logits + (long long)b * C;
    const long long y = labels[b];
    if (y < 0 || y >= C) {
      if (dlogits)
        for (int c = 0; c < C; ++c) dlogits[(long long)b * C + c] = 0.f;
      continue;
    }
    float mx = -INFINITY;
    for (int c = 0; c < C; ++c) mx = fmaxf(mx, z[c]);
    float se = 0.f;
    for (int c = 0; c < C; ++c) se += expf(z[c] - mx);
    const float lse = logf(se);
    local += -(z[y] - mx - lse);
    if (dlogits) {
      float* d = dlogits + (long long)b * C;
      for (int c = 0; c < C; ++c) {
        const float sm = expf(z[c] - mx - lse);
        d[c] = (sm - (c == (int)y ? 1.f : 0.f)) * inv;
      }
    }
  }
  const float tot = block_sum<float>(local, red);
  if (threadIdx.x == 0) loss[0] = nvalid > 0.f ? tot * inv : __int_as_float(0x7fc00000);   // all ignored: nan, as torch
}

}  // namespace b200

extern "C" int b200ddp_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits, int B,
                                  int C, int* err_flag_dev, int device, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(logits && labels && loss && B > 0 && C > 0, "softmax_ce: bad argument");
  B200_DEVICE(device);
  int threads = ((B + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  softmax_ce_kernel<<<1, threads, 0, as_stream(stream)>>>(logits, labels, loss, dlogits, B, C, err_flag_dev);
  B200_LAUNCH_OK();
  return 0;
}
