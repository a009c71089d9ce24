// f1 — the evaluate() path: head + loss + accuracy in one launch, metrics accumulated on the device.
//
// Reference evaluate() (multi-GPU-training-torch.py:136-153, multi-GPU-training-accelerate.py:60-75) runs, per test
// batch: model(inputs) -> criterion -> loss.item() (sync) -> torch.max(outputs, 1) -> (predicted == labels).sum().item()
// (second sync) and adds into three 1-element tensors. Here the 10-class head (classifier.6), the softmax
// cross-entropy, the argmax and the comparison with the labels are ONE kernel that adds
//     metrics[0] += sum_i CE_i      metrics[1] += #(argmax_i == y_i)      metrics[2] += B
// into a 3-float device buffer: no host synchronisation per batch, and at the end of the epoch one 3-float
// all-reduce replaces the reference's separate scalar all-reduces (:198-204).
//
// Semantics kept from torch: loss per row = logsumexp(z) - z[y] (the reference adds mean * batch_size, i.e. the
// same sum up to rounding); argmax = index of the FIRST maximal logit, NaN wins (torch.max); a label outside
// [0, C) gives zero loss and is never "correct" but still counts in the total. Deterministic: per-row results go to
// scratch, the last CTA to finish adds them up in row order.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {
namespace {

constexpr int kEvalMaxN = 16;

__device__ __forceinline__ void row_metrics(const float* z, int C, long long y, float& loss, float& correct) {
  float mx = -INFINITY;
  int arg = 0;
  bool nan_seen = false;
  for (int c = 0; c < C; ++c) {
    const float v = z[c];
    if (!nan_seen && (v > mx || v != v)) { mx = v; arg = c; nan_seen = (v != v); }
  }
  loss = 0.f;
  if (y >= 0 && y < C) {
    float m2 = -INFINITY;
    for (int c = 0; c < C; ++c) m2 = fmaxf(m2, z[c]);
    float se = 0.f;
    for (int c = 0; c < C; ++c) se += expf(z[c] - m2);
    loss = -(z[y] - m2 - logf(se));
  }
  correct = ((long long)arg == y) ? 1.f : 0.f;
}

// last-ticket merge: rows[2*b] = loss_b, rows[2*b+1] = correct_b; metrics += (sum loss, sum correct, B)
__device__ __forceinline__ void merge_rows(const float* rows, int B, float* metrics) {
  float l = 0.f, c = 0.f;
  for (int b = 0; b < B; ++b) {
    l += ((volatile const float*)rows)[2 * b];
    c += ((volatile const float*)rows)[2 * b + 1];
  }
  metrics[0] += l;
  metrics[1] += c;
  metrics[2] += (float)B;
}

// One CTA (8 warps) per batch row: logits of the row (N <= 16 dot products over K), then its loss / hit.
__global__ void __launch_bounds__(256) head_eval_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                                        const float* __restrict__ bias, const long long* __restrict__ labels,
                                                        float* __restrict__ logits_out, int B, int N, int K,
                                                        float* __restrict__ rows, unsigned int* __restrict__ ticket,
                                                        float* __restrict__ metrics) {
  __shared__ float part[8][kEvalMaxN];
  __shared__ float z[kEvalMaxN];
  __shared__ bool last;
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc[kEvalMaxN];
#pragma unroll
  for (int n = 0; n < kEvalMaxN; ++n) acc[n] = 0.f;
  const float* xr = h + (long long)b * K;
  if ((K & 3) == 0 && (((uintptr_t)h | (uintptr_t)w) & 15) == 0) {
    for (int k = (warp * 32 + lane) * 4; k < K; k += 256 * 4) {
      const float4 xv = *reinterpret_cast<const float4*>(xr + k);
#pragma unroll
      for (int n = 0; n < kEvalMaxN; ++n)
        if (n < N) {
          const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (long long)n * K + k));
          acc[n] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[n]))));
        }
    }
  } else {
    for (int k = warp * 32 + lane; k < K; k += 256) {
      const float xv = xr[k];
#pragma unroll
      for (int n = 0; n < kEvalMaxN; ++n)
        if (n < N) acc[n] = fmaf(xv, __ldg(w + (long long)n * K + k), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < kEvalMaxN; ++n)
    if (n < N) {
      const float s = warp_sum(acc[n]);
      if (lane == 0) part[warp][n] = s;
    }
  __syncthreads();
  if (threadIdx.x < N) {
    float s = bias ? bias[threadIdx.x] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += part[j][threadIdx.x];
    z[threadIdx.x] = s;
    if (logits_out) logits_out[(long long)b * N + threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float loss, correct;
    row_metrics(z, N, labels[b], loss, correct);
    rows[2 * b] = loss;
    rows[2 * b + 1] = correct;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == (unsigned)B - 1);
    if (last) {
      __threadfence();
      merge_rows(rows, B, metrics);
      *ticket = 0u;
    }
  }
}

// Same metrics from logits that already exist (a model whose head is not a b200ddp Linear): one CTA.
__global__ void __launch_bounds__(1024) logits_eval_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                           int B, int C, float* __restrict__ rows, float* __restrict__ metrics) {
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float loss, correct;
    row_metrics(logits + (long long)b * C, C, labels[b], loss, correct);
    rows[2 * b] = loss;
    rows[2 * b + 1] = correct;
  }
  __syncthreads();
  if (threadIdx.x == 0) merge_rows(rows, B, metrics);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

long long b200ddp_eval_scratch_bytes(int B) { return 256 + 8ll * (B > 0 ? B : 0); }

int b200ddp_head_eval(const float* h, const float* w, const float* bias, const long long* labels, float* logits_out, int B,
                      int N, int K, float* metrics, void* scratch, long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(h && w && labels && metrics && scratch && B > 0 && K > 0, "head_eval: bad argument");
  B200_CHECK_ARG(N >= 1 && N <= kEvalMaxN, "head_eval: 1..%d classes (got %d)", kEvalMaxN, N);
  B200_CHECK_ARG(scratch_bytes >= b200ddp_eval_scratch_bytes(B) && ((uintptr_t)scratch & 255) == 0,
                 "head_eval: scratch too small or not 256-byte aligned");
  B200_DEVICE(device);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);      // zero on first use, reset by the kernel
  float* rows = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  head_eval_kernel<<<B, 256, 0, as_stream(stream)>>>(h, w, bias, labels, logits_out, B, N, K, rows, ticket, metrics);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_logits_eval(const float* logits, const long long* labels, int B, int C, float* metrics, void* scratch,
                        long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(logits && labels && metrics && scratch && B > 0 && C > 0, "logits_eval: bad argument");
  B200_CHECK_ARG(scratch_bytes >= b200ddp_eval_scratch_bytes(B) && ((uintptr_t)scratch & 255) == 0,
                 "logits_eval: scratch too small or not 256-byte aligned");
  B200_DEVICE(device);
  float* rows = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  int threads = ((B + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  logits_eval_kernel<<<1, threads, 0, as_stream(stream)>>>(logits, labels, B, C, rows, metrics);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
