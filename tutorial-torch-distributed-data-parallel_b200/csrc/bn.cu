// B1 — SyncBatchNorm kernels (README.md:79-81 of the reference prescribes convert_sync_batchnorm).
//
// Restates the five ATen kernels torch/nn/modules/_functions.py:7-209 calls:
//   batch_norm_stats                      -> b200ddp_bn_stats        (local mean / invstd per channel)
//   batch_norm_gather_stats_with_counts   -> b200ddp_bn_gather_stats (merge all ranks + running stats)
//   batch_norm_elemt                      -> b200ddp_bn_elemt        (normalise + affine [+ ReLU])
//   batch_norm_backward_reduce            -> b200ddp_bn_bwd_reduce   (sum_dy, sum_dy_xmu, dweight, dbias)
//   batch_norm_backward_elemt             -> b200ddp_bn_bwd_elemt    (grad_input)
// The two collectives in between (all_gather of 2C+1 floats, all_reduce of 2C floats) are issued
// by the host layer through b200ddp_comm_*.
//
// Reductions: per-thread fp32 partial sums over a few dozen elements, warp-shuffle + block reduction
// in fp64, one (count, mean, M2) partial per CTA in a scratch buffer, and the CTA that takes the last
// ticket of a channel merges that channel's partials in index order (Chan et al.) -> deterministic.
// HBM-bound: grid = S x C CTAs with S chosen so that S*C covers >= 4 CTAs per SM.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

constexpr int kBnThreads = 256;
constexpr int kBnMaxSplit = 64;

static inline int bn_splits(int B, int C) {
  int s = (kNumSMs * 4 + C - 1) / C;
  if (s > B) s = B;
  if (s > kBnMaxSplit) s = kBnMaxSplit;
  return s < 1 ? 1 : s;
}

// Scratch layout: [tickets: kBnMaxChannels x u32][partials]. The ticket region has a FIXED size so that a
// scratch buffer shared by layers with different channel counts never sees one layer's partial sums where
// another layer expects zeroed tickets (tickets are reset to zero by the kernel that used them).
constexpr int kBnMaxChannels = 4096;
static inline long long bn_ticket_bytes(int) { return (long long)kBnMaxChannels * 4; }

__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const float* __restrict__ x, int B, int C, int HW, float eps, double* __restrict__ partials,
                unsigned int* __restrict__ tickets, float* __restrict__ mean, float* __restrict__ invstd) {
  __shared__ double red[32];
  __shared__ bool last;
  const int s = blockIdx.x, S = gridDim.x, c = blockIdx.y;
  float sum = 0.f, sq = 0.f;
  long long cnt = 0;
  for (int b = s; b < B; b += S) {
    const float* p = x + ((long long)b * C + c) * HW;
    for (int i = threadIdx.x; i < HW; i += kBnThreads) {
      const float v = __ldg(p + i);
      sum += v;
      sq = fmaf(v, v, sq);
    }
    cnt += HW;
  }
  const double dsum = block_sum<double>((double)sum, red);
  const double dsq = block_sum<double>((double)sq, red);
  if (threadIdx.x == 0) {
    const double n = (double)cnt;
    const double m = n > 0 ? dsum / n : 0.0;
    double* out = partials + ((long long)c * S + s) * 3;
    out[0] = n; out[1] = m; out[2] = n > 0 ? fmax(dsq - dsum * m, 0.0) : 0.0;
    __threadfence();
    last = (atomicAdd(&tickets[c], 1u) == (unsigned)S - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double n = 0, m = 0, M2 = 0;
    for (int j = 0; j < S; ++j) {
      const volatile double* in = partials + ((long long)c * S + j) * 3;
      const double nb = in[0], mb = in[1], M2b = in[2];
      if (nb == 0) continue;
      const double nn = n + nb, d = mb - m;
      m += d * nb / nn;
      M2 += M2b + d * d * n * nb / nn;
      n = nn;
    }
    mean[c] = (float)m;
    invstd[c] = (float)(1.0 / sqrt(M2 / n + (double)eps));
    tickets[c] = 0u;
  }
}

// all: [world][2C+1] = (mean[C], invstd[C], count). One thread per channel.
__global__ void bn_gather_stats_kernel(const float* __restrict__ all, int world, int C, float eps, float momentum,
                                       float* __restrict__ running_mean, float* __restrict__ running_var,
                                       float* __restrict__ mean, float* __restrict__ invstd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int stride = 2 * C + 1;
  double n = 0, avg = 0, var_n = 0;
  for (int r = 0; r < world; ++r) {
    const double m = all[(long long)r * stride + c];
    const double is = all[(long long)r * stride + C + c];
    const double nr = all[(long long)r * stride + 2 * C];
    if (nr == 0) continue;
    const double v = (1.0 / is) * (1.0 / is) - (double)eps;
    const double nn = n + nr, d = m - avg;
    avg += d * nr / nn;
    var_n += v * nr + d * d * n * nr / nn;
    n = nn;
  }
  mean[c] = (float)avg;
  invstd[c] = (float)(1.0 / sqrt(var_n / n + (double)eps));
  if (running_mean) running_mean[c] = (float)((1.0 - momentum) * running_mean[c] + momentum * avg);
  if (running_var) running_var[c] = (float)((1.0 - momentum) * running_var[c] + momentum * (var_n / (n - 1.0)));
}

// y = (x - mean[c]) * invstd[c] * w[c] + b[c], optional ReLU
__global__ void __launch_bounds__(256)
bn_elemt_kernel(const float* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ invstd,
                const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ y, long long planes,
                int C, int HW, int relu) {
  // one CTA walks whole (b,c) planes so the per-channel scalars are loaded once per plane
  for (long long pl = blockIdx.x; pl < planes; pl += gridDim.x) {
    const int c = (int)(pl % C);
    const float m = __ldg(mean + c);
    const float a = __ldg(invstd + c) * (w ? __ldg(w + c) : 1.f);
    const float bb = bias ? __ldg(bias + c) : 0.f;
    const float* xp = x + pl * HW;
    float* yp = y + pl * HW;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) {
      float v = fmaf(xp[i] - m, a, bb);
      if (relu) v = v > 0.f ? v : 0.f;
      yp[i] = v;
    }
  }
}

__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y_relu,
                     const float* __restrict__ mean, const float* __restrict__ invstd, int B, int C, int HW,
                     double* __restrict__ partials, unsigned int* __restrict__ tickets, float* __restrict__ sum_dy,
                     float* __restrict__ sum_dy_xmu, float* __restrict__ grad_weight, float* __restrict__ grad_bias) {
  __shared__ double red[32];
  __shared__ bool last;
  const int s = blockIdx.x, S = gridDim.x, c = blockIdx.y;
  const float m = __ldg(mean + c);
  float a = 0.f, bsum = 0.f;
  for (int b = s; b < B; b += S) {
    const long long off = ((long long)b * C + c) * HW;
    for (int i = threadIdx.x; i < HW; i += kBnThreads) {
      float g = __ldg(dy + off + i);
      if (y_relu && !(__ldg(y_relu + off + i) > 0.f)) g = 0.f;
      a += g;
      bsum = fmaf(g, __ldg(x + off + i) - m, bsum);
    }
  }
  const double da = block_sum<double>((double)a, red);
  const double db = block_sum<double>((double)bsum, red);
  if (threadIdx.x == 0) {
    double* out = partials + ((long long)c * S + s) * 2;
    out[0] = da; out[1] = db;
    __threadfence();
    last = (atomicAdd(&tickets[c], 1u) == (unsigned)S - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double ta = 0, tb = 0;
    for (int j = 0; j < S; ++j) {
      const volatile double* in = partials + ((long long)c * S + j) * 2;
      ta += in[0]; tb += in[1];
    }
    sum_dy[c] = (float)ta;
    sum_dy_xmu[c] = (float)tb;
    if (grad_weight) grad_weight[c] = (float)(tb * (double)invstd[c]);
    if (grad_bias) grad_bias[c] = (float)ta;
    tickets[c] = 0u;
  }
}

__global__ void __launch_bounds__(256)
bn_bwd_elemt_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y_relu,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ w,
                    const float* __restrict__ sum_dy, const float* __restrict__ sum_dy_xmu, float inv_count,
                    float* __restrict__ dx, long long planes, int C, int HW) {
  for (long long pl = blockIdx.x; pl < planes; pl += gridDim.x) {
    const int c = (int)(pl % C);
    const float m = __ldg(mean + c), is = __ldg(invstd + c);
    const float m_dy = __ldg(sum_dy + c) * inv_count;
    const float f1 = is * is * __ldg(sum_dy_xmu + c) * inv_count;
    const float f2 = (w ? __ldg(w + c) : 1.f) * is;
    const long long off = pl * HW;
    for (int i = threadIdx.x; i < HW; i += blockDim.x) {
      float g = dy[off + i];
      if (y_relu && !(y_relu[off + i] > 0.f)) g = 0.f;
      dx[off + i] = (g - m_dy - (x[off + i] - m) * f1) * f2;
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

long long b200ddp_bn_scratch_bytes(int B, int C) {
  return bn_ticket_bytes(C) + (long long)C * bn_splits(B, C) * 3 * 8;
}

int b200ddp_bn_stats(const float* x, int B, int C, int HW, float eps, float* mean, float* invstd, void* scratch,
                     long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(x && mean && invstd && scratch && B > 0 && C > 0 && HW > 0, "bn_stats: bad argument");
  B200_CHECK_ARG(C <= kBnMaxChannels, "bn_stats: at most %d channels", kBnMaxChannels);
  B200_CHECK_ARG(scratch_bytes >= b200ddp_bn_scratch_bytes(B, C), "bn_stats: scratch too small");
  B200_CHECK_ARG(((uintptr_t)scratch & 255) == 0, "bn_stats: scratch must be 256-byte aligned");
  B200_DEVICE(device);
  const int S = bn_splits(B, C);
  unsigned int* tickets = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + bn_ticket_bytes(C));
  bn_stats_kernel<<<dim3(S, C), kBnThreads, 0, as_stream(stream)>>>(x, B, C, HW, eps, partials, tickets, mean, invstd);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_bn_gather_stats(const float* all_stats, int world, int C, float eps, float momentum, float* running_mean,
                            float* running_var, float* mean, float* invstd, int device, void* stream) {
  B200_CHECK_ARG(all_stats && mean && invstd && world > 0 && C > 0, "bn_gather_stats: bad argument");
  B200_DEVICE(device);
  bn_gather_stats_kernel<<<ceil_div(C, 128), 128, 0, as_stream(stream)>>>(all_stats, world, C, eps, momentum,
                                                                          running_mean, running_var, mean, invstd);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_bn_elemt(const float* x, const float* mean, const float* invstd, const float* weight, const float* bias,
                     float* y, int B, int C, int HW, int relu, int device, void* stream) {
  B200_CHECK_ARG(x && mean && invstd && y && B > 0 && C > 0 && HW > 0, "bn_elemt: bad argument");
  B200_DEVICE(device);
  const long long planes = (long long)B * C;
  const int grid = (int)(planes < num_sms() * 8 ? planes : num_sms() * 8);
  bn_elemt_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, mean, invstd, weight, bias, y, planes, C, HW, relu);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_bn_bwd_reduce(const float* dy, const float* x, const float* y_relu, const float* mean, const float* invstd,
                          int B, int C, int HW, float* sum_dy, float* sum_dy_xmu, float* grad_weight, float* grad_bias,
                          void* scratch, long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(dy && x && mean && invstd && sum_dy && sum_dy_xmu && scratch && B > 0 && C > 0 && HW > 0,
                 "bn_bwd_reduce: bad argument");
  B200_CHECK_ARG(C <= kBnMaxChannels, "bn_bwd_reduce: at most %d channels", kBnMaxChannels);
  B200_CHECK_ARG(scratch_bytes >= b200ddp_bn_scratch_bytes(B, C), "bn_bwd_reduce: scratch too small");
  B200_CHECK_ARG(((uintptr_t)scratch & 255) == 0, "bn_bwd_reduce: scratch must be 256-byte aligned");
  B200_DEVICE(device);
  const int S = bn_splits(B, C);
  unsigned int* tickets = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + bn_ticket_bytes(C));
  bn_bwd_reduce_kernel<<<dim3(S, C), kBnThreads, 0, as_stream(stream)>>>(dy, x, y_relu, mean, invstd, B, C, HW, partials,
                                                                        tickets, sum_dy, sum_dy_xmu, grad_weight,
                                                                        grad_bias);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_bn_bwd_elemt(const float* dy, const float* x, const float* y_relu, const float* mean, const float* invstd,
                         const float* weight, const float* sum_dy, const float* sum_dy_xmu, float inv_count, float* dx,
                         int B, int C, int HW, int device, void* stream) {
  B200_CHECK_ARG(dy && x && mean && invstd && sum_dy && sum_dy_xmu && dx && B > 0 && C > 0 && HW > 0,
                 "bn_bwd_elemt: bad argument");
  B200_DEVICE(device);
  const long long planes = (long long)B * C;
  const int grid = (int)(planes < num_sms() * 8 ? planes : num_sms() * 8);
  bn_bwd_elemt_kernel<<<grid, 256, 0, as_stream(stream)>>>(dy, x, y_relu, mean, invstd, weight, sum_dy, sum_dy_xmu,
                                                           inv_count, dx, planes, C, HW);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
