// C6 — activation-side elementwise kernels of the feature extractor and classifier
// (reference alexnet.py:23-24,26-27,29,31,33-34,38,41): ReLU fused with MaxPool2d(3,2) forward and
// backward (one read of the conv output instead of relu + pool = two), stand-alone ReLU for the two
// conv layers without a pool, and Philox dropout with a saved gate.
// All HBM-bound; NCHW fp32; one thread per output element, coalesced along W.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

constexpr int kPoolK = 3, kPoolS = 2;

// y[b,c,oh,ow] = max(0, max_{3x3 window} x); idx = window slot 0..8 of the first maximum, 255 when the
// maximum is <= 0 (ReLU closed: no gradient flows).
__global__ void __launch_bounds__(256) relu_maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               unsigned char* __restrict__ idx, long long planes,
                                                               int H, int W, int OH, int OW, int relu) {
  const long long total = planes * OH * OW, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int ow = (int)(i % OW);
    const long long t = i / OW;
    const int oh = (int)(t % OH);
    const long long pl = t / OH;
    const float* xp = x + pl * (long long)H * W + (long long)(oh * kPoolS) * W + ow * kPoolS;
    float best = -INFINITY;
    int bi = 0;
#pragma unroll
    for (int kh = 0; kh < kPoolK; ++kh)
#pragma unroll
      for (int kw = 0; kw < kPoolK; ++kw) {
        const float v = __ldg(xp + kh * W + kw);
        if (v > best || v != v) {   // torch semantics: NaN wins, otherwise first strict maximum
          best = v;
          bi = kh * kPoolK + kw;
        }
      }
    unsigned char code = (unsigned char)bi;
    if (relu && !(best > 0.f)) { best = 0.f; code = 255; }
    y[i] = best;
    idx[i] = code;
  }
}

// dx[b,c,h,w] = sum over the (<= 4) windows that cover (h,w) and selected it of dy[window].
// One CTA per (b,c) plane: the plane's dy and argmax codes (<= 27x27) are staged in shared memory, then
// every input pixel gathers from them (deterministic, no atomics). 32-bit index math only.
constexpr int kPoolMaxOut = 27 * 27;
__global__ void __launch_bounds__(256) relu_maxpool_bwd_kernel(const float* __restrict__ dy,
                                                               const unsigned char* __restrict__ idx,
                                                               float* __restrict__ dx, long long planes, int H, int W,
                                                               int OH, int OW) {
  __shared__ float s_dy[kPoolMaxOut];
  __shared__ unsigned char s_idx[kPoolMaxOut];
  const int ohw = OH * OW, hw = H * W;
  for (long long pl = blockIdx.x; pl < planes; pl += gridDim.x) {
    const float* dyp = dy + pl * ohw;
    const unsigned char* ip = idx + pl * ohw;
    for (int i = threadIdx.x; i < ohw; i += blockDim.x) {
      s_dy[i] = __ldg(dyp + i);
      s_idx[i] = __ldg(ip + i);
    }
    __syncthreads();
    float* dxp = dx + pl * hw;
    // warp <-> input row, lane <-> column: no integer division, straight-line window tests. A pixel (h,w) lies
    // in window rows {h>>1} and, when h is even, {(h>>1)-1} (same for columns): at most 2x2 candidates.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int h = warp; h < H; h += nwarps) {
      const int oh1 = h >> 1, oh0 = oh1 - 1, hodd = h & 1;
      const bool vh1 = oh1 < OH, vh0 = !hodd && oh0 >= 0;
      for (int w = lane; w < W; w += 32) {
        const int ow1 = w >> 1, ow0 = ow1 - 1, wodd = w & 1;
        const bool vw1 = ow1 < OW, vw0 = !wodd && ow0 >= 0;
        float g = 0.f;
        if (vh1 && vw1) { const int o = oh1 * OW + ow1; if ((int)s_idx[o] == hodd * 3 + wodd) g += s_dy[o]; }
        if (vh1 && vw0) { const int o = oh1 * OW + ow0; if ((int)s_idx[o] == hodd * 3 + 2) g += s_dy[o]; }
        if (vh0 && vw1) { const int o = oh0 * OW + ow1; if ((int)s_idx[o] == 6 + wodd) g += s_dy[o]; }
        if (vh0 && vw0) { const int o = oh0 * OW + ow0; if ((int)s_idx[o] == 8) g += s_dy[o]; }
        dxp[h * W + w] = g;
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) relu_fwd_kernel(float* __restrict__ x, long long n) {
  const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
  const bool vec = ((uintptr_t)x & 15) == 0;
  if (vec) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      float4 v = reinterpret_cast<float4*>(x)[i];
      v.x = v.x > 0.f ? v.x : 0.f; v.y = v.y > 0.f ? v.y : 0.f;
      v.z = v.z > 0.f ? v.z : 0.f; v.w = v.w > 0.f ? v.w : 0.f;
      reinterpret_cast<float4*>(x)[i] = v;
    }
    const long long t = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) x[t] = x[t] > 0.f ? x[t] : 0.f;
  } else {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
      x[i] = x[i] > 0.f ? x[i] : 0.f;
  }
}

// dx = dy * (y > 0); dx may alias dy
__global__ void __launch_bounds__(256) relu_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                       float* __restrict__ dx, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}

// y = keep ? x/(1-p) : 0 ; gate = keep ? 2 : 0 (same code as the GEMM epilogue gate)
__global__ void __launch_bounds__(256) dropout_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                          unsigned char* __restrict__ gate, long long n,
                                                          uint32_t p_threshold, float scale, unsigned long long seed) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const bool keep = dropout_keep(seed, (uint64_t)i, p_threshold);
    y[i] = keep ? x[i] * scale : 0.f;
    gate[i] = keep ? 2 : 0;
  }
}

static inline int ew_grid(long long n) {
  long long want = (n + 255) / 256;
  if (want < 1) want = 1;
  const long long cap = (long long)kNumSMs * 16;
  return (int)(want < cap ? want : cap);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200ddp_relu_maxpool_fwd(const float* x, float* y, unsigned char* idx, int B, int C, int H, int W, int relu,
                             int device, void* stream) {
  B200_CHECK_ARG(x && y && idx && B > 0 && C > 0 && H >= 3 && W >= 3, "relu_maxpool_fwd: bad argument");
  B200_DEVICE(device);
  const int OH = (H - kPoolK) / kPoolS + 1, OW = (W - kPoolK) / kPoolS + 1;
  const long long planes = (long long)B * C;
  relu_maxpool_fwd_kernel<<<ew_grid(planes * OH * OW), 256, 0, as_stream(stream)>>>(x, y, idx, planes, H, W, OH, OW, relu);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_relu_maxpool_bwd(const float* dy, const unsigned char* idx, float* dx, int B, int C, int H, int W,
                             int device, void* stream) {
  B200_CHECK_ARG(dy && idx && dx && B > 0 && C > 0 && H >= 3 && W >= 3, "relu_maxpool_bwd: bad argument");
  B200_DEVICE(device);
  const int OH = (H - kPoolK) / kPoolS + 1, OW = (W - kPoolK) / kPoolS + 1;
  const long long planes = (long long)B * C;
  B200_CHECK_ARG(OH * OW <= kPoolMaxOut, "relu_maxpool_bwd: pooled plane %dx%d exceeds %d elements", OH, OW, kPoolMaxOut);
  const int grid = (int)(planes < (long long)kNumSMs * 16 ? planes : (long long)kNumSMs * 16);
  relu_maxpool_bwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(dy, idx, dx, planes, H, W, OH, OW);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_relu_fwd(float* x, long long n, int device, void* stream) {
  B200_CHECK_ARG(x && n >= 0, "relu_fwd: bad argument");
  if (n == 0) return 0;
  B200_DEVICE(device);
  relu_fwd_kernel<<<ew_grid((n + 3) / 4), 256, 0, as_stream(stream)>>>(x, n);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_relu_bwd(const float* dy, const float* y, float* dx, long long n, int device, void* stream) {
  B200_CHECK_ARG(dy && y && dx && n >= 0, "relu_bwd: bad argument");
  if (n == 0) return 0;
  B200_DEVICE(device);
  relu_bwd_kernel<<<ew_grid(n), 256, 0, as_stream(stream)>>>(dy, y, dx, n);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_dropout_fwd(const float* x, float* y, unsigned char* gate, long long n, float p,
                        unsigned long long seed, int device, void* stream) {
  B200_CHECK_ARG(x && y && gate && n >= 0 && p >= 0.f && p < 1.f, "dropout_fwd: bad argument");
  if (n == 0) return 0;
  B200_DEVICE(device);
  double thr = (double)p * 4294967296.0;
  const uint32_t pt = thr >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)thr;
  dropout_fwd_kernel<<<ew_grid(n), 256, 0, as_stream(stream)>>>(x, y, gate, n, pt, 1.0f / (1.0f - p), seed);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
