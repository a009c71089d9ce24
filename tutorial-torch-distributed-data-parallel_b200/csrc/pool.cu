// C6 — activation-side elementwise kernels of the feature extractor and classifier
// (reference alexnet.py:23-24,26-27,29,31,33-34,38,41): ReLU fused with MaxPool2d(3,2) forward and
// backward (one read of the conv output instead of relu + pool = two), stand-alone ReLU for the two
// conv layers without a pool, and Philox dropout with a saved gate.
// All HBM-bound; NCHW fp32; one thread per output element, coalesced along W.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

constexpr int kPoolK = 3, kPoolS = 2;

// Index walker for "flat index over [planes][rows][cols], advanced by a fixed step": one division at set-up, then
// additions with carries (these kernels were instruction-bound on 64-bit divisions: ncu, issue slots 70 % busy).
struct Walk3 {
  int pl, r, c;            // current (plane, row, col)
  int dpl, dr, dc;         // the step, decomposed
  int rows, cols;
  __device__ __forceinline__ Walk3(int start, int step, int rows_, int cols_) : rows(rows_), cols(cols_) {
    const int per = rows_ * cols_;
    pl = start / per; int t = start - pl * per; r = t / cols_; c = t - r * cols_;
    dpl = step / per; t = step - dpl * per; dr = t / cols_; dc = t - dr * cols_;
  }
  __device__ __forceinline__ void next() {
    c += dc;
    const int k1 = c >= cols; c -= k1 ? cols : 0;
    r += dr + k1;
    const int k2 = r >= rows; r -= k2 ? rows : 0;
    pl += dpl + k2;
  }
};

// y[b,c,oh,ow] = max(0, max_{3x3 window} x); idx = window slot 0..8 of the first maximum, 255 when the
// maximum is <= 0 (ReLU closed: no gradient flows). One CTA per group of P whole planes, one thread per output.
__global__ void __launch_bounds__(256) relu_maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               unsigned char* __restrict__ idx, long long planes,
                                                               int H, int W, int OH, int OW, int relu, int P) {
  const long long pl0 = (long long)blockIdx.x * P;
  const int np = (int)(planes - pl0 < P ? planes - pl0 : P);
  const int ohw = OH * OW, hw = H * W, n = np * ohw;
  const float* xb = x + pl0 * hw;
  float* yb = y + pl0 * ohw;
  unsigned char* ib = idx ? idx + pl0 * ohw : nullptr;    // null in inference: the argmax codes are not needed
  Walk3 wk(threadIdx.x, blockDim.x, OH, OW);
  for (int i = threadIdx.x; i < n; i += blockDim.x, wk.next()) {
    const float* xp = xb + wk.pl * hw + (wk.r * kPoolS) * W + wk.c * kPoolS;
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
    yb[i] = best;
    if (ib) ib[i] = code;
  }
}

// dx[b,c,h,w] = sum over the (<= 4) windows that cover (h,w) and selected it of dy[window].
// Backward of ReLU + 3x3/2 max pool, optionally with the bias gradient of the convolution that fed it.
// A CTA stages the dy and argmax codes of P whole planes (<= kPoolChunk pooled elements, contiguous in NCHW) in
// shared memory and then produces the P input planes as one contiguous, coalesced stream: every input pixel
// gathers from the <= 2x2 windows that contain it (deterministic, no atomics). With `plane_sums` the CTA also
// writes sum(dx) of each plane - which is just the sum of the open dy entries - and bias_fold_kernel adds them over
// the batch in a fixed order into db[c] (fp64 accumulation): the five channel_sum passes over the conv outputs'
// gradients disappear from the step.
constexpr int kPoolChunk = 2048;

// db[c] = sum_b plane_sums[b][c] in fp64, fixed order: one warp per channel, lane l adds b = l, l+32, ... and the
// 32 partials are combined by a shuffle tree. A separate small launch: folding inside the producing kernel by its
// last CTA (ticket) was measured at 13-60 us of single-SM tail per layer, this costs 3-4 us.
__global__ void __launch_bounds__(256) bias_fold_kernel(const double* __restrict__ plane_sums, float* __restrict__ db,
                                                        int B, int C) {
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0;
  for (int bI = lane; bI < B; bI += 32) s += __ldcg(plane_sums + (long long)bI * C + c);
#pragma unroll
  for (int sft = 16; sft > 0; sft >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sft);
  if (lane == 0) db[c] = (float)s;
}

__global__ void __launch_bounds__(256) relu_maxpool_bwd_kernel(const float* __restrict__ dy,
                                                               const unsigned char* __restrict__ idx,
                                                               float* __restrict__ dx, double* __restrict__ plane_sums,
                                                               int B, int C, int H, int W, int OH, int OW, int P) {
  __shared__ float s_dy[kPoolChunk + 1];              // [n_in] = dummy entry (gradient 0, slot 255) for windows that
  __shared__ unsigned char s_idx[kPoolChunk + 4];     // do not exist: no predicated loads in the quad loop
  const int ohw = OH * OW, hw = H * W;
  const long long planes = (long long)B * C;
  const long long nchunks = (planes + P - 1) / P;
  const int QH = (H + 1) >> 1, QW = (W + 1) >> 1, qhw = QH * QW;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (long long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
    const long long pl0 = chunk * P;
    const int np = (int)(planes - pl0 < P ? planes - pl0 : P);
    const int n_in = np * ohw;
    const float* dyp = dy + pl0 * ohw;
    const unsigned char* ip = idx + pl0 * ohw;
    for (int i = threadIdx.x; i < n_in; i += blockDim.x) {
      s_dy[i] = __ldg(dyp + i);
      s_idx[i] = __ldg(ip + i);
    }
    if (threadIdx.x == 0) { s_dy[n_in] = 0.f; s_idx[n_in] = 255; }
    __syncthreads();
    if (plane_sums) {
      for (int pl = warp; pl < np; pl += nwarps) {
        float acc = 0.f;
        for (int o = lane; o < ohw; o += 32) acc += s_idx[pl * ohw + o] != 255 ? s_dy[pl * ohw + o] : 0.f;
        double d = (double)acc;
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) d += __shfl_xor_sync(0xffffffffu, d, sft);
        if (lane == 0) plane_sums[pl0 + pl] = d;
      }
    }
    // One thread per 2x2 input quad (rows 2a, 2a+1; columns 2b, 2b+1). Window (a,b) covers the whole quad (slots
    // 0,1,3,4), window (a,b-1) its left column (slots 2,5), window (a-1,b) its top row (slots 6,7) and window
    // (a-1,b-1) the top-left pixel (slot 8): four (dy, slot) pairs from shared memory, nine compares, four stores.
    float* dxp = dx + pl0 * hw;
    const int n_quads = np * qhw;
    Walk3 wk(threadIdx.x, blockDim.x, QH, QW);
    for (int i = threadIdx.x; i < n_quads; i += blockDim.x, wk.next()) {
      const int pl = wk.pl, a = wk.r, bq = wk.c;
      const bool va = a < OH, vb = bq < OW, va0 = a > 0, vb0 = bq > 0;
      const int base = pl * ohw + a * OW + bq;                         // window (a, bq) of this plane
      const int i11 = (va && vb) ? base : n_in, i10 = (va && vb0) ? base - 1 : n_in;
      const int i01 = (va0 && vb) ? base - OW : n_in, i00 = (va0 && vb0) ? base - OW - 1 : n_in;
      const float d11 = s_dy[i11], d10 = s_dy[i10], d01 = s_dy[i01], d00 = s_dy[i00];
      const int c11 = s_idx[i11], c10 = s_idx[i10], c01 = s_idx[i01], c00 = s_idx[i00];
      const float g00 = (c11 == 0 ? d11 : 0.f) + (c10 == 2 ? d10 : 0.f) + (c01 == 6 ? d01 : 0.f) + (c00 == 8 ? d00 : 0.f);
      const float g01 = (c11 == 1 ? d11 : 0.f) + (c01 == 7 ? d01 : 0.f);
      const float g10 = (c11 == 3 ? d11 : 0.f) + (c10 == 5 ? d10 : 0.f);
      const float g11 = (c11 == 4 ? d11 : 0.f);
      const int h = 2 * a, w = 2 * bq;
      float* o = dxp + pl * hw + h * W + w;
      const bool w1 = w + 1 < W, h1 = h + 1 < H;
      o[0] = g00;
      if (w1) o[1] = g01;
      if (h1) { o[W] = g10; if (w1) o[W + 1] = g11; }
    }
    __syncthreads();
  }
}

// dx = dy * (y > 0) for [B][C][HW] tensors plus the per-plane sums of dx, folded into db[c] by bias_fold_kernel (the
// bias gradient of a convolution whose ReLU is fused into its epilogue). One warp per plane, eight independent
// (y, dy) load pairs in flight per lane, no shared memory, no block barriers.
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                            float* __restrict__ dx, double* __restrict__ plane_sums,
                                                            long long planes, int HW) {
  const int lane = threadIdx.x & 31;
  const long long nw = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long pl = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); pl < planes; pl += nw) {
    const long long base = pl * HW;
    float acc = 0.f;
    for (int o0 = 0; o0 < HW; o0 += 256) {
      float yv[8], gv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int o = o0 + u * 32 + lane;
        yv[u] = o < HW ? __ldg(y + base + o) : 0.f;
        gv[u] = o < HW ? __ldg(dy + base + o) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int o = o0 + u * 32 + lane;
        const float v = yv[u] > 0.f ? gv[u] : 0.f;
        if (o < HW) dx[base + o] = v;
        acc += v;
      }
    }
    double d = (double)acc;
#pragma unroll
    for (int sft = 16; sft > 0; sft >>= 1) d += __shfl_xor_sync(0xffffffffu, d, sft);
    if (lane == 0) plane_sums[pl] = d;
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
  const long long cap = (long long)num_sms() * 16;
  return (int)(want < cap ? want : cap);
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200ddp_relu_maxpool_fwd(const float* x, float* y, unsigned char* idx, int B, int C, int H, int W, int relu,
                             int device, void* stream) {
  B200_CHECK_ARG(x && y && B > 0 && C > 0 && H >= 3 && W >= 3, "relu_maxpool_fwd: bad argument");
  B200_DEVICE(device);
  const int OH = (H - kPoolK) / kPoolS + 1, OW = (W - kPoolK) / kPoolS + 1;
  const long long planes = (long long)B * C;
  B200_CHECK_ARG((long long)H * W < (1 << 24) , "relu_maxpool_fwd: plane too large");
  int P = kPoolChunk / (OH * OW);
  if (P < 1) P = 1;
  while (P > 1 && (long long)P * H * W >= (1LL << 30)) --P;
  const long long nblocks = (planes + P - 1) / P;
  B200_CHECK_ARG(nblocks < (1LL << 31), "relu_maxpool_fwd: too many planes");
  relu_maxpool_fwd_kernel<<<(unsigned)nblocks, 256, 0, as_stream(stream)>>>(x, y, idx, planes, H, W, OH, OW, relu, P);
  B200_LAUNCH_OK();
  return 0;
}

static int pool_bwd_launch(const float* dy, const unsigned char* idx, float* dx, float* db, int B, int C, int H, int W,
                           void* scratch, long long scratch_bytes, int device, void* stream, const char* who) {
  B200_CHECK_ARG(dy && idx && dx && B > 0 && C > 0 && H >= 3 && W >= 3, "%s: bad argument", who);
  const int OH = (H - kPoolK) / kPoolS + 1, OW = (W - kPoolK) / kPoolS + 1;
  B200_CHECK_ARG(OH * OW <= kPoolChunk, "%s: pooled plane %dx%d exceeds %d elements", who, OH, OW, kPoolChunk);
  int P = kPoolChunk / (OH * OW);
  while (P > 1 && (long long)P * H * W >= (1 << 22)) --P;      // keeps the float-reciprocal index split exact
  B200_CHECK_ARG((long long)P * H * W < (1 << 23), "%s: plane %dx%d too large", who, H, W);
  double* plane_sums = nullptr;
  if (db) {
    B200_CHECK_ARG(scratch && ((uintptr_t)scratch & 7) == 0 && scratch_bytes >= (long long)B * C * 8,
                   "%s: scratch too small or unaligned (need %lld bytes, 8-byte aligned)", who, (long long)B * C * 8);
    plane_sums = reinterpret_cast<double*>(scratch);
  }
  B200_DEVICE(device);
  const long long planes = (long long)B * C, nchunks = (planes + P - 1) / P;
  const int grid = (int)(nchunks < (long long)num_sms() * 8 ? nchunks : (long long)num_sms() * 8);
  relu_maxpool_bwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(dy, idx, dx, plane_sums, B, C, H, W, OH, OW, P);
  B200_LAUNCH_OK();
  if (db) {
    bias_fold_kernel<<<(C + 7) / 8, 256, 0, as_stream(stream)>>>(plane_sums, db, B, C);
    B200_LAUNCH_OK();
  }
  return 0;
}

int b200ddp_relu_maxpool_bwd(const float* dy, const unsigned char* idx, float* dx, int B, int C, int H, int W,
                             int device, void* stream) {
  return pool_bwd_launch(dy, idx, dx, nullptr, B, C, H, W, nullptr, 0, device, stream, "relu_maxpool_bwd");
}

long long b200ddp_bias_grad_scratch_bytes(int B, int C) { return (long long)B * C * 8; }

int b200ddp_relu_maxpool_bwd_bias(const float* dy, const unsigned char* idx, float* dx, float* db, int B, int C, int H,
                                  int W, void* scratch, long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(db, "relu_maxpool_bwd_bias: db is null (use relu_maxpool_bwd)");
  return pool_bwd_launch(dy, idx, dx, db, B, C, H, W, scratch, scratch_bytes, device, stream, "relu_maxpool_bwd_bias");
}

int b200ddp_relu_bwd_bias(const float* dy, const float* y, float* dx, float* db, int B, int C, int HW, void* scratch,
                          long long scratch_bytes, int device, void* stream) {
  B200_CHECK_ARG(dy && y && dx && db && B > 0 && C > 0 && HW > 0, "relu_bwd_bias: bad argument");

  B200_CHECK_ARG(scratch && ((uintptr_t)scratch & 7) == 0 && scratch_bytes >= (long long)B * C * 8,
                 "relu_bwd_bias: scratch too small or unaligned (need %lld bytes, 8-byte aligned)", (long long)B * C * 8);
  B200_DEVICE(device);
  const long long planes = (long long)B * C;
  const long long want = (planes + 7) / 8;                 // 8 warps per CTA, one plane per warp and pass
  const int grid = (int)(want < (long long)num_sms() * 16 ? want : (long long)num_sms() * 16);
  double* plane_sums = reinterpret_cast<double*>(scratch);
  relu_bwd_bias_kernel<<<grid, 256, 0, as_stream(stream)>>>(dy, y, dx, plane_sums, planes, HW);
  B200_LAUNCH_OK();
  bias_fold_kernel<<<(C + 7) / 8, 256, 0, as_stream(stream)>>>(plane_sums, db, B, C);
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
