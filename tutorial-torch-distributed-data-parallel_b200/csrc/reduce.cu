// R2 / R3 — gradient pack, scale, clip-before-aggregation and squared-norm kernels.
//
// R2 restates c10d::Reducer::mark_variable_ready_dense (bucket_view = grad * 1/ws,
// torch/csrc/distributed/c10d/reducer.hpp:276 + default_comm_hooks.hpp:36-40).
// R3 restates torch/nn/utils/clip_grad.py:165-169 (norm) and :291-299 (value), applied to
// the LOCAL gradients before the all-reduce as reference README.md:92-95 prescribes.
// All kernels are HBM-bound streaming passes: float4, grid = k x 148 CTAs.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

constexpr int kRedThreads = 256;
constexpr int kRedMaxBlocks = kNumSMs * 4;  // 592 partials at most

static inline int stream_grid(long long n) {
  long long want = ((n >> 2) + kRedThreads - 1) / kRedThreads;
  if (want < 1) want = 1;
  const long long cap = (long long)num_sms() * 8;
  return (int)(want < cap ? want : cap);
}

__device__ __forceinline__ float clamp_keep_nan(float g, float c, int nan_to_zero) {
  if (g != g) return nan_to_zero ? 0.f : g;
  return fminf(fmaxf(g, -c), c);
}

// dst = clamp(src, +-clip) * scale   (clip <= 0: no clamp). dst may equal src.
__global__ void __launch_bounds__(kRedThreads)
pack_scale_clip_kernel(float* __restrict__ dst, const float* __restrict__ src, long long n, float scale,
                       float clip, int nan_to_zero) {
  const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
  const bool vec = ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0);
  if (vec) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      float4 v = reinterpret_cast<const float4*>(src)[i];
      if (clip > 0.f) {
        v.x = clamp_keep_nan(v.x, clip, nan_to_zero);
        v.y = clamp_keep_nan(v.y, clip, nan_to_zero);
        v.z = clamp_keep_nan(v.z, clip, nan_to_zero);
        v.w = clamp_keep_nan(v.w, clip, nan_to_zero);
      }
      v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
      reinterpret_cast<float4*>(dst)[i] = v;
    }
    const long long t = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) {
      float g = src[t];
      if (clip > 0.f) g = clamp_keep_nan(g, clip, nan_to_zero);
      dst[t] = g * scale;
    }
  } else {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      float g = src[i];
      if (clip > 0.f) g = clamp_keep_nan(g, clip, nan_to_zero);
      dst[i] = g * scale;
    }
  }
}

// g *= *coef, skipping all traffic when *coef == 1. With exact_zero, *coef == 0 writes exact
// zeros (so a rank whose gradient is non-finite contributes nothing instead of NaN*0 = NaN).
__global__ void __launch_bounds__(kRedThreads)
scale_by_dev_kernel(float* __restrict__ g, long long n, const float* __restrict__ coef, int exact_zero) {
  float c = __ldg(coef);
  if (c == 1.0f) return;
  const bool zero = exact_zero && (c == 0.f);
  const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 v = reinterpret_cast<float4*>(g)[i];
    if (zero) { v = make_float4(0.f, 0.f, 0.f, 0.f); }
    else { v.x *= c; v.y *= c; v.z *= c; v.w *= c; }
    reinterpret_cast<float4*>(g)[i] = v;
  }
  const long long t = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) g[t] = zero ? 0.f : g[t] * c;
}

// Sum of squares of g[0..n). Each CTA writes one double partial into `partials`; the CTA that
// takes the last ticket adds them up in index order (deterministic) and stores / accumulates
// into out_sq. `ticket` must be zero on entry and is reset to zero on exit.
__global__ void __launch_bounds__(kRedThreads)
sqnorm_kernel(const float* __restrict__ g, long long n, double* __restrict__ partials,
              unsigned int* __restrict__ ticket, float* __restrict__ out_sq, int accumulate) {
  __shared__ double red[32];
  __shared__ bool last;
  const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
  // fp64 accumulation: the pass is HBM-bound, and squares of large (but finite) gradients must not
  // overflow before the clip coefficient can tame them
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = ld_stream(reinterpret_cast<const float4*>(g) + i);
    a0 = fma((double)v.x, (double)v.x, a0); a1 = fma((double)v.y, (double)v.y, a1);
    a2 = fma((double)v.z, (double)v.z, a2); a3 = fma((double)v.w, (double)v.w, a3);
  }
  const long long t = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) a0 = fma((double)g[t], (double)g[t], a0);
  double acc = a0 + a1 + a2 + a3;
  acc = block_sum<double>(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = acc;
    __threadfence();
    const unsigned int tk = atomicAdd(ticket, 1u);
    last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += ((volatile double*)partials)[i];
    // fixed-shape tree: per-thread strided sums then block_sum -> run-to-run deterministic
    s = block_sum<double>(s, red);
    if (threadIdx.x == 0) {
      const float r = (float)s;
      out_sq[0] = accumulate ? out_sq[0] + r : r;
      *ticket = 0u;
    }
  }
}

// coef = min(1, max_norm / (sqrt(sq) + 1e-6)); non-finite norm: policy 0 -> propagate (stock
// torch: coef = NaN), policy 1 -> coef = 0 (drop this rank's contribution).
__global__ void clip_coef_kernel(const float* __restrict__ sq, float max_norm, int nonfinite_policy,
                                 float* __restrict__ coef_out, float* __restrict__ norm_out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const float norm = sqrtf(sq[0]);
    float coef = max_norm / (norm + 1e-6f);
    const bool finite = (norm == norm) && !isinf(norm);
    if (!finite && nonfinite_policy) coef = 0.f;
    else if (coef == coef) coef = fminf(coef, 1.0f);   // NaN stays NaN like torch.clamp
    coef_out[0] = coef;
    if (norm_out) norm_out[0] = norm;
  }
}

}  // namespace b200

extern "C" {

int b200ddp_pack_scale_clip(float* dst, const float* src, long long n, float scale, float clip_value,
                            int nan_to_zero, int device, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(n >= 0, "pack_scale_clip: n < 0");
  if (n == 0) return 0;
  B200_CHECK_ARG(dst && src, "pack_scale_clip: null buffer");
  B200_DEVICE(device);
  pack_scale_clip_kernel<<<stream_grid(n), kRedThreads, 0, as_stream(stream)>>>(dst, src, n, scale,
                                                                                 clip_value, nan_to_zero);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_scale_by_device_scalar(float* g, long long n, const float* coef_dev, int exact_zero, int device,
                                   void* stream) {
  using namespace b200;
  B200_CHECK_ARG(n >= 0, "scale_by_device_scalar: n < 0");
  if (n == 0) return 0;
  B200_CHECK_ARG(g && coef_dev, "scale_by_device_scalar: null buffer");
  B200_CHECK_ARG(((uintptr_t)g & 15) == 0, "scale_by_device_scalar: g must be 16-byte aligned");
  B200_DEVICE(device);
  scale_by_dev_kernel<<<stream_grid(n), kRedThreads, 0, as_stream(stream)>>>(g, n, coef_dev, exact_zero);
  B200_LAUNCH_OK();
  return 0;
}

long long b200ddp_sqnorm_scratch_bytes(void) { return 256 + 8ll * b200::kRedMaxBlocks; }

int b200ddp_sqnorm(const float* g, long long n, float* out_sq, int accumulate, void* scratch,
                   long long scratch_bytes, int device, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(n >= 0, "sqnorm: n < 0");
  B200_CHECK_ARG(g && out_sq && scratch, "sqnorm: null buffer");
  B200_CHECK_ARG(scratch_bytes >= b200ddp_sqnorm_scratch_bytes(), "sqnorm: scratch too small (%lld < %lld)",
                 scratch_bytes, b200ddp_sqnorm_scratch_bytes());
  B200_CHECK_ARG(((uintptr_t)g & 15) == 0 && ((uintptr_t)scratch & 255) == 0,
                 "sqnorm: g must be 16-byte and scratch 256-byte aligned");
  B200_DEVICE(device);
  long long want = ((n >> 2) + kRedThreads - 1) / kRedThreads;
  if (want < 1) want = 1;
  const int grid = (int)(want < kRedMaxBlocks ? want : kRedMaxBlocks);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 256);
  sqnorm_kernel<<<grid, kRedThreads, 0, as_stream(stream)>>>(g, n, partials, ticket, out_sq, accumulate);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_clip_coef(const float* sq_dev, float max_norm, int nonfinite_policy, float* coef_out_dev,
                      float* norm_out_dev, int device, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(sq_dev && coef_out_dev, "clip_coef: null buffer");
  B200_DEVICE(device);
  clip_coef_kernel<<<1, 32, 0, as_stream(stream)>>>(sq_dev, max_norm, nonfinite_policy, coef_out_dev,
                                                    norm_out_dev);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
