// O1 — single-pass Adam over flat parameter / gradient / moment arenas.
//
// Replaces torch.optim.Adam's 7-kernel foreach chain (torch/optim/adam.py:705-800,
// called from reference multi-GPU-training-torch.py:126) by ONE launch that reads
// p,g,m,v once and writes p,m,v once: 28 B/param, the algorithmic minimum.
// HBM-bound: float4 streaming accesses, grid = 148 SMs x 8 CTAs, grid-stride loop.
#include "common.cuh"
#include "../../include/b200ddp.h"

#include <math.h>

namespace b200 {

struct AdamScalars {
  float one_minus_beta1;
  float beta2;
  float one_minus_beta2;
  float neg_step_size;   // -(lr / (1 - beta1^t))
  float inv_bc2_sqrt;    // unused when dividing; kept for the fast path
  float bc2_sqrt;        // sqrt(1 - beta2^t)
  float eps;
  float weight_decay;
};

// The optimizer streams 28 bytes per parameter exactly once while (inside backward) the convolution kernels next to it
// live off activations that sit in the 126 MB L2: every access carries an evict-first L2 policy so that the stream does
// not push those activations out.
__device__ __forceinline__ uint64_t evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ float4 ld_ef(const float4* p, uint64_t pol) {
  float4 r;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void st_ef(float4* p, const float4& v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "l"(pol)
               : "memory");
}

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamScalars& s,
                                         float gscale) {
  g *= gscale;
  if (s.weight_decay != 0.f) g = fmaf(s.weight_decay, p, g);
  m = fmaf(g - m, s.one_minus_beta1, m);                 // lerp(m, g, 1-beta1)
  v = fmaf(s.one_minus_beta2 * g, g, v * s.beta2);       // v*beta2 + (1-beta2)*g*g
  const float denom = sqrtf(v) / s.bc2_sqrt + s.eps;     // IEEE sqrt/div (no fast-math) for parity
  p = fmaf(s.neg_step_size, m / denom, p);
}

// 128-thread CTAs on purpose: when the optimizer runs per bucket INSIDE backward (reducer._fused_adam), its CTAs have to
// squeeze in beside a convolution CTA that holds 57,344 of the SM's 65,536 registers; 128 threads x 40 registers fit
// the remaining 8,192, 256 threads do not (they would only run in the gaps between convolution kernels).
constexpr int kAdamThreads = 128;
__global__ void __launch_bounds__(kAdamThreads) adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        long long n, AdamScalars s,
                                                        const float* __restrict__ gscale_dev) {
  const float gscale = gscale_dev ? __ldg(gscale_dev) : 1.0f;
  const uint64_t pol = evict_first_policy();
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = ld_ef(p4 + i, pol);
    const float4 gg = ld_ef(g4 + i, pol);
    float4 mm = ld_ef(m4 + i, pol);
    float4 vv = ld_ef(v4 + i, pol);
    adam_one(pp.x, gg.x, mm.x, vv.x, s, gscale);
    adam_one(pp.y, gg.y, mm.y, vv.y, s, gscale);
    adam_one(pp.z, gg.z, mm.z, vv.z, s, gscale);
    adam_one(pp.w, gg.w, mm.w, vv.w, s, gscale);
    st_ef(p4 + i, pp, pol);
    st_ef(m4 + i, mm, pol);
    st_ef(v4 + i, vv, pol);
  }
  // scalar tail (n % 4 elements)
  const long long tail0 = n4 << 2;
  const long long t = tail0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    float pp = p[t], mm = m[t], vv = v[t];
    adam_one(pp, g[t], mm, vv, s, gscale);
    p[t] = pp; m[t] = mm; v[t] = vv;
  }
}

}  // namespace b200

extern "C" int b200ddp_adam_step_flat(float* p, const float* g, float* m, float* v, long long n, double lr,
                                      double beta1, double beta2, double eps, double weight_decay,
                                      long long step, const float* grad_scale_dev, int device,
                                      void* stream) {
  using namespace b200;
  B200_CHECK_ARG(n >= 0, "adam_step_flat: n=%lld < 0", n);
  if (n == 0) return 0;
  B200_CHECK_ARG(p && g && m && v, "adam_step_flat: null buffer");
  B200_CHECK_ARG(step >= 1, "adam_step_flat: step=%lld must be >= 1", step);
  B200_CHECK_ARG((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0,
                 "adam_step_flat: buffers must be 16-byte aligned");
  B200_DEVICE(device);
  AdamScalars s;
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  s.one_minus_beta1 = (float)(1.0 - beta1);
  s.beta2 = (float)beta2;
  s.one_minus_beta2 = (float)(1.0 - beta2);
  s.neg_step_size = (float)(-(lr / bc1));
  s.bc2_sqrt = (float)sqrt(bc2);
  s.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
  s.eps = (float)eps;
  s.weight_decay = (float)weight_decay;
  const long long n4 = n >> 2;
  long long want = (n4 + kAdamThreads - 1) / kAdamThreads;
  if (want < 1) want = 1;
  const long long cap = (long long)num_sms() * 16;
  const int grid = (int)(want < cap ? want : cap);
  adam_flat_kernel<<<grid, kAdamThreads, 0, as_stream(stream)>>>(p, g, m, v, n, s, grad_scale_dev);
  B200_LAUNCH_OK();
  return 0;
}
