// C1-C5, strict mode — the five convolutions in true fp32 (SIMT FMA, no tensor cores, no TF32 rounding).
//
// The product path (conv.cu) computes with TF32 operands like the reference's cuDNN default. SURVEY 8(d) config 2
// also asks for a STRICT comparison (loss rtol 1e-5, gradients rel-L2 1e-5 against torch with
// `torch.backends.cudnn.allow_tf32 = False`): these three kernels are that mode for torchvision alexnet.py:22-32
// (forward), and for the dgrad / wgrad halves of loss.backward() (reference multi-GPU-training-torch.py:125).
// `AlexNet(mode="fp32")` routes every Conv2d here. They are direct (gather-form, atomics-free, deterministic)
// convolutions written for exactness, not speed: products in fp32, row sums in fp32, sums of rows in fp64.
// Weights are read in torch's own layout [Cout][Cin][KH][KW]; any stride / padding.
#include "common.cuh"
#include "../../include/b200ddp.h"

namespace b200 {
namespace {

struct ConvGeom {
  int Nb, Cin, H, W, Cout, OH, OW, KH, KW, stride, pad;
};

// y[n][co][oh][ow] = bias[co] + sum_{ci,kh,kw} x[n][ci][oh*s-p+kh][ow*s-p+kw] * w[co][ci][kh][kw]
// one thread per output element, consecutive threads = consecutive ow (coalesced x rows, broadcast weights)
__global__ void __launch_bounds__(256) conv_fwd_fp32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ bias, float* __restrict__ y,
                                                            ConvGeom g, int relu) {
  const long long total = (long long)g.Nb * g.Cout * g.OH * g.OW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int ow = (int)(i % g.OW);
    long long t = i / g.OW;
    const int oh = (int)(t % g.OH);
    t /= g.OH;
    const int co = (int)(t % g.Cout), n = (int)(t / g.Cout);
    const int ih0 = oh * g.stride - g.pad, iw0 = ow * g.stride - g.pad;
    double acc = 0.0;
    for (int ci = 0; ci < g.Cin; ++ci) {
      const float* xp = x + ((long long)n * g.Cin + ci) * g.H * g.W;
      const float* wp = w + ((long long)co * g.Cin + ci) * g.KH * g.KW;
      float part = 0.f;
      for (int kh = 0; kh < g.KH; ++kh) {
        const int ih = ih0 + kh;
        if ((unsigned)ih >= (unsigned)g.H) continue;
        for (int kw = 0; kw < g.KW; ++kw) {
          const int iw = iw0 + kw;
          if ((unsigned)iw < (unsigned)g.W) part = fmaf(__ldg(xp + (long long)ih * g.W + iw), __ldg(wp + kh * g.KW + kw), part);
        }
      }
      acc += (double)part;
    }
    float v = (float)acc;
    if (bias) v += __ldg(bias + co);
    if (relu) v = v > 0.f ? v : 0.f;
    y[i] = v;
  }
}

// dx[n][ci][ih][iw] = sum_{co,kh,kw : (ih+p-kh) % s == 0, (iw+p-kw) % s == 0} dy[n][co][(ih+p-kh)/s][(iw+p-kw)/s] * w[co][ci][kh][kw]
__global__ void __launch_bounds__(256) conv_dgrad_fp32_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                              float* __restrict__ dx, ConvGeom g) {
  const long long total = (long long)g.Nb * g.Cin * g.H * g.W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int iw = (int)(i % g.W);
    long long t = i / g.W;
    const int ih = (int)(t % g.H);
    t /= g.H;
    const int ci = (int)(t % g.Cin), n = (int)(t / g.Cin);
    double acc = 0.0;
    for (int co = 0; co < g.Cout; ++co) {
      const float* dp = dy + ((long long)n * g.Cout + co) * g.OH * g.OW;
      const float* wp = w + ((long long)co * g.Cin + ci) * g.KH * g.KW;
      float part = 0.f;
      for (int kh = 0; kh < g.KH; ++kh) {
        const int a = ih + g.pad - kh;
        if (a < 0 || a % g.stride) continue;
        const int oh = a / g.stride;
        if (oh >= g.OH) continue;
        for (int kw = 0; kw < g.KW; ++kw) {
          const int b = iw + g.pad - kw;
          if (b < 0 || b % g.stride) continue;
          const int ow = b / g.stride;
          if (ow < g.OW) part = fmaf(__ldg(dp + (long long)oh * g.OW + ow), __ldg(wp + kh * g.KW + kw), part);
        }
      }
      acc += (double)part;
    }
    dx[i] = (float)acc;
  }
}

// dw[co][ci][kh][kw] = sum_{n,oh,ow} dy[n][co][oh][ow] * x[n][ci][oh*s-p+kh][ow*s-p+kw]
// one CTA per weight element (co, ci, kh, kw): threads stride over (n, oh) rows, fp32 along a row, fp64 across rows,
// deterministic block reduction.
__global__ void __launch_bounds__(128) conv_wgrad_fp32_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                              float* __restrict__ dw, ConvGeom g) {
  __shared__ double red[32];
  const long long e = blockIdx.x;
  const int kw = (int)(e % g.KW);
  long long t = e / g.KW;
  const int kh = (int)(t % g.KH);
  t /= g.KH;
  const int ci = (int)(t % g.Cin), co = (int)(t / g.Cin);
  double acc = 0.0;
  const int rows = g.Nb * g.OH;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const int n = r / g.OH, oh = r - n * g.OH;
    const int ih = oh * g.stride - g.pad + kh;
    if ((unsigned)ih >= (unsigned)g.H) continue;
    const float* dp = dy + (((long long)n * g.Cout + co) * g.OH + oh) * g.OW;
    const float* xp = x + (((long long)n * g.Cin + ci) * g.H + ih) * g.W;
    float part = 0.f;
    for (int ow = 0; ow < g.OW; ++ow) {
      const int iw = ow * g.stride - g.pad + kw;
      if ((unsigned)iw < (unsigned)g.W) part = fmaf(__ldg(dp + ow), __ldg(xp + iw), part);
    }
    acc += (double)part;
  }
  const double tot = block_sum<double>(acc, red);
  if (threadIdx.x == 0) dw[e] = (float)tot;
}

int geom(ConvGeom& g, int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride, int pad, const char* who) {
  B200_CHECK_ARG(Nb > 0 && Cin > 0 && Cout > 0 && H > 0 && W > 0 && KH > 0 && KW > 0 && stride > 0 && pad >= 0,
                 "%s: bad shape", who);
  g.Nb = Nb; g.Cin = Cin; g.H = H; g.W = W; g.Cout = Cout; g.KH = KH; g.KW = KW; g.stride = stride; g.pad = pad;
  g.OH = (H + 2 * pad - KH) / stride + 1;
  g.OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(g.OH > 0 && g.OW > 0, "%s: empty output", who);
  return 0;
}

inline int grid_for(long long total) {
  long long want = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 32;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

int b200ddp_conv2d_fwd_fp32(const float* x, const float* w, const float* bias, float* y, int Nb, int Cin, int H, int W,
                            int Cout, int KH, int KW, int stride, int pad, int relu, int device, void* stream) {
  B200_CHECK_ARG(x && w && y, "conv2d_fwd_fp32: null argument");
  ConvGeom g;
  if (int rc = geom(g, Nb, Cin, H, W, Cout, KH, KW, stride, pad, "conv2d_fwd_fp32")) return rc;
  B200_DEVICE(device);
  const long long total = (long long)Nb * Cout * g.OH * g.OW;
  conv_fwd_fp32_kernel<<<grid_for(total), 256, 0, as_stream(stream)>>>(x, w, bias, y, g, relu);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_conv2d_dgrad_fp32(const float* dy, const float* w, float* dx, int Nb, int Cin, int H, int W, int Cout, int KH,
                              int KW, int stride, int pad, int device, void* stream) {
  B200_CHECK_ARG(dy && w && dx, "conv2d_dgrad_fp32: null argument");
  ConvGeom g;
  if (int rc = geom(g, Nb, Cin, H, W, Cout, KH, KW, stride, pad, "conv2d_dgrad_fp32")) return rc;
  B200_DEVICE(device);
  const long long total = (long long)Nb * Cin * H * W;
  conv_dgrad_fp32_kernel<<<grid_for(total), 256, 0, as_stream(stream)>>>(dy, w, dx, g);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_conv2d_wgrad_fp32(const float* x, const float* dy, float* dw, int Nb, int Cin, int H, int W, int Cout, int KH,
                              int KW, int stride, int pad, int device, void* stream) {
  B200_CHECK_ARG(x && dy && dw, "conv2d_wgrad_fp32: null argument");
  ConvGeom g;
  if (int rc = geom(g, Nb, Cin, H, W, Cout, KH, KW, stride, pad, "conv2d_wgrad_fp32")) return rc;
  const long long total = (long long)Cout * Cin * KH * KW;
  B200_CHECK_ARG(total < (1ll << 31), "conv2d_wgrad_fp32: weight tensor too large");
  B200_DEVICE(device);
  conv_wgrad_fp32_kernel<<<(int)total, 128, 0, as_stream(stream)>>>(x, dy, dw, g);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
