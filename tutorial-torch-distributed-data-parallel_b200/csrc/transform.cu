// f2 — the input transform of the reference on the GPU.
//
// Reference: data_and_toy_model.py:13-36 — transforms.Resize(224) -> RandomHorizontalFlip -> ToTensor -> Normalize
// on CIFAR-10's 32x32 RGB uint8 images, executed by two DataLoader workers with PIL on the host, followed by a
// blocking 77 MB host-to-device copy per batch (multi-GPU-training-torch.py:110). Here the batch travels as the
// dataset stores it (uint8 [B][32][32][3], 393 KB for 128 images) and ONE kernel produces the normalised fp32 NCHW
// batch in HBM.
//
// Bit-exact with the stock path, which fixes the arithmetic:
//   * Pillow's ImagingResample (src/libImaging/Resample.c): a horizontal then a vertical pass, per-output-pixel
//     normalised bilinear coefficients in 22-bit fixed point, rounding to uint8 BETWEEN the passes;
//     b200ddp_resize_coeffs() restates precompute_coeffs + normalize_coeffs_8bpc (host, double precision);
//   * torchvision to_tensor (x / 255 in fp32) and normalize ((x - mean) / std in fp32, IEEE division).
// The flip decision is the caller's (one byte per image): torchvision draws it from torch's CPU generator inside
// each DataLoader worker, which no other implementation can reproduce; the arithmetic after the decision is exact.
//
// HBM-bound on the fp32 output (4 * 3 * OH * OW bytes per image; the uint8 input is 1/196 of that): one CTA per
// (image, channel, band of output rows), coalesced stores, a 256-entry lookup table per CTA for the float stage.
#include "common.cuh"
#include "../../include/b200ddp.h"

#include <math.h>

namespace b200 {
namespace {

constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kMaxIn = 64;        // input side (CIFAR: 32)
constexpr int kMaxOut = 512;      // output side (reference: 224)

__device__ __forceinline__ unsigned char clip8(int v) {
  v >>= kPrecisionBits;
  return (unsigned char)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

struct XformParams {
  const unsigned char* img;   // [B][H][W][3]
  const unsigned char* flip;  // [B] or null
  const int* tab_h;           // horizontal: [OW] xmin | [OW] count | [OW][3] coef
  const int* tab_v;           // vertical:   [OH] ymin | [OH] count | [OH][3] coef
  float* out;                 // [B][3][OH][OW]
  int B, H, W, OH, OW, band;
  float mean[3], stdv[3];
};

__global__ void __launch_bounds__(256) image_transform_kernel(XformParams p) {
  extern __shared__ unsigned char hbuf[];               // [H][OW] horizontally resampled rows of this channel
  __shared__ float lut[256];
  const int c = blockIdx.y, b = blockIdx.z;
  const int y0 = blockIdx.x * p.band;
  const int y1 = min(y0 + p.band, p.OH);
  lut[threadIdx.x] = ((float)threadIdx.x / 255.0f - p.mean[c]) / p.stdv[c];   // blockDim.x == 256
  // input rows this band needs
  const int r0 = __ldg(p.tab_v + y0);
  const int r1 = __ldg(p.tab_v + (y1 - 1)) + __ldg(p.tab_v + p.OH + (y1 - 1));
  const unsigned char* src = p.img + (long long)b * p.H * p.W * 3 + c;
  const int* hx = p.tab_h;
  const int* hn = p.tab_h + p.OW;
  const int* hk = p.tab_h + 2 * p.OW;
  for (int i = threadIdx.x; i < (r1 - r0) * p.OW; i += blockDim.x) {
    const int r = r0 + i / p.OW, xx = i - (i / p.OW) * p.OW;
    const int x0 = __ldg(hx + xx), n = __ldg(hn + xx);
    int acc = 1 << (kPrecisionBits - 1);
    for (int t = 0; t < n; ++t) acc += (int)__ldg(src + ((long long)r * p.W + x0 + t) * 3) * __ldg(hk + xx * 3 + t);
    hbuf[r * p.OW + xx] = clip8(acc);
  }
  __syncthreads();
  const bool fl = p.flip && p.flip[b];
  const int* vy = p.tab_v;
  const int* vn = p.tab_v + p.OH;
  const int* vk = p.tab_v + 2 * p.OH;
  float* dst = p.out + ((long long)b * 3 + c) * p.OH * p.OW;
  for (int i = threadIdx.x; i < (y1 - y0) * p.OW; i += blockDim.x) {
    const int y = y0 + i / p.OW, x = i - (i / p.OW) * p.OW;
    const int xs = fl ? p.OW - 1 - x : x;                   // flip AFTER the resize, as the reference's Compose does
    const int ry = __ldg(vy + y), n = __ldg(vn + y);
    int acc = 1 << (kPrecisionBits - 1);
    for (int t = 0; t < n; ++t) acc += (int)hbuf[(ry + t) * p.OW + xs] * __ldg(vk + y * 3 + t);
    dst[(long long)y * p.OW + x] = lut[clip8(acc)];
  }
}

double bilinear_filter(double x) {
  if (x < 0.0) x = -x;
  return x < 1.0 ? 1.0 - x : 0.0;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

// Host only. table = [out] first input index | [out] tap count (<= 3) | [out][3] fixed-point coefficients.
int b200ddp_resize_coeffs(int in_size, int out_size, int* table) {
  B200_CHECK_ARG(table && in_size > 0 && out_size > 0, "resize_coeffs: bad argument");
  B200_CHECK_ARG(in_size <= out_size, "resize_coeffs: only enlargements are implemented (support 1.0, 3 taps); got %d -> %d",
                 in_size, out_size);
  const double scale = (double)in_size / out_size;
  const double filterscale = 1.0, support = 1.0, ss = 1.0 / filterscale;
  int* xmins = table;
  int* counts = table + out_size;
  int* coefs = table + 2 * out_size;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    B200_CHECK_ARG(xmax >= 0 && xmax <= 3, "resize_coeffs: unexpected tap count %d", xmax);
    double k[3] = {0, 0, 0}, ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      k[x] = bilinear_filter((x + xmin - center + 0.5) * ss);
      ww += k[x];
    }
    for (int x = 0; x < 3; ++x) {
      double v = (x < xmax && ww != 0.0) ? k[x] / ww : (x < xmax ? k[x] : 0.0);
      coefs[xx * 3 + x] = v < 0 ? (int)(-0.5 + v * (1 << kPrecisionBits)) : (int)(0.5 + v * (1 << kPrecisionBits));
    }
    xmins[xx] = xmin;
    counts[xx] = xmax;
  }
  return 0;
}

int b200ddp_image_transform(const unsigned char* img, const unsigned char* flip, float* out, int B, int H, int W, int OH,
                            int OW, const int* tab_h, const int* tab_v, const float* mean3, const float* std3, int device,
                            void* stream) {
  B200_CHECK_ARG(img && out && tab_h && tab_v && mean3 && std3 && B > 0, "image_transform: bad argument");
  B200_CHECK_ARG(H > 0 && W > 0 && H <= kMaxIn && W <= kMaxIn, "image_transform: input side must be 1..%d (got %dx%d)", kMaxIn, H, W);
  B200_CHECK_ARG(OH >= H && OW >= W && OH <= kMaxOut && OW <= kMaxOut, "image_transform: output side must be input..%d (got %dx%d)",
                 kMaxOut, OH, OW);
  B200_CHECK_ARG(B <= 65535, "image_transform: at most 65535 images per call");
  B200_DEVICE(device);
  XformParams p;
  p.img = img; p.flip = flip; p.tab_h = tab_h; p.tab_v = tab_v; p.out = out;
  p.B = B; p.H = H; p.W = W; p.OH = OH; p.OW = OW;
  for (int c = 0; c < 3; ++c) { p.mean[c] = mean3[c]; p.stdv[c] = std3[c]; }
  // bands: enough CTAs to fill the machine a few times over, at least 8 output rows each
  int bands = (int)((4ll * num_sms() + 3ll * B - 1) / (3ll * B));
  if (bands < 1) bands = 1;
  int band = (OH + bands - 1) / bands;
  if (band < 8) band = 8;
  if (band > OH) band = OH;
  p.band = band;
  bands = (OH + band - 1) / band;
  const size_t smem = (size_t)H * OW;
  image_transform_kernel<<<dim3(bands, 3, B), 256, smem, as_stream(stream)>>>(p);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
