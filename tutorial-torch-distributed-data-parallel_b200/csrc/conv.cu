// C1-C5 — the five convolutions of the feature extractor (torchvision alexnet.py:22,25,28,30,32, reached
// from reference multi-GPU-training-torch.py:121/:125) as implicit GEMMs on the 5th-generation tensor
// cores: TF32 operands read straight from the fp32 NCHW tensors, fp32 accumulation in TMEM.
// (The reference runs these through cuDNN, also in TF32 — torch's default for convolutions.)
//
// No im2col matrix is ever materialised and no NCHW<->NHWC conversion pass runs (the stock path spends
// ~1 ms per step in cudnn::nchwToNhwcKernel / nhwcToNchwKernel): four producer warps gather the im2col
// tile of each k-block directly into shared memory in the swizzled layout tcgen05.mma expects, the
// weight tile arrives by TMA, one thread issues the MMAs, and the same four warps run the epilogue.
//
//   forward / dgrad   D[m = pixel][n = co] = sum_k  col(pixel, k) * W[co][k]        k = (ci, kh, kw)
//       A = col tile, MN-major (32 consecutive output pixels are contiguous in smem and — for stride 1 —
//           in the input row), layout SWIZZLE_128B_BASE32B; B = W rows, K-major, TMA SWIZZLE_128B.
//       Epilogue: + bias[co], ReLU, coalesced stores y[n][co][pix] (lanes = consecutive pixels).
//       dgrad of a stride-1 convolution is the same kernel run on dY with the flipped / transposed
//       weights produced by conv_weight_prep (pad' = KH-1-pad).
//   wgrad             D[m = k][n = co]     = sum_pixel col(pixel, k) * dY[co][pixel]
//       both operands K-major (32 consecutive pixels per smem row), gathered by the producer warps;
//       the pixel range is split over CTAs (deterministic partials + finalize) because only
//       ceil(K/128) x ceil(Cout/BN) output tiles exist.
#include "gemm.h"
#include "tc.cuh"
#include "../../include/b200ddp.h"

namespace b200 {

namespace {

using namespace tc;

constexpr int CBM = 128;            // UMMA_M
constexpr int CBK = 32;             // fp32 per k-block (128 bytes)
constexpr int CUK = 8;              // K per tcgen05.mma kind::tf32
constexpr int kConvThreads = 160;   // warps 0-3: gather producers, then epilogue; warp 4: TMEM + MMA issue
constexpr int kInvalidPos = 0x7fff7fff;

template <int BN>
constexpr uint32_t tmem_cols() { return BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256; }

struct ConvFwdParams {
  const float* x;
  float* y;
  const float* bias;
  const int* koff;     // [kblocks*32] ci*H*W + kh*W + kw
  const int* kpos;     // [kblocks*32] kh | kw << 16 (kInvalidPos beyond K)
  int Nb, Cin, H, W, Cout, OH, OW, stride, pad;
  int kblocks, M, OHW, relu, stages;
};

template <int BN>
__global__ void __launch_bounds__(kConvThreads)
conv_fwd_kernel(const __grid_constant__ CUtensorMap tmW, ConvFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = CBM * CBK * 4;
  constexpr uint32_t B_BYTES = BN * CBK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * CBM, n0 = blockIdx.y * BN;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 128);     // every producer thread arrives once per phase (+ TMA bytes)
      mbar_init(&empty_bar[s], 1);      // tcgen05.commit
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
  }
  if (warp == 4) tmem_alloc(tmem_slot, tmem_cols<BN>());
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ---------------- im2col gather producer: thread <-> output pixel ----------------
    const int m = m0 + tid;
    const bool valid_m = m < p.M;
    const int mm = valid_m ? m : 0;
    const int n_img = mm / p.OHW, pix = mm - n_img * p.OHW;
    const int oh = pix / p.OW, ow = pix - oh * p.OW;
    const int ih0 = oh * p.stride - p.pad, iw0 = ow * p.stride - p.pad;
    const long long xbase = (long long)n_img * p.Cin * p.H * p.W + (long long)ih0 * p.W + iw0;
    // MN-major SWIZZLE_128B_BASE32B: k-row r at r*128 B, 32-byte chunk c stored at chunk c ^ (r & 3);
    // the 32-pixel atom `warp` starts at warp*4096 B.
    const uint32_t chunk = (uint32_t)lane >> 3;
    const uint32_t s_thread = (uint32_t)warp * 4096u + ((uint32_t)lane & 7u) * 4u;
    for (int kb = 0; kb < p.kblocks; ++kb) {
      const int s = kb % stages;
      const uint32_t ph = (uint32_t)(kb / stages) & 1u;
      mbar_wait(&empty_bar[s], ph ^ 1u);
      uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
      if (tid == 0) {
        mbar_expect_tx(&full_bar[s], B_BYTES);
        tma_load_2d(sa + A_BYTES, &tmW, &full_bar[s], kb * CBK, n0);       // box {32 k, BN co}
      }
      const int my_off = __ldg(p.koff + kb * CBK + lane);
      const int my_pos = __ldg(p.kpos + kb * CBK + lane);
      const uint32_t sdst = smem_u32(sa) + s_thread;
#pragma unroll 8
      for (int r = 0; r < CBK; ++r) {
        const int off = __shfl_sync(0xffffffffu, my_off, r);
        const int pos = __shfl_sync(0xffffffffu, my_pos, r);
        const int ih = ih0 + (pos & 0xffff), iw = iw0 + (pos >> 16);
        const bool ok = valid_m && (unsigned)ih < (unsigned)p.H && (unsigned)iw < (unsigned)p.W;
        float v = 0.f;
        if (ok) v = __ldg(p.x + xbase + off);
        const uint32_t addr = sdst + (uint32_t)r * 128u + ((chunk ^ ((uint32_t)r & 3u)) << 5);
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
      }
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[s]);
    }
    // ---------------- epilogue: TMEM lane = pixel, columns = output channels ----------------
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    float* yb = p.y + (long long)n_img * p.Cout * p.OHW + pix;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, r);
      if (valid_m) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int co = n0 + c + j;
          if (co < p.Cout) {
            float v = __uint_as_float(r[j]);
            if (p.bias) v += __ldg(p.bias + co);
            if (p.relu) v = v > 0.f ? v : 0.f;
            yb[(long long)co * p.OHW] = v;
          }
        }
      }
    }
  } else if (lane == 0) {
    // ---------------- MMA issuer ----------------
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) /* A MN-major */ |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(CBM >> 4) << 24);
    for (int kb = 0; kb < p.kblocks; ++kb) {
      const int s = kb % stages;
      const uint32_t ph = (uint32_t)(kb / stages) & 1u;
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
      const uint32_t sb = sa + A_BYTES;
#pragma unroll
      for (int k = 0; k < CBK / CUK; ++k) {
        const uint64_t adesc = make_smem_desc(sa + k * 1024u, 4096u, 512u, 1u);
        const uint64_t bdesc = make_smem_desc(sb + k * (CUK * 4), 16u, 1024u, 2u);
        umma_tf32(tmem_base, adesc, bdesc, idesc, (kb > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit(&empty_bar[s]);
    }
    umma_commit(tmem_full_bar);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, tmem_cols<BN>());
}

struct ConvWgradParams {
  const float* x;
  const float* dy;
  float* out;          // dW [Cout][K] (splits == 1) or partials [z][Cout][K]
  const int* koff;     // [m_tiles*128]
  const int* kpos;
  int Nb, Cin, H, W, Cout, OH, OW, stride, pad;
  int K, M, OHW, kblocks_total, kblocks_per_split, splits, stages;
};

template <int BN>
__global__ void __launch_bounds__(kConvThreads)
conv_wgrad_kernel(ConvWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = CBM * CBK * 4;
  constexpr uint32_t B_BYTES = BN * CBK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  int* s_koff = reinterpret_cast<int*>(tmem_slot + 4);
  int* s_kpos = s_koff + CBM;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * CBM, n0 = blockIdx.y * BN, z = blockIdx.z;
  const int kb_begin = z * p.kblocks_per_split;
  int kb_end = kb_begin + p.kblocks_per_split;
  if (kb_end > p.kblocks_total) kb_end = p.kblocks_total;
  const int nkb = kb_end - kb_begin;

  if (tid < CBM) {
    s_koff[tid] = __ldg(p.koff + m0 + tid);
    s_kpos[tid] = __ldg(p.kpos + m0 + tid);
  }
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 128);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) tmem_alloc(tmem_slot, tmem_cols<BN>());
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ---------------- producers: lane <-> pixel of the k-block, warp <-> quarter of the tile rows ----------------
    constexpr int B_ROWS = BN / 4;
    // K-major SWIZZLE_128B: row i at i*128 B, 16-byte chunk c stored at chunk c ^ (i & 7)
    const uint32_t c16 = (uint32_t)lane >> 2, e16 = ((uint32_t)lane & 3u) * 4u;
    for (int i = 0; i < nkb; ++i) {
      const int s = i % stages;
      const uint32_t ph = (uint32_t)(i / stages) & 1u;
      mbar_wait(&empty_bar[s], ph ^ 1u);
      const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
      const uint32_t sb = sa + A_BYTES;
      const int pxl = (kb_begin + i) * CBK + lane;
      const bool valid_p = pxl < p.M;
      const int pp = valid_p ? pxl : 0;
      const int n_img = pp / p.OHW, pix = pp - n_img * p.OHW;
      const int oh = pix / p.OW, ow = pix - oh * p.OW;
      const int ih0 = oh * p.stride - p.pad, iw0 = ow * p.stride - p.pad;
      const long long xbase = (long long)n_img * p.Cin * p.H * p.W + (long long)ih0 * p.W + iw0;
      const float* dyb = p.dy + (long long)n_img * p.Cout * p.OHW + pix;
#pragma unroll 8
      for (int j = 0; j < 32; ++j) {
        const int row = warp * 32 + j;
        const int off = s_koff[row], pos = s_kpos[row];
        const int ih = ih0 + (pos & 0xffff), iw = iw0 + (pos >> 16);
        const bool ok = valid_p && (unsigned)ih < (unsigned)p.H && (unsigned)iw < (unsigned)p.W;
        float v = 0.f;
        if (ok) v = __ldg(p.x + xbase + off);
        const uint32_t addr = sa + (uint32_t)row * 128u + ((c16 ^ ((uint32_t)row & 7u)) << 4) + e16;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
      }
#pragma unroll 8
      for (int j = 0; j < B_ROWS; ++j) {
        const int row = warp * B_ROWS + j;
        const int co = n0 + row;
        float v = 0.f;
        if (valid_p && co < p.Cout) v = __ldg(dyb + (long long)co * p.OHW);
        const uint32_t addr = sb + (uint32_t)row * 128u + ((c16 ^ ((uint32_t)row & 7u)) << 4) + e16;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
      }
      fence_proxy_async_smem();
      mbar_arrive(&full_bar[s]);
    }
    // ---------------- epilogue: TMEM lane = k index, columns = output channels ----------------
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const int m = m0 + tid;
    float* ob = p.out + (long long)z * p.Cout * p.K + m;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, r);
      if (m < p.K) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int co = n0 + c + j;
          if (co < p.Cout) ob[(long long)co * p.K] = __uint_as_float(r[j]);
        }
      }
    }
  } else if (lane == 0) {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(CBM >> 4) << 24);
    for (int i = 0; i < nkb; ++i) {
      const int s = i % stages;
      const uint32_t ph = (uint32_t)(i / stages) & 1u;
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
      const uint32_t sb = sa + A_BYTES;
#pragma unroll
      for (int k = 0; k < CBK / CUK; ++k) {
        const uint64_t adesc = make_smem_desc(sa + k * (CUK * 4), 16u, 1024u, 2u);
        const uint64_t bdesc = make_smem_desc(sb + k * (CUK * 4), 16u, 1024u, 2u);
        umma_tf32(tmem_base, adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit(&empty_bar[s]);
    }
    umma_commit(tmem_full_bar);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, tmem_cols<BN>());
}

// mode 0: out[co][k] = w[co][k] for k < K, zero padding up to ld_out (row stride multiple of 4 for TMA)
// mode 1: dgrad weights  out[ci][co*KH*KW + (KH-1-kh)*KW + (KW-1-kw)] = w[co][ci][kh][kw]
__global__ void __launch_bounds__(256) conv_weight_prep_kernel(const float* __restrict__ w, float* __restrict__ out,
                                                               int Cout, int Cin, int KH, int KW, long long ld_out,
                                                               int mode) {
  const int KHW = KH * KW;
  if (mode == 0) {
    const int K = Cin * KHW;
    const long long total = (long long)Cout * ld_out, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
      const int co = (int)(i / ld_out), k = (int)(i - (long long)co * ld_out);
      out[i] = k < K ? __ldg(w + (long long)co * K + k) : 0.f;
    }
  } else {
    const long long total = (long long)Cout * Cin * KHW, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
      // i enumerates the OUTPUT: [ci][co][kh'][kw']
      const int kwp = (int)(i % KW);
      long long t = i / KW;
      const int khp = (int)(t % KH);
      t /= KH;
      const int co = (int)(t % Cout), ci = (int)(t / Cout);
      const long long src = (((long long)co * Cin + ci) * KH + (KH - 1 - khp)) * KW + (KW - 1 - kwp);
      out[(long long)ci * ld_out + (long long)co * KHW + khp * KW + kwp] = __ldg(w + src);
    }
  }
}

// out[c] = sum over (n, pix) of dy[n][c][pix]  (bias gradient). S x C CTAs, deterministic last-ticket merge.
constexpr int kCsThreads = 256;
constexpr int kCsMaxChannels = 4096;
__global__ void __launch_bounds__(kCsThreads) channel_sum_kernel(const float* __restrict__ dy, int Nb, int C, int HW,
                                                                 double* __restrict__ partials,
                                                                 unsigned int* __restrict__ tickets,
                                                                 float* __restrict__ out) {
  __shared__ double red[32];
  __shared__ bool last;
  const int s = blockIdx.x, S = gridDim.x, c = blockIdx.y;
  float a = 0.f;
  for (int b = s; b < Nb; b += S) {
    const float* pch = dy + ((long long)b * C + c) * HW;
    for (int i = threadIdx.x; i < HW; i += kCsThreads) a += __ldg(pch + i);
  }
  const double da = block_sum<double>((double)a, red);
  if (threadIdx.x == 0) {
    partials[(long long)c * S + s] = da;
    __threadfence();
    last = (atomicAdd(&tickets[c], 1u) == (unsigned)S - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double t = 0;
    for (int j = 0; j < S; ++j) t += ((volatile double*)partials)[(long long)c * S + j];
    out[c] = (float)t;
    tickets[c] = 0u;
  }
}

inline int cs_splits(int Nb, int C) {
  int s = (kNumSMs * 4 + C - 1) / C;
  if (s > Nb) s = Nb;
  if (s > 64) s = 64;
  return s < 1 ? 1 : s;
}

template <int BN>
int launch_fwd(const CUtensorMap& tmW, const ConvFwdParams& p, cudaStream_t stream) {
  const size_t stage_bytes = (size_t)(CBM + BN) * CBK * 4;
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 16;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_fwd_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv_fwd_kernel<%d>) -> %s", BN, cudaGetErrorString(e));
      return kCudaErrBase + (int)e;
    }
    attr_set = true;
  }
  dim3 grid((p.M + CBM - 1) / CBM, (p.Cout + BN - 1) / BN, 1);
  conv_fwd_kernel<BN><<<grid, kConvThreads, smem, stream>>>(tmW, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv_fwd_kernel<%d> launch (grid %d,%d smem %zu) -> %s", BN, grid.x, grid.y, smem, cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

template <int BN>
int launch_wgrad(const ConvWgradParams& p, cudaStream_t stream) {
  const size_t stage_bytes = (size_t)(CBM + BN) * CBK * 4;
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 32 + 2 * CBM * 4;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv_wgrad_kernel<%d>) -> %s", BN, cudaGetErrorString(e));
      return kCudaErrBase + (int)e;
    }
    attr_set = true;
  }
  dim3 grid((p.K + CBM - 1) / CBM, (p.Cout + BN - 1) / BN, p.splits);
  conv_wgrad_kernel<BN><<<grid, kConvThreads, smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv_wgrad_kernel<%d> launch (grid %d,%d,%d smem %zu) -> %s", BN, grid.x, grid.y, grid.z, smem,
              cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

// output-channel tile: the whole Cout when it fits one MMA (<= 256), else the largest divisor-friendly tile
inline int pick_bn(int Cout) {
  if (Cout <= 64) return 64;
  if (Cout <= 128) return 128;
  if (Cout <= 192) return 192;
  if (Cout <= 256) return 256;
  if (Cout % 192 == 0) return 192;
  return 256;
}

struct WgradPlan {
  int bn, m_tiles, n_tiles, kblocks_total, splits, per;
};

inline WgradPlan plan_wgrad(int Nb, int Cin, int OH, int OW, int Cout, int KH, int KW) {
  WgradPlan w;
  const int K = Cin * KH * KW;
  w.bn = pick_bn(Cout);
  w.m_tiles = (K + CBM - 1) / CBM;
  w.n_tiles = (Cout + w.bn - 1) / w.bn;
  const long long M = (long long)Nb * OH * OW;
  w.kblocks_total = (int)((M + CBK - 1) / CBK);
  int splits = kNumSMs / (w.m_tiles * w.n_tiles);
  if (splits < 1) splits = 1;
  if (splits > 64) splits = 64;
  if (splits > w.kblocks_total) splits = w.kblocks_total;
  w.per = (w.kblocks_total + splits - 1) / splits;
  w.splits = (w.kblocks_total + w.per - 1) / w.per;
  return w;
}

}  // namespace

}  // namespace b200

using namespace b200;

extern "C" {

int b200ddp_conv2d_fwd(const float* x, const float* wmat, long long ldw, const float* bias, float* y, const int* koff,
                       const int* kpos, int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride, int pad,
                       int relu, int device, void* stream) {
  B200_CHECK_ARG(x && wmat && y && koff && kpos, "conv2d_fwd: null argument");
  B200_CHECK_ARG(Nb > 0 && Cin > 0 && Cout > 0 && KH > 0 && KW > 0 && stride > 0 && pad >= 0, "conv2d_fwd: bad shape");
  B200_CHECK_ARG((ldw & 3) == 0 && ((uintptr_t)wmat & 15) == 0, "conv2d_fwd: weight rows must be 16-byte aligned (ldw %% 4 == 0)");
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(OH > 0 && OW > 0, "conv2d_fwd: empty output");
  B200_CHECK_ARG((long long)Nb * OH * OW < (1ll << 31) && (long long)Cin * H * W < (1ll << 31), "conv2d_fwd: tensor too large");
  B200_DEVICE(device);
  const int K = Cin * KH * KW;
  ConvFwdParams p;
  p.x = x; p.y = y; p.bias = bias; p.koff = koff; p.kpos = kpos;
  p.Nb = Nb; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.OH = OH; p.OW = OW; p.stride = stride; p.pad = pad;
  p.kblocks = (K + CBK - 1) / CBK;
  p.M = Nb * OH * OW; p.OHW = OH * OW; p.relu = relu;
  const int bn = pick_bn(Cout);
  const size_t stage_bytes = (size_t)(CBM + bn) * CBK * 4;
  int stages = (int)((200 * 1024) / stage_bytes);
  if (stages > 6) stages = 6;
  if (stages > p.kblocks) stages = p.kblocks;
  if (stages < 2) stages = 2;
  p.stages = stages;
  CUtensorMap tmW;
  int rc = tma_map_2d(&tmW, wmat, ldw, Cout, ldw, CBK, bn, 0);
  if (rc) return rc;
  cudaStream_t s = as_stream(stream);
  switch (bn) {
    case 64: return launch_fwd<64>(tmW, p, s);
    case 128: return launch_fwd<128>(tmW, p, s);
    case 192: return launch_fwd<192>(tmW, p, s);
    default: return launch_fwd<256>(tmW, p, s);
  }
}

long long b200ddp_conv2d_wgrad_workspace_bytes(int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride,
                                               int pad) {
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  const WgradPlan w = plan_wgrad(Nb, Cin, OH, OW, Cout, KH, KW);
  return w.splits > 1 ? (long long)w.splits * Cout * Cin * KH * KW * 4 : 0;
}

int b200ddp_conv2d_wgrad(const float* x, const float* dy, float* dw, const int* koff, const int* kpos, int Nb, int Cin,
                         int H, int W, int Cout, int KH, int KW, int stride, int pad, void* ws, long long ws_bytes,
                         int device, void* stream) {
  B200_CHECK_ARG(x && dy && dw && koff && kpos, "conv2d_wgrad: null argument");
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(OH > 0 && OW > 0 && Nb > 0, "conv2d_wgrad: bad shape");
  B200_CHECK_ARG((long long)Nb * OH * OW < (1ll << 31) && (long long)Cin * H * W < (1ll << 31), "conv2d_wgrad: tensor too large");
  B200_DEVICE(device);
  const WgradPlan w = plan_wgrad(Nb, Cin, OH, OW, Cout, KH, KW);
  const int K = Cin * KH * KW;
  B200_CHECK_ARG(w.splits == 1 || (ws && ws_bytes >= (long long)w.splits * Cout * K * 4), "conv2d_wgrad: workspace too small");
  ConvWgradParams p;
  p.x = x; p.dy = dy; p.koff = koff; p.kpos = kpos;
  p.out = w.splits > 1 ? reinterpret_cast<float*>(ws) : dw;
  p.Nb = Nb; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.OH = OH; p.OW = OW; p.stride = stride; p.pad = pad;
  p.K = K; p.M = Nb * OH * OW; p.OHW = OH * OW;
  p.kblocks_total = w.kblocks_total; p.kblocks_per_split = w.per; p.splits = w.splits;
  const size_t stage_bytes = (size_t)(CBM + w.bn) * CBK * 4;
  int stages = (int)((200 * 1024) / stage_bytes);
  if (stages > 4) stages = 4;
  if (stages > w.per) stages = w.per;
  if (stages < 2) stages = 2;
  p.stages = stages;
  cudaStream_t s = as_stream(stream);
  int rc;
  switch (w.bn) {
    case 64: rc = launch_wgrad<64>(p, s); break;
    case 128: rc = launch_wgrad<128>(p, s); break;
    case 192: rc = launch_wgrad<192>(p, s); break;
    default: rc = launch_wgrad<256>(p, s); break;
  }
  if (rc) return rc;
  if (w.splits > 1) {
    Epilogue ep;
    return splitk_finalize_launch(reinterpret_cast<const float*>(ws), w.splits, Cout, K, dw, K, ep, s);
  }
  return 0;
}

int b200ddp_conv2d_weight_prep(const float* w, float* out, int Cout, int Cin, int KH, int KW, long long ld_out, int mode,
                               int device, void* stream) {
  B200_CHECK_ARG(w && out && Cout > 0 && Cin > 0 && KH > 0 && KW > 0, "conv2d_weight_prep: bad argument");
  B200_CHECK_ARG(mode == 0 ? ld_out >= (long long)Cin * KH * KW : ld_out >= (long long)Cout * KH * KW,
                 "conv2d_weight_prep: ld_out too small");
  B200_DEVICE(device);
  const long long total = mode == 0 ? (long long)Cout * ld_out : (long long)Cout * Cin * KH * KW;
  long long want = (total + 255) / 256;
  const int grid = (int)(want < kNumSMs * 8 ? (want < 1 ? 1 : want) : kNumSMs * 8);
  conv_weight_prep_kernel<<<grid, 256, 0, as_stream(stream)>>>(w, out, Cout, Cin, KH, KW, ld_out, mode);
  B200_LAUNCH_OK();
  return 0;
}

long long b200ddp_channel_sum_scratch_bytes(int Nb, int C) {
  return (long long)kCsMaxChannels * 4 + (long long)C * cs_splits(Nb, C) * 8;
}

int b200ddp_channel_sum(const float* dy, float* out, int Nb, int C, int HW, void* scratch, long long scratch_bytes,
                        int device, void* stream) {
  B200_CHECK_ARG(dy && out && scratch && Nb > 0 && C > 0 && HW > 0, "channel_sum: bad argument");
  B200_CHECK_ARG(C <= kCsMaxChannels, "channel_sum: at most %d channels", kCsMaxChannels);
  B200_CHECK_ARG(scratch_bytes >= b200ddp_channel_sum_scratch_bytes(Nb, C), "channel_sum: scratch too small");
  B200_CHECK_ARG(((uintptr_t)scratch & 255) == 0, "channel_sum: scratch must be 256-byte aligned");
  B200_DEVICE(device);
  const int S = cs_splits(Nb, C);
  unsigned int* tickets = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + (long long)kCsMaxChannels * 4);
  channel_sum_kernel<<<dim3(S, C), kCsThreads, 0, as_stream(stream)>>>(dy, Nb, C, HW, partials, tickets, out);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
