// F1 / F2 / F3 — Linear layers of the classifier (reference alexnet.py:38-44 via
// data_and_toy_model.py:41-45): forward with bias + ReLU + dropout + gate epilogue, dgrad, wgrad,
// the gate/bias backward, and the tiny N=10 head as dedicated kernels.
//
// Two engines sit behind the same entry points:
//   mode 0  strict fp32: register-tiled SIMT sgemm (128x128x16 tiles, 8x8 per thread). This is the
//           parity path: the reference runs these layers as true-fp32 cuBLAS sgemm.
//   mode 1  TF32 tcgen05/TMA (gemm_tf32.cu): the throughput path, HBM-bound at batch 128.
// Both use deterministic split-K (partials in a caller-provided workspace + one finalize kernel
// that also applies the epilogue) so that 148 SMs have work at M = batch = 128.
#include "gemm.h"
#include "../../include/b200ddp.h"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// strict-fp32 SIMT sgemm:  C[m][n] = sum_r A(m,r) * B(r,n)
//   A_RC: A stored [M][R] (r contiguous, lda)   else stored [R][M] (m contiguous, lda)
//   B_RC: B stored [N][R] (r contiguous, ldb)   else stored [R][N] (n contiguous, ldb)
// ------------------------------------------------------------------------------------------------
constexpr int SBM = 128, SBN = 128, SBK = 16, SPAD = 4;

template <bool RC>
__device__ __forceinline__ void load_tile_regs(const float* __restrict__ base, long long ld, int mn0, int MN, int r0,
                                               int r_end, bool vec, float (&v)[2][4]) {
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int f = t + i * 256;
    if (RC) {
      const int row = f >> 2, c4 = f & 3;
      const int g_mn = mn0 + row, g_r = r0 + c4 * 4;
      const float* src = base + (long long)g_mn * ld + g_r;
      if (g_mn < MN && g_r + 3 < r_end && vec) {
        const float4 q = *reinterpret_cast<const float4*>(src);
        v[i][0] = q.x; v[i][1] = q.y; v[i][2] = q.z; v[i][3] = q.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[i][j] = (g_mn < MN && g_r + j < r_end) ? src[j] : 0.f;
      }
    } else {
      const int r = f >> 5, c4 = f & 31;
      const int g_r = r0 + r, g_mn = mn0 + c4 * 4;
      const float* src = base + (long long)g_r * ld + g_mn;
      if (g_r < r_end && g_mn + 3 < MN && vec) {
        const float4 q = *reinterpret_cast<const float4*>(src);
        v[i][0] = q.x; v[i][1] = q.y; v[i][2] = q.z; v[i][3] = q.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[i][j] = (g_r < r_end && g_mn + j < MN) ? src[j] : 0.f;
      }
    }
  }
}

template <bool RC>
__device__ __forceinline__ void store_tile_smem(float (*S)[SBM + SPAD], const float (&v)[2][4]) {
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int f = t + i * 256;
    if (RC) {
      const int row = f >> 2, c4 = f & 3;
#pragma unroll
      for (int j = 0; j < 4; ++j) S[c4 * 4 + j][row] = v[i][j];
    } else {
      const int r = f >> 5, c4 = f & 31;
      *reinterpret_cast<float4*>(&S[r][c4 * 4]) = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
    }
  }
}

struct SgemmParams {
  const float* A; long long lda;
  const float* B; long long ldb;
  float* C; long long ldc;
  float* ws;            // split partials [z][M][N] when splits > 1
  int M, N, R;
  int r_per_split, splits;
  int a_vec, b_vec, c_vec;
  Epilogue ep;
};

template <bool A_RC, bool B_RC>
__global__ void __launch_bounds__(256) sgemm_kernel(SgemmParams p) {
  __shared__ __align__(16) float As[SBK][SBM + SPAD];
  __shared__ __align__(16) float Bs[SBK][SBN + SPAD];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * SBM, n0 = blockIdx.x * SBN, z = blockIdx.z;
  const int r_begin = z * p.r_per_split;
  const int r_end = min(p.R, r_begin + p.r_per_split);

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float ra[2][4], rb[2][4];
  load_tile_regs<A_RC>(p.A, p.lda, m0, p.M, r_begin, r_end, p.a_vec, ra);
  load_tile_regs<B_RC>(p.B, p.ldb, n0, p.N, r_begin, r_end, p.b_vec, rb);
  for (int r0 = r_begin; r0 < r_end; r0 += SBK) {
    store_tile_smem<A_RC>(As, ra);
    store_tile_smem<B_RC>(Bs, rb);
    __syncthreads();
    if (r0 + SBK < r_end) {   // prefetch the next k-tile into registers while computing this one
      load_tile_regs<A_RC>(p.A, p.lda, m0, p.M, r0 + SBK, r_end, p.a_vec, ra);
      load_tile_regs<B_RC>(p.B, p.ldb, n0, p.N, r0 + SBK, r_end, p.b_vec, rb);
    }
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 4 + 64]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4 + 64]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const bool partial = p.splits > 1;
  float* dst = partial ? p.ws + (size_t)z * (size_t)p.M * (size_t)p.N : p.C;
  const long long ld = partial ? p.N : p.ldc;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= p.M) continue;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int n = n0 + tx * 4 + h * 64;
      float v[4] = {acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]};
      if (!partial) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n + j < p.N) v[j] = apply_epilogue(v[j], p.ep, m, n + j, ld);
      }
      float* o = dst + (long long)m * ld + n;
      if (n + 3 < p.N && (partial ? ((p.N & 3) == 0) : p.c_vec)) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (n + j < p.N) o[j] = v[j];
      }
    }
  }
}

// out[row*ldo + col] = epilogue(sum_z ws[z][row][col]); ws slices dense [rows x cols].
__global__ void __launch_bounds__(256) splitk_finalize_kernel(const float* __restrict__ ws, int splits, int rows,
                                                              int cols, float* __restrict__ out, long long ldo,
                                                              Epilogue ep) {
  const long long total = (long long)rows * cols;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    float s = 0.f;
    for (int z = 0; z < splits; ++z) s += __ldg(ws + (long long)z * total + i);
    const long long row = i / cols, col = i - row * cols;
    out[row * ldo + col] = apply_epilogue(s, ep, row, col, ldo);
  }
}

// Same, four consecutive columns per thread (cols % 4 == 0, 16-byte aligned slices and output rows): 16-byte loads,
// four partial slices in flight at a time, 32-bit index arithmetic. The partial sums are added in slice order, as in
// the scalar kernel (bit-identical results).
__global__ void __launch_bounds__(256) splitk_finalize_vec4_kernel(const float* __restrict__ ws, int splits, int rows,
                                                                   int cols, float* __restrict__ out, long long ldo,
                                                                   Epilogue ep) {
  const int cols4 = cols >> 2;
  const int total4 = rows * cols4;                    // host guarantees rows*cols < 2^31
  const long long slice = (long long)rows * cols;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const int row = i / cols4, c4 = i - row * cols4;
    const float4* src = reinterpret_cast<const float4*>(ws) + i;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int z = 0;
    for (; z + 4 <= splits; z += 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(src + (long long)(z + u) * (slice >> 2));
#pragma unroll
      for (int u = 0; u < 4; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    }
    for (; z < splits; ++z) {
      const float4 v = __ldg(src + (long long)z * (slice >> 2));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    const long long col = (long long)c4 << 2;
    float4 o;
    o.x = apply_epilogue(acc.x, ep, row, col, ldo);
    o.y = apply_epilogue(acc.y, ep, row, col + 1, ldo);
    o.z = apply_epilogue(acc.z, ep, row, col + 2, ldo);
    o.w = apply_epilogue(acc.w, ep, row, col + 3, ldo);
    *reinterpret_cast<float4*>(out + (long long)row * ldo + col) = o;
  }
}

int splitk_finalize_launch(const float* ws, int splits, int rows, int cols, float* out, long long ldo,
                           const Epilogue& ep, cudaStream_t stream) {
  const long long total = (long long)rows * cols;
  const bool vec = (cols & 3) == 0 && (ldo & 3) == 0 && total < (1LL << 31) && (((uintptr_t)ws | (uintptr_t)out) & 15) == 0;
  if (vec) {
    long long want = (total / 4 + 255) / 256;
    const long long cap = (long long)num_sms() * 16;
    const int grid = (int)(want < cap ? (want < 1 ? 1 : want) : cap);
    splitk_finalize_vec4_kernel<<<grid, 256, 0, stream>>>(ws, splits, rows, cols, out, ldo, ep);
  } else {
    long long want = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    const int grid = (int)(want < cap ? (want < 1 ? 1 : want) : cap);
    splitk_finalize_kernel<<<grid, 256, 0, stream>>>(ws, splits, rows, cols, out, ldo, ep);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("splitk_finalize launch -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

static int sgemm_launch(bool a_rc, bool b_rc, const float* A, long long lda, const float* B, long long ldb, float* C,
                        long long ldc, int M, int N, int R, int splits, float* ws, const Epilogue& ep,
                        cudaStream_t stream) {
  SgemmParams p;
  p.A = A; p.lda = lda; p.B = B; p.ldb = ldb; p.C = C; p.ldc = ldc; p.ws = ws;
  p.M = M; p.N = N; p.R = R;
  if (splits < 1) splits = 1;
  int per = ((R + splits - 1) / splits + SBK - 1) / SBK * SBK;
  if (per < SBK) per = SBK;
  p.r_per_split = per;
  p.splits = (R + per - 1) / per;
  p.a_vec = (((uintptr_t)A & 15) == 0) && ((lda & 3) == 0);
  p.b_vec = (((uintptr_t)B & 15) == 0) && ((ldb & 3) == 0);
  p.c_vec = (((uintptr_t)C & 15) == 0) && ((ldc & 3) == 0);
  p.ep = ep;
  if (p.splits > 1 && !ws) {
    set_error("sgemm: split-K needs a workspace");
    return kInvalidArg;
  }
  dim3 grid((N + SBN - 1) / SBN, (M + SBM - 1) / SBM, p.splits);
  if (a_rc && b_rc) sgemm_kernel<true, true><<<grid, 256, 0, stream>>>(p);
  else if (a_rc && !b_rc) sgemm_kernel<true, false><<<grid, 256, 0, stream>>>(p);
  else if (!a_rc && b_rc) sgemm_kernel<false, true><<<grid, 256, 0, stream>>>(p);
  else sgemm_kernel<false, false><<<grid, 256, 0, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("sgemm launch -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  if (p.splits > 1) return splitk_finalize_launch(ws, p.splits, M, N, C, ldc, ep, stream);
  return 0;
}

static inline int sgemm_effective_splits(int R, int splits) {
  if (splits < 1) splits = 1;
  int per = ((R + splits - 1) / splits + SBK - 1) / SBK * SBK;
  if (per < SBK) per = SBK;
  return (R + per - 1) / per;
}

// pick a split count so that tiles * splits fits in ONE wave of 148 SMs (a 149th CTA would double the
// kernel time: these kernels run one CTA per SM), keeping >= 128 of K per split
constexpr int kMaxSplits = 16;
static inline int pick_splits(int tiles, int R) {
  int s = sm_budget() / tiles;
  if (s > kMaxSplits) s = kMaxSplits;
  const int max_by_k = R / 128 > 0 ? R / 128 : 1;
  if (s > max_by_k) s = max_by_k;
  return s < 1 ? 1 : s;
}

// ------------------------------------------------------------------------------------------------
// F3: the 10-class head (N <= 16) — not a GEMM; one warp per batch row / simple streaming kernels
// ------------------------------------------------------------------------------------------------
constexpr int kSmallN = 16;

// One CTA (8 warps) per batch row: warp w covers K-slice w, partial sums meet in shared memory.
__global__ void __launch_bounds__(256) linear_smalln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                                const float* __restrict__ bias, float* __restrict__ y,
                                                                int B, int N, int K) {
  __shared__ float part[8][kSmallN];
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc[kSmallN];
#pragma unroll
  for (int n = 0; n < kSmallN; ++n) acc[n] = 0.f;
  const float* xr = x + (long long)b * K;
  if ((K & 3) == 0 && (((uintptr_t)x | (uintptr_t)w) & 15) == 0) {
    for (int k = (warp * 32 + lane) * 4; k < K; k += 256 * 4) {
      const float4 xv = *reinterpret_cast<const float4*>(xr + k);
#pragma unroll
      for (int n = 0; n < kSmallN; ++n) {
        if (n < N) {
          const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (long long)n * K + k));
          acc[n] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[n]))));
        }
      }
    }
  } else {
    for (int k = warp * 32 + lane; k < K; k += 256) {
      const float xv = xr[k];
#pragma unroll
      for (int n = 0; n < kSmallN; ++n)
        if (n < N) acc[n] = fmaf(xv, __ldg(w + (long long)n * K + k), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < kSmallN; ++n) {
    if (n < N) {
      const float s = warp_sum(acc[n]);
      if (lane == 0) part[warp][n] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x < N) {
    float s = bias ? bias[threadIdx.x] : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += part[j][threadIdx.x];
    y[(long long)b * N + threadIdx.x] = s;
  }
}

// dx[b][k] = sum_n dy[b][n] * w[n][k]
__global__ void __launch_bounds__(256) linear_smalln_dx_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                               float* __restrict__ dx, int B, int N, int K) {
  const long long total = (long long)B * K, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int b = (int)(i / K), k = (int)(i - (long long)b * K);
    float s = 0.f;
    for (int n = 0; n < N; ++n) s = fmaf(__ldg(dy + (long long)b * N + n), __ldg(w + (long long)n * K + k), s);
    dx[i] = s;
  }
}

// dw[n][k] = sum_b dy[b][n] * x[b][k];  db[n] = sum_b dy[b][n] (block 0 of the k-grid does db)
__global__ void __launch_bounds__(256) linear_smalln_dw_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                               float* __restrict__ dw, float* __restrict__ db, int B,
                                                               int N, int K) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = blockIdx.y;
  if (k < K) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s = fmaf(__ldg(dy + (long long)b * N + n), __ldg(x + (long long)b * K + k), s);
    dw[(long long)n * K + k] = s;
  }
  if (db && blockIdx.x == 0 && threadIdx.x == 0) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dy[(long long)b * N + n];
    db[n] = s;
  }
}

// dz = dy * gate-scale, db[n] = sum_b dz[b][n]. Block = 32 columns x 32 row lanes (the kernel is a few MB of traffic:
// latency bound, so as many independent loads in flight as possible).
constexpr int kGbRows = 32;
__global__ void __launch_bounds__(32 * kGbRows) gate_bias_bwd_kernel(const float* __restrict__ dy,
                                                            const unsigned char* __restrict__ gate, float drop_scale,
                                                            float* __restrict__ dz, float* __restrict__ db, int B, int N) {
  __shared__ float red[kGbRows][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float s = 0.f;
  if (n < N) {
#pragma unroll 4
    for (int b = ry; b < B; b += kGbRows) {
      const long long i = (long long)b * N + n;
      float v = dy[i];
      if (gate) {
        const unsigned char g = gate[i];
        v = (g == 0) ? 0.f : (g == 2 ? v * drop_scale : v);
      }
      if (dz) dz[i] = v;
      s += v;
    }
  }
  red[ry][cx] = s;
  __syncthreads();
  if (ry == 0 && n < N && db) {
    float tot = 0.f;
#pragma unroll
    for (int j = 0; j < kGbRows; ++j) tot += red[j][cx];
    db[n] = tot;
  }
}

static Epilogue make_epilogue(const float* bias, int relu, float dropout_p, unsigned long long seed,
                              unsigned char* gate) {
  Epilogue ep;
  ep.bias = bias;
  ep.relu = relu;
  ep.dropout = dropout_p > 0.f;
  if (ep.dropout) {
    double thr = (double)dropout_p * 4294967296.0;
    ep.p_threshold = thr >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)thr;
    ep.drop_scale = 1.0f / (1.0f - dropout_p);
  }
  ep.seed = seed;
  ep.gate = gate;
  return ep;
}

}  // namespace b200

using namespace b200;

extern "C" {

long long b200ddp_linear_workspace_bytes(int B, int N, int K) {
  // split-K partials: at most kMaxSplits slices of the largest of the forward ([B x N]) and dgrad ([B x K]) outputs.
  const long long a = (long long)B * N, b = (long long)B * K;
  return (long long)kMaxSplits * (a > b ? a : b) * 4ll;
}

int b200ddp_linear_fwd(const float* x, const float* w, const float* bias, float* y, unsigned char* gate, int B, int N,
                       int K, int relu, float dropout_p, unsigned long long seed, int mode, void* ws,
                       long long ws_bytes, int device, void* stream) {
  B200_CHECK_ARG(x && w && y && B > 0 && N > 0 && K > 0, "linear_fwd: bad argument");
  B200_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "linear_fwd: dropout_p=%f out of [0,1)", dropout_p);
  B200_DEVICE(device);
  cudaStream_t s = as_stream(stream);
  if (N <= kSmallN) {
    B200_CHECK_ARG(!relu && dropout_p == 0.f && !gate, "linear_fwd: no epilogue on the small-N head");
    linear_smalln_fwd_kernel<<<B, 256, 0, s>>>(x, w, bias, y, B, N, K);
    B200_LAUNCH_OK();
    return 0;
  }
  const Epilogue ep = make_epilogue(bias, relu, dropout_p, seed, gate);
  if (mode == 1) {
    const int tiles = ceil_div(N, 128 * tf32_row_subtiles(N, K)) * ceil_div(B, 128);
    const int splits = tf32_effective_splits(K, pick_splits(tiles, K));
    B200_CHECK_ARG(splits == 1 || (ws && ws_bytes >= (long long)splits * B * N * 4), "linear_fwd: workspace too small");
    int rc = tf32_gemm_launch(w, K, 0, x, K, 0, /*M=*/N, /*N=*/B, /*R=*/K, y, N, splits, (float*)ws, ep, s);
    if (rc) return rc;
    if (splits > 1) return splitk_finalize_launch((const float*)ws, splits, B, N, y, N, ep, s);
    return 0;
  }
  const int tiles = ceil_div(N, SBN) * ceil_div(B, SBM);
  const int splits = sgemm_effective_splits(K, pick_splits(tiles, K));
  B200_CHECK_ARG(splits == 1 || (ws && ws_bytes >= (long long)splits * B * N * 4), "linear_fwd: workspace too small");
  return sgemm_launch(true, true, x, K, w, K, y, N, B, N, K, splits, (float*)ws, ep, s);
}

int b200ddp_linear_bwd_dx(const float* dz, const float* w, float* dx, int B, int N, int K, int mode, void* ws,
                          long long ws_bytes, int device, void* stream) {
  B200_CHECK_ARG(dz && w && dx && B > 0 && N > 0 && K > 0, "linear_bwd_dx: bad argument");
  B200_DEVICE(device);
  cudaStream_t s = as_stream(stream);
  if (N <= kSmallN) {
    const long long total = (long long)B * K;
    long long want = (total + 255) / 256;
    const int grid = (int)(want < num_sms() * 8 ? want : num_sms() * 8);
    linear_smalln_dx_kernel<<<grid, 256, 0, s>>>(dz, w, dx, B, N, K);
    B200_LAUNCH_OK();
    return 0;
  }
  Epilogue ep;
  if (mode == 1) {
    const int tiles = ceil_div(K, 128 * tf32_row_subtiles(K, N)) * ceil_div(B, 128);
    const int splits = tf32_effective_splits(N, pick_splits(tiles, N));
    B200_CHECK_ARG(splits == 1 || (ws && ws_bytes >= (long long)splits * B * K * 4), "linear_bwd_dx: workspace too small");
    // D[m=k][n=b] = sum_f W[f][k] * dZ[b][f]
    int rc = tf32_gemm_launch(w, K, 1, dz, N, 0, /*M=*/K, /*N=*/B, /*R=*/N, dx, K, splits, (float*)ws, ep, s);
    if (rc) return rc;
    if (splits > 1) return splitk_finalize_launch((const float*)ws, splits, B, K, dx, K, ep, s);
    return 0;
  }
  const int tiles = ceil_div(K, SBN) * ceil_div(B, SBM);
  const int splits = sgemm_effective_splits(N, pick_splits(tiles, N));
  B200_CHECK_ARG(splits == 1 || (ws && ws_bytes >= (long long)splits * B * K * 4), "linear_bwd_dx: workspace too small");
  // C[m=b][n=k] = sum_f dZ[b][f] * W[f][k]
  return sgemm_launch(true, false, dz, N, w, K, dx, K, B, K, N, splits, (float*)ws, ep, s);
}

int b200ddp_linear_bwd_dw(const float* dz, const float* x, float* dw, float* db, int B, int N, int K, int mode,
                          int device, void* stream) {
  B200_CHECK_ARG(dz && x && dw && B > 0 && N > 0 && K > 0, "linear_bwd_dw: bad argument");
  B200_DEVICE(device);
  cudaStream_t s = as_stream(stream);
  if (N <= kSmallN) {
    dim3 grid(ceil_div(K, 256), N);
    linear_smalln_dw_kernel<<<grid, 256, 0, s>>>(dz, x, dw, db, B, N, K);
    B200_LAUNCH_OK();
    return 0;
  }
  if (db) {
    gate_bias_bwd_kernel<<<ceil_div(N, 32), 32 * kGbRows, 0, s>>>(dz, nullptr, 1.f, nullptr, db, B, N);
    B200_LAUNCH_OK();
  }
  Epilogue ep;
  if (mode == 1) {
    // D[m=k][n=f] = sum_b X[b][k] * dZ[b][f]  ->  dW[f][k]; reduction length = batch, no split
    return tf32_gemm_launch(x, K, 1, dz, N, 1, /*M=*/K, /*N=*/N, /*R=*/B, dw, K, 1, nullptr, ep, s);
  }
  // C[m=f][n=k] = sum_b dZ[b][f] * X[b][k]
  return sgemm_launch(false, false, dz, N, x, K, dw, K, N, K, B, 1, nullptr, ep, s);
}

int b200ddp_gate_bias_bwd(const float* dy, const unsigned char* gate, float dropout_p, float* dz, float* db, int B,
                          int N, int device, void* stream) {
  B200_CHECK_ARG(dy && B > 0 && N > 0 && (dz || db), "gate_bias_bwd: bad argument");
  B200_DEVICE(device);
  const float scale = dropout_p > 0.f ? 1.0f / (1.0f - dropout_p) : 1.0f;
  gate_bias_bwd_kernel<<<ceil_div(N, 32), 32 * kGbRows, 0, as_stream(stream)>>>(dy, gate, scale, dz, db, B, N);
  B200_LAUNCH_OK();
  return 0;
}

}  // extern "C"
