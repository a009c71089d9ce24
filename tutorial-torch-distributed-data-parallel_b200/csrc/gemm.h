// Private interface between the linear-layer entry points (linear.cu) and the two GEMM
// engines: the strict-fp32 SIMT kernel (linear.cu) and the TF32 tcgen05/TMA kernel (gemm_tf32.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

namespace b200 {

// Fused epilogue applied to an output matrix stored row-major as [rows x cols] with leading
// dimension ld: v = acc + bias[col]; relu; dropout (Philox keyed by row*ld+col); gate byte.
struct Epilogue {
  const float* bias = nullptr;      // per column, may be null
  int relu = 0;
  int dropout = 0;
  uint32_t p_threshold = 0;         // keep iff philox >= p_threshold
  float drop_scale = 1.f;           // 1/(1-p)
  unsigned long long seed = 0;
  unsigned char* gate = nullptr;    // optional [rows x cols] (same ld): 0 = off, 1 = on, 2 = on*drop_scale
};

__device__ __forceinline__ float apply_epilogue(float v, const Epilogue& ep, long long row, long long col,
                                                long long ld) {
  if (ep.bias) v += __ldg(ep.bias + col);
  unsigned char gate = 1;
  if (ep.relu && !(v > 0.f)) { v = 0.f; gate = 0; }
  if (ep.dropout) {
    if (dropout_keep(ep.seed, (uint64_t)(row * ld + col), ep.p_threshold)) {
      v *= ep.drop_scale;
      if (gate) gate = 2;
    } else {
      v = 0.f; gate = 0;
    }
  }
  if (ep.gate) ep.gate[row * ld + col] = gate;
  return v;
}

// TF32 tensor-core GEMM  D[m][n] = sum_r A(m,r) * B(n,r), written TRANSPOSED: out[n*ldo + m].
//   a_mn_major = 0: A stored [M x R] (r contiguous, lda = row stride); 1: A stored [R x M] (m contiguous).
//   b_mn_major likewise for B ([N x R] or [R x N]).
// splits > 1 writes raw partials to ws[z][n][m] (ld = M) and the caller runs splitk_finalize.
int tf32_gemm_launch(const float* A, long long lda, int a_mn_major, const float* B, long long ldb,
                     int b_mn_major, int M, int N, int R, float* out, long long ldo, int splits, float* ws,
                     const Epilogue& ep, cudaStream_t stream);

// 2-D fp32 TMA tensor map over a row-major [outer x inner] matrix with row stride ld (elements).
// swizzle_base32b = 0: SWIZZLE_128B (K-major operands); 1: SWIZZLE_128B_ATOM_32B (MN-major 32-bit operands).
int tma_map_2d(CUtensorMap* map, const float* base, long long inner, long long outer, long long ld, int box_inner,
               int box_outer, int swizzle_base32b);

// N-D (2..5) fp32 tiled tensor map without swizzle: gdim[0] is the contiguous dimension, gstr[i] = byte stride of dimension
// i+1, box = elements per dimension. Out-of-range elements of a box read as zero.
int tma_map_nd(CUtensorMap* map, const float* base, int rank, const long long* gdim, const long long* gstr_bytes, const int* box);

// First-layer kernel (conv1.cu): k_order 2 of b200ddp_conv2d_fwd.
int conv1_fwd_launch(const float* x, const float* wmat, long long ldw, const float* bias, float* y, int Nb, int Cin, int H,
                     int W, int Cout, int KH, int KW, int stride, int pad, int relu, cudaStream_t stream);

long long conv1_wgrad_workspace_bytes(int Nb, int H, int W, int pad);
int conv1_wgrad_launch(const float* x, const float* dy, float* dw, int Nb, int Cin, int H, int W, int Cout, int KH, int KW,
                       int stride, int pad, void* ws, long long ws_bytes, cudaStream_t stream);

// 128-row sub-tiles per CTA that tf32_gemm_launch uses for an M x N x R problem. The kernel supports 2 (two TMEM
// accumulators sharing one B tile, parity-tested), but on the classifier GEMMs it measured SLOWER than 1
// (FC1 forward 57 vs 50 us, dX 53 vs 47 us on B200). ncu (profiles/r02_gemm_ncu_summary.md): at 1 the weight stream runs
// at 0.71 of the measured HBM copy rate with one CTA per SM and L2 at 41 % of its throughput although every CTA re-reads
// the activation tile (L2->SM bytes = 2x the weight bytes) - L2->SM traffic is not the limiter, the number of SMs that
// stream is (128 / 144 CTAs on 148 SMs), and fewer, fatter CTAs only lose parallelism.
inline int tf32_row_subtiles(int M, int R) { (void)M; (void)R; return 1; }

// Number of K splits tf32_gemm_launch will really use for a request (no empty splits).
inline int tf32_effective_splits(int R, int splits) {
  const int total_kb = (R + 31) / 32;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  const int per = (total_kb + splits - 1) / splits;
  return (total_kb + per - 1) / per;
}

// out[row*ldo + col] = epilogue(sum_z ws[z][row][col]), ws slices are dense [rows x cols].
int splitk_finalize_launch(const float* ws, int splits, int rows, int cols, float* out, long long ldo,
                           const Epilogue& ep, cudaStream_t stream);

}  // namespace b200
