// F1 / F2 — TF32 tensor-core GEMM for the classifier's Linear layers, hand-written for sm_100a:
// TMA (cp.async.bulk.tensor, 128B swizzle) -> shared memory ring -> tcgen05.mma kind::tf32 with
// the fp32 accumulator in TMEM -> tcgen05.ld epilogue with bias / ReLU / dropout / gate folded in.
//
// The operands are the fp32 master weights and activations exactly as they sit in HBM: the tensor
// core reads the top 19 bits of each fp32 word, so there is no conversion pass and no low-precision
// copy of the 151 MB / 67 MB weight matrices (reference alexnet.py:39,42 run as fp32 cuBLAS sgemm;
// at batch 128 these GEMMs are weight-streaming bound).
//
// Shape convention:  D[m][n] = sum_r A(m,r) * B(n,r), stored transposed: out[n*ldo + m]. The M role
// is always given to the dimension that is contiguous in the OUTPUT, so that the 32 lanes of an
// epilogue warp (= 32 consecutive TMEM lanes = 32 consecutive m) write 128 contiguous bytes.
//   forward   Y[b][f]  : M=f (A = W  [f][k], K-major),  N=b (B = X  [b][k], K-major),  R=k
//   dgrad     dX[b][k] : M=k (A = W  [f][k], MN-major), N=b (B = dZ [b][f], K-major),  R=f
//   wgrad     dW[f][k] : M=k (A = X  [b][k], MN-major), N=f (B = dZ [b][f], MN-major), R=b
//
// One CTA computes one 128 x BN output tile for one K split. Warp roles: warp 0 = TMA producer,
// warp 1 = TMEM allocator + single-thread MMA issuer, warps 2..5 = epilogue (TMEM lane quarters).
#include "gemm.h"
#include "tc.cuh"

#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

namespace b200 {

namespace {

using namespace tc;

constexpr int BM = 128;          // UMMA_M (cta_group::1)
constexpr int BK = 32;           // fp32 elements per k-block = one 128-byte swizzle span
constexpr int UMMA_K = 8;        // kind::tf32: 32 bytes of K per instruction
constexpr int kThreads = 192;
struct KParams {
  int M, N, R;
  int kblocks_per_split;
  int splits;
  int stages;
  int a_mn, b_mn;
  float* out;
  long long ldo;
  float* ws;
  Epilogue ep;
};

// EPI = false: raw accumulator store (split-K partials, dgrad, wgrad) — no epilogue code is compiled in.
// EPI = true : bias / ReLU / Philox dropout / gate byte applied while the tile leaves TMEM.
// RS = 128-row sub-tiles of the output per CTA (two accumulators in TMEM share one B tile): in the classifier GEMMs the
// M side is the streamed weight matrix and the B tile (activations) is re-read from L2 by every CTA - at RS = 1 that
// doubles the L2->SM traffic of each streamed weight byte (the measured L2 cap, not HBM, bounded those kernels).
template <int BN, bool EPI, int RS>
__global__ void __launch_bounds__(kThreads)
tf32_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, KParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128B swizzle pattern (TMA and UMMA must agree on it).
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_SUB_BYTES = BM * BK * 4;   // 16 KB
  constexpr uint32_t A_BYTES = RS * A_SUB_BYTES;
  constexpr uint32_t B_BYTES = BN * BK * 4;
  constexpr uint32_t TMEM_COLS = RS * BN;         // 64, 128 or 256: powers of two
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * (BM * RS), n0 = blockIdx.y * BN, z = blockIdx.z;
  // Split-K is INTERLEAVED: split z owns k-blocks z, z+splits, z+2*splits, ... so that the CTAs working on the
  // same rows of the streamed operand touch adjacent 128-byte pieces at the same time (DRAM page locality)
  // instead of 'splits' far-apart column ranges.
  const int total_kb = (p.R + BK - 1) / BK;
  const int nkb = (total_kb - z + p.splits - 1) / p.splits;   // host guarantees splits <= total_kb
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ---------------- TMA producer ----------------
    for (int i = 0; i < nkb; ++i) {
      const int s = i % stages;
      const uint32_t ph = (uint32_t)(i / stages) & 1u;
      mbar_wait(&empty_bar[s], ph ^ 1u);
      mbar_arrive_expect_tx(&full_bar[s], STAGE_BYTES);
      uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
      uint8_t* sb = sa + A_BYTES;
      const int r0 = (z + i * p.splits) * BK;
      if (!p.a_mn) {
#pragma unroll
        for (int j = 0; j < RS; ++j) tma_load_2d(sa + j * A_SUB_BYTES, &tmA, &full_bar[s], r0, m0 + j * BM);  // box {32 r, 128 m}
      } else {
#pragma unroll
        for (int j = 0; j < RS * BM / 32; ++j) tma_load_2d(sa + j * 4096, &tmA, &full_bar[s], m0 + 32 * j, r0);  // box {32 m, 32 r}
      }
      if (!p.b_mn) {
        tma_load_2d(sb, &tmB, &full_bar[s], r0, n0);                       // box {32 r, BN n}
      } else {
#pragma unroll
        for (int j = 0; j < BN / 32; ++j) tma_load_2d(sb + j * 4096, &tmB, &full_bar[s], n0 + 32 * j, r0);
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (one thread) ----------------
    const uint32_t idesc = (1u << 4)                      // D format  = F32
                           | (2u << 7) | (2u << 10)       // A, B format = TF32
                           | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16)
                           | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    mma_issue_loop<BK / UMMA_K, RS>(tmem_base, idesc, smem_u32(smem), STAGE_BYTES, A_BYTES,
                                    p.a_mn ? operand_mn_major() : operand_k_major(),
                                    p.b_mn ? operand_mn_major() : operand_k_major(), smem_u32(&full_bar[0]),
                                    smem_u32(&empty_bar[0]), smem_u32(tmem_full_bar), stages, nkb, nullptr, A_SUB_BYTES, BN);
  } else if (warp >= 2) {
    // ---------------- epilogue: TMEM -> registers -> global ----------------
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const int q = warp & 3;                              // TMEM lane quarter this warp may touch
    float* dst;
    long long ld;
    if (p.splits > 1) { dst = p.ws + (size_t)z * (size_t)p.N * (size_t)p.M; ld = p.M; }
    else { dst = p.out; ld = p.ldo; }
#pragma unroll 1
    for (int cc = 0; cc < RS * BN; cc += 32) {
      const int sub = cc / BN, c = cc - sub * BN;          // accumulator `sub` = rows m0 + sub*128 ..
      const int m = m0 + sub * BM + q * 32 + lane;
      const bool m_ok = m < p.M;
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)cc, r);
      if (m_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = n0 + c + j;
          if (n < p.N) {
            float v = __uint_as_float(r[j]);
            if (EPI) v = apply_epilogue(v, p.ep, n, m, ld);
            dst[(long long)n * ld + m] = v;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
}  // namespace

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

static int get_encode() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
  if (!g_encode) {
    set_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return kNotInitialised;
  }
  return 0;
}

// 2-D fp32 tensor map over a row-major [outer x inner] matrix with row stride ld (elements).
int tma_map_2d(CUtensorMap* map, const float* base, long long inner, long long outer, long long ld, int box_inner,
               int box_outer, int swizzle_base32b) {
  int rc0 = get_encode();
  if (rc0) return rc0;
  const CUtensorMapSwizzle swizzle = swizzle_base32b ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 4ull};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstr, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d) inner=%lld outer=%lld ld=%lld box=%dx%d", (int)r, inner,
              outer, ld, box_inner, box_outer);
    return kDriverErrBase + (int)r;
  }
  return 0;
}

int tma_map_nd(CUtensorMap* map, const float* base, int rank, const long long* gdim, const long long* gstr_bytes, const int* box) {
  int rc0 = get_encode();
  if (rc0) return rc0;
  if (rank < 2 || rank > 5) {
    set_error("tma_map_nd: rank %d", rank);
    return kInvalidArg;
  }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = (cuuint64_t)gdim[i];
    bx[i] = (cuuint32_t)box[i];
    es[i] = 1;
    if (i + 1 < rank) gs[i] = (cuuint64_t)gstr_bytes[i];
  }
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<float*>(base), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (rank %d) failed (CUresult %d)", rank, (int)r);
    return kDriverErrBase + (int)r;
  }
  return 0;
}

namespace {

template <int BN, bool EPI, int RS>
int launch_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, const KParams& p, cudaStream_t stream) {
  const size_t stage_bytes = (size_t)(BM * RS + BN) * BK * 4;
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 16;
  static std::atomic<unsigned long long> attr_devices{0};   // per template instance, one bit per device
  {
    cudaError_t e = ensure_max_dyn_smem(tf32_gemm_kernel<BN, EPI, RS>, attr_devices);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(tf32_gemm_kernel) -> %s", cudaGetErrorString(e));
      return kCudaErrBase + (int)e;
    }
  }
  dim3 grid((p.M + BM * RS - 1) / (BM * RS), (p.N + BN - 1) / BN, p.splits);
  tf32_gemm_kernel<BN, EPI, RS><<<grid, kThreads, smem, stream>>>(tmA, tmB, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("tf32_gemm_kernel<%d> launch (grid %d,%d,%d smem %zu) -> %s", BN, grid.x, grid.y, grid.z, smem,
              cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

}  // namespace

int tf32_gemm_launch(const float* A, long long lda, int a_mn, const float* B, long long ldb, int b_mn, int M, int N,
                     int R, float* out, long long ldo, int splits, float* ws, const Epilogue& ep,
                     cudaStream_t stream) {
  int rc = 0;
  if ((((uintptr_t)A | (uintptr_t)B) & 15) || (lda & 3) || (ldb & 3)) {
    set_error("tf32 gemm: operands must be 16-byte aligned with leading dimensions multiple of 4 (lda=%lld ldb=%lld)",
              lda, ldb);
    return kInvalidArg;
  }
  const int BN = (N > 64) ? 128 : 64;
  CUtensorMap tmA, tmB;
  rc = a_mn ? tma_map_2d(&tmA, A, M, R, lda, 32, 32, 1) : tma_map_2d(&tmA, A, R, M, lda, BK, BM, 0);
  if (rc) return rc;
  rc = b_mn ? tma_map_2d(&tmB, B, N, R, ldb, 32, 32, 1) : tma_map_2d(&tmB, B, R, N, ldb, BK, BN, 0);
  if (rc) return rc;
  const int total_kb = (R + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  KParams p;
  p.M = M; p.N = N; p.R = R;
  p.kblocks_per_split = (total_kb + splits - 1) / splits;
  p.splits = (total_kb + p.kblocks_per_split - 1) / p.kblocks_per_split;   // no empty splits
  const int rs = tf32_row_subtiles(M, R);
  const int max_stages = (int)((200 * 1024) / ((size_t)(BM * rs + BN) * BK * 4)) > 8 ? 8 : (int)((200 * 1024) / ((size_t)(BM * rs + BN) * BK * 4));
  p.stages = p.kblocks_per_split < max_stages ? p.kblocks_per_split : max_stages;
  // short K loops (wgrad: K = batch): 2 stages keep the CTA at ~64 KB so 3 CTAs share an SM and one
  // CTA's epilogue overlaps its neighbours' loads and MMAs
  if (p.kblocks_per_split <= 8) p.stages = 2;
  if (p.stages < 2) p.stages = 2;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.out = out; p.ldo = ldo; p.ws = ws; p.ep = ep;
  if (p.splits > 1 && !ws) {
    set_error("tf32 gemm: split-K needs a workspace");
    return kInvalidArg;
  }
  const bool epi = p.splits == 1 && (ep.bias || ep.relu || ep.dropout || ep.gate);
  if (rs == 2) {
    if (BN == 128) return epi ? launch_bn<128, true, 2>(tmA, tmB, p, stream) : launch_bn<128, false, 2>(tmA, tmB, p, stream);
    return epi ? launch_bn<64, true, 2>(tmA, tmB, p, stream) : launch_bn<64, false, 2>(tmA, tmB, p, stream);
  }
  if (BN == 128) return epi ? launch_bn<128, true, 1>(tmA, tmB, p, stream) : launch_bn<128, false, 1>(tmA, tmB, p, stream);
  return epi ? launch_bn<64, true, 1>(tmA, tmB, p, stream) : launch_bn<64, false, 1>(tmA, tmB, p, stream);
}

}  // namespace b200
