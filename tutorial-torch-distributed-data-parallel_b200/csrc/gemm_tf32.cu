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

#include <cstdlib>
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
// Persistent variant for problems with MANY output tiles and a SHORT reduction (dW of the Linear layers: R = batch, so
// a tile is 4 k-blocks and 64 KB of output; FC1 has 2304 of them). One CTA per SM walks tiles blockIdx.x, +gridDim.x, ..:
// the TMA ring runs ahead across tile boundaries and two TMEM accumulators alternate, so the loads and MMAs of tile
// i+1 run while eight epilogue warps store tile i. In the one-tile-per-CTA kernel above those three phases are serial
// per CTA (profiles/r02_gemm_ncu_summary.md: half of all warp samples wait for another phase of the same CTA).
// Raw accumulator store only (no split-K, no fused epilogue).
// ------------------------------------------------------------------------------------------------
// EPIW epilogue warps = 4 TMEM lane quarters x EPIW/4 column parts; MINB CTAs per SM (register cap for the compiler).
template <int BN, int EPIW, int MINB>
__global__ void __launch_bounds__(64 + 32 * EPIW, MINB)
tf32_gemm_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, KParams p,
                            int m_tiles, int tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = BM * BK * 4;
  constexpr uint32_t B_BYTES = BN * BK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* d_full = empty_bar + stages;         // [2] MMA -> epilogue
  uint64_t* d_empty = d_full + 2;                // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (p.R + BK - 1) / BK;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&d_full[a], 1);
      mbar_init(&d_empty[a], 32 * EPIW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = ((int)blockIdx.x < tiles) ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    if (lane == 0) {
      // ---------------- TMA producer: the ring does not know about tile boundaries ----------------
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const int tile = (int)blockIdx.x + i * (int)gridDim.x;
        const int nt = tile / m_tiles, mt = tile - nt * m_tiles;   // neighbouring CTAs share the B tile in L2
        const int m0 = mt * BM, n0 = nt * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          mbar_arrive_expect_tx(&full_bar[s], STAGE_BYTES);
          uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const int r0 = kb * BK;
          if (!p.a_mn) {
            tma_load_2d(sa, &tmA, &full_bar[s], r0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 32; ++j) tma_load_2d(sa + j * 4096, &tmA, &full_bar[s], m0 + 32 * j, r0);
          }
          if (!p.b_mn) {
            tma_load_2d(sb, &tmB, &full_bar[s], r0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 32; ++j) tma_load_2d(sb + j * 4096, &tmB, &full_bar[s], n0 + 32 * j, r0);
          }
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: all 32 lanes walk the loop, one elected lane issues ----------------
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16)
                           | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    const OperandDesc a = p.a_mn ? operand_mn_major() : operand_k_major();
    const OperandDesc b = p.b_mn ? operand_mn_major() : operand_k_major();
    const uint32_t smem_saddr = smem_u32(smem);
    const uint32_t a_lo0 = ((smem_saddr & 0x3FFFFu) >> 4) | ((a.lbo >> 4) << 16);
    const uint32_t b_lo0 = (((smem_saddr + A_BYTES) & 0x3FFFFu) >> 4) | ((b.lbo >> 4) << 16);
    const uint32_t a_hi = (a.sbo >> 4) | (1u << 14) | (a.layout_type << 29);
    const uint32_t b_hi = (b.sbo >> 4) | (1u << 14) | (b.layout_type << 29);
    const uint32_t a_step = a.k_advance >> 4, b_step = b.k_advance >> 4, stage_step = STAGE_BYTES >> 4;
    const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]);
    uint32_t a_lo = a_lo0, b_lo = b_lo0, fb = full0, eb = empty0, ph = 0;
    int s = 0;
#pragma unroll 1
    for (int i = 0; i < my_tiles; ++i) {
      const int acc = i & 1;
      mbar_wait(&d_empty[acc], (((uint32_t)i >> 1) & 1u) ^ 1u);      // the epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t d_addr = tmem_base + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait_addr(fb, ph);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_tf32_lohi(d_addr, a_lo + k * a_step, a_hi, b_lo + k * b_step, b_hi, idesc, (kb | k) ? 1u : 0u);
          umma_commit_addr(eb);
          if (kb == nkb - 1) umma_commit(&d_full[acc]);
        }
        __syncwarp();
        if (++s == stages) {
          s = 0; ph ^= 1u; a_lo = a_lo0; b_lo = b_lo0; fb = full0; eb = empty0;
        } else {
          a_lo += stage_step; b_lo += stage_step; fb += 8u; eb += 8u;
        }
      }
    }
  } else {
    // ---------------- epilogue: 4 TMEM lane quarters x EPIW/4 column parts ----------------
    constexpr int PART = BN / (EPIW / 4);                  // columns per warp: 64 or 32
    const int q = warp & 3;
    const int part = (warp - 2) >> 2;
#pragma unroll 1
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int nt = tile / m_tiles, mt = tile - nt * m_tiles;
      const int m = mt * BM + q * 32 + lane;
      const int n_base = nt * BN + part * PART;
      const int acc = i & 1;
      mbar_wait(&d_full[acc], ((uint32_t)i >> 1) & 1u);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < PART; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + part * PART + c), r);
        if (m < p.M) {
          float* dst = p.out + (long long)(n_base + c) * p.ldo + m;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n_base + c + j < p.N) dst[(long long)j * p.ldo] = __uint_as_float(r[j]);
        }
      }
      tc_fence_before();
      mbar_arrive(&d_empty[acc]);
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

// Whether the persistent kernel takes the short-reduction raw-store GEMMs. Default: when no SMs are set aside for a
// resident collective (b200ddp_set_sm_budget) - its CTAs stay on their SMs for the whole kernel, and that has only been
// measured next to the all-reduce kernel's SM-owning CTAs on one GPU. B200DDP_GEMM_PERSISTENT=1 / 0 forces it on / off
// (read per call: a benchmark flips it).
inline bool tf32_persistent_mode() {
  const char* e = getenv("B200DDP_GEMM_PERSISTENT");
  if (e && e[0]) return e[0] != '0';
  return sm_budget() >= num_sms();
}

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

template <int BN, int EPIW, int MINB>
int launch_persistent(const CUtensorMap& tmA, const CUtensorMap& tmB, const KParams& p, int m_tiles, int tiles, int grid,
                      cudaStream_t stream) {
  const size_t stage_bytes = (size_t)(BM + BN) * BK * 4;
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 4) * 8 + 16;
  static std::atomic<unsigned long long> attr_devices{0};
  cudaError_t e = ensure_max_dyn_smem(tf32_gemm_persistent_kernel<BN, EPIW, MINB>, attr_devices);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(tf32_gemm_persistent_kernel) -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  tf32_gemm_persistent_kernel<BN, EPIW, MINB><<<grid, 64 + 32 * EPIW, smem, stream>>>(tmA, tmB, p, m_tiles, tiles);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("tf32_gemm_persistent_kernel launch (grid %d smem %zu) -> %s", grid, smem, cudaGetErrorString(e));
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
  {
    // many tiles, short reduction, raw store: persistent CTAs (see tf32_gemm_persistent_kernel)
    const int m_tiles = (M + BM - 1) / BM, n_tiles = (N + 127) / 128;
    const long long tiles = (long long)m_tiles * n_tiles;
    const int sms = sm_budget() > 0 ? sm_budget() : num_sms();
    if (tf32_persistent_mode() && p.splits == 1 && !epi && BN == 128 && total_kb <= 8 && tiles >= 2ll * sms &&
        tiles < (1ll << 30)) {
      const char* st_env = getenv("B200DDP_GEMM_PERS_STAGES");
      int st = st_env ? atoi(st_env) : 3;
      if (st < 2) st = 2;
      if (st > 6) st = 6;
      p.stages = st;
      const int per_sm = st <= 3 ? 2 : 1;                 // 3 stages = 97 KB: two CTAs (2 x 256 TMEM columns) share an SM
      long long grid = (long long)sms * per_sm;
      if (grid > tiles) grid = tiles;
      const char* ew_env = getenv("B200DDP_GEMM_PERS_EPI");
      const int ew = ew_env ? atoi(ew_env) : 16;
      if (ew == 16)
        return per_sm == 2 ? launch_persistent<128, 16, 2>(tmA, tmB, p, m_tiles, (int)tiles, (int)grid, stream)
                           : launch_persistent<128, 16, 1>(tmA, tmB, p, m_tiles, (int)tiles, (int)grid, stream);
      return launch_persistent<128, 8, 2>(tmA, tmB, p, m_tiles, (int)tiles, (int)grid, stream);
    }
  }
  if (rs == 2) {
    if (BN == 128) return epi ? launch_bn<128, true, 2>(tmA, tmB, p, stream) : launch_bn<128, false, 2>(tmA, tmB, p, stream);
    return epi ? launch_bn<64, true, 2>(tmA, tmB, p, stream) : launch_bn<64, false, 2>(tmA, tmB, p, stream);
  }
  if (BN == 128) return epi ? launch_bn<128, true, 1>(tmA, tmB, p, stream) : launch_bn<128, false, 1>(tmA, tmB, p, stream);
  return epi ? launch_bn<64, true, 1>(tmA, tmB, p, stream) : launch_bn<64, false, 1>(tmA, tmB, p, stream);
}

}  // namespace b200
