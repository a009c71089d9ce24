// C1 — the first convolution (torchvision alexnet.py:22: Conv2d(3, 64, kernel_size=11, stride=4, padding=2)) as its own
// kernel. Reached from reference multi-GPU-training-torch.py:121.
//
// Why: in the generic implicit-GEMM kernel (conv.cu) the im2col gather of a stride-4 convolution reads, per warp load,
// 32 pixels that sit 16 bytes apart — four 128-byte lines for 128 useful bytes — and it does so 363 times per pixel;
// the layer ran at 13 % of the TF32 rate with the L1 wavefront pipe 70-80 % busy (profiles/r01_summary.md).
// The layer has only 3 input channels: everything a tile of output pixels needs is a thin slab of the image, so
//   * a PERSISTENT CTA (one per SM) keeps the whole weight matrix (64 x 408 TF32, 104 KB) in shared memory,
//   * TMA brings the slab of the image for two output rows (3 channels x 15 input rows x 228 padded columns, 41 KB,
//     out-of-image elements zero-filled by the TMA unit: no bounds tests anywhere) into a double buffer,
//   * the gather becomes shared-memory reads: a thread (= output pixel) reads the 11 taps of a kernel row as three
//     aligned 16-byte loads + one word — consecutive pixels are 16 bytes apart, so a quarter-warp reads 128 contiguous
//     bytes, conflict-free (the slab starts 4 columns left of the image, not `pad` = 2: a TMA tile must start at a
//     16-byte multiple of the innermost dimension — an odd start raised "illegal instruction" on B200 — so a pixel's
//     window begins 2 words into its first 16-byte load) — rounds to TF32 and writes them to TENSOR MEMORY (tcgen05.st); the MMA is the TS form
//     tcgen05.mma [d], [a_tmem], b_desc, as in conv.cu,
//   * two accumulators in TMEM let the epilogue of tile t (separate warps) overlap the gather and MMAs of tile t+1.
// Reduction index: k' = ci * 136 + kh * 12 + kw  (kw < 11; column 11 of each row and the last 4 columns of a channel
// are zero in both operands), K' = 408 = 51 MMAs of K = 8. The matching weight matrix comes from conv_weight_prep
// mode 5 (conv.cu).
//
// Geometry limits (checked by the host code): stride 4, KW <= 12, Cin * KH rows fit the staging buffers, W + 2 pad <= 256.
#include "gemm.h"
#include "tc.cuh"
#include "../../include/b200ddp.h"

#include <stdlib.h>

namespace b200 {
namespace {

using namespace tc;

constexpr int kC1Cin = 3, kC1KH = 11, kC1KW = 11, kC1Stride = 4;
constexpr int kC1RowCols = 12;                     // taps per kernel row in the padded K order (11 real + 1 zero)
constexpr int kC1StageCols = 136;                  // one input channel: 11 rows x 12 + 4 zero columns = 17 MMAs
constexpr int kC1K = kC1Cin * kC1StageCols;        // 408
constexpr int kC1KBlocks = (kC1K + 31) / 32;       // 13 weight k-blocks of 32 (the last one uses 24 columns)
constexpr int kC1BN = 64;
constexpr int kC1TileRows = 2;                     // output rows per tile
constexpr int kC1InRows = (kC1TileRows - 1) * kC1Stride + kC1KH;   // 15 input rows
constexpr int kC1Threads = 32 * 10;                // TMA, MMA, 4 gather warps, 4 epilogue warps
constexpr uint32_t kC1TmemCols = 512;
constexpr int kC1XOff = 4;                         // the staged slab starts at image column -4 (TMA: 16-byte multiple)
constexpr int kC1Shift = 2;                        // = kC1XOff - pad: a pixel's window starts this many words into its first load
constexpr uint32_t kC1ACol0 = 128;                 // accumulators in [0,64) and [64,128); A stages from 128

struct Conv1Params {
  const float* bias;
  float* y;
  int Nb, H, W, OH, OW, pad, relu;
  int in_cols;          // W + 2 pad rounded up to a multiple of 4 (row pitch of the staged slab, floats)
  int tiles_per_image, tiles;
  int dbg;              // debugging switches (B200DDP_C1_DBG): 1 no MMA, 2 no tcgen05.st, 4 no tcgen05.ld, 8 no x-TMA
};

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st_x4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ float lds32(uint32_t saddr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint32_t rtf32(float v) { return __float_as_uint(v) + 0x1000u; }   // round to TF32 (see conv.cu)

__global__ void __launch_bounds__(kC1Threads, 1)
conv1_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, Conv1Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t W_BYTES = kC1KBlocks * kC1BN * 32 * 4;                 // 13 x 8 KB
  const uint32_t in_bytes = (uint32_t)(kC1Cin * kC1InRows * p.in_cols * 4);   // one staged slab
  const uint32_t in_stride = (in_bytes + 1023u) & ~1023u;
  uint8_t* s_w = smem;
  uint8_t* s_in = smem + W_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_in + 2 * in_stride);
  uint64_t* w_full = bars;            // weights resident
  uint64_t* in_full = bars + 1;       // [2] TMA -> gather
  uint64_t* in_empty = bars + 3;      // [2] gather -> TMA
  uint64_t* a_full = bars + 5;        // [2] gather -> MMA
  uint64_t* a_empty = bars + 7;       // [2] MMA -> gather
  uint64_t* d_full = bars + 9;        // [2] MMA -> epilogue
  uint64_t* d_empty = bars + 11;      // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(w_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], 128);
      mbar_init(&a_full[s], 128);
      mbar_init(&a_empty[s], 1);
      mbar_init(&d_full[s], 1);
      mbar_init(&d_empty[s], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, kC1TmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = ((int)blockIdx.x < p.tiles) ? (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    if (lane == 0) {
      // ---------------- TMA: the weight matrix once, then one image slab per tile ----------------
      mbar_arrive_expect_tx(w_full, W_BYTES);
      for (int kb = 0; kb < kC1KBlocks; ++kb) tma_load_2d(s_w + kb * (kC1BN * 128), &tmW, w_full, kb * 32, 0);
      for (int i = 0; i < my_tiles; ++i) {
        const int tile = (int)blockIdx.x + i * (int)gridDim.x;
        const int n = tile / p.tiles_per_image, t = tile - n * p.tiles_per_image;
        const int s = i & 1;
        mbar_wait(&in_empty[s], (((uint32_t)i >> 1) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&in_full[s], in_bytes);
        // box {in_cols, 15 rows, 3 channels} starting at column -4, row 4*oh0 - pad of image n: the TMA unit fills
        // everything outside the image with zeros
        tma_load_3d(s_in + s * in_stride, &tmX, &in_full[s], -kC1XOff, t * kC1TileRows * kC1Stride - p.pad, n * kC1Cin);
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kC1BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t b_hi = (1024u >> 4) | (1u << 14) | (2u << 29);      // SBO 1024, version 1, SWIZZLE_128B
    const uint32_t w_saddr = smem_u32(s_w);
    mbar_wait(w_full, 0);
    uint32_t it = 0;                                                   // A-stage counter (3 per tile)
    for (int i = 0; i < my_tiles; ++i) {
      const int acc = i & 1;
      mbar_wait(&d_empty[acc], (((uint32_t)i >> 1) & 1u) ^ 1u);
      tc_fence_after();
      for (int ci = 0; ci < kC1Cin; ++ci, ++it) {
        const int st = (int)(it & 1u);
        mbar_wait(&a_full[st], (it >> 1) & 1u);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < ((p.dbg & 1) ? 0 : kC1StageCols / 8); ++k) {
            const int kk = ci * kC1StageCols + k * 8;                  // global k' of this MMA
            const uint32_t b_addr = w_saddr + (uint32_t)(kk >> 5) * (kC1BN * 128) + (uint32_t)(kk & 31) * 4u;
            const uint32_t b_lo = ((b_addr & 0x3FFFFu) >> 4) | ((16u >> 4) << 16);
            umma_tf32_ts(tmem_base + (uint32_t)acc * kC1BN, tmem_base + kC1ACol0 + (uint32_t)st * kC1StageCols + (uint32_t)k * 8u,
                         b_lo, b_hi, idesc, (ci | k) ? 1u : 0u);
          }
          umma_commit(&a_empty[st]);
          if (ci == kC1Cin - 1) umma_commit(&d_full[acc]);
        }
        __syncwarp();
      }
    }
  } else if (warp < 6) {
    // ---------------- gather: thread <-> output pixel of the tile; smem slab -> TMEM ----------------
    const int q = warp & 3;                                            // TMEM lane quarter of this warp
    const int pix = q * 32 + lane;                                     // pixel of the tile (two output rows)
    const int r_local = pix >= p.OW ? 1 : 0;
    int ow = pix - r_local * p.OW;
    if (pix >= kC1TileRows * p.OW) ow = 0;                             // idle lanes read valid data, results unused
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + kC1ACol0;
    const uint32_t pitch = (uint32_t)p.in_cols * 4u;
    const uint32_t px_off = ((uint32_t)(r_local * kC1Stride) * (uint32_t)p.in_cols + (uint32_t)(ow * kC1Stride)) * 4u;
    uint32_t it = 0;
    for (int i = 0; i < my_tiles; ++i) {
      const int s = i & 1;
      mbar_wait(&in_full[s], ((uint32_t)i >> 1) & 1u);
      const uint32_t slab = smem_u32(s_in + s * in_stride) + px_off;
      for (int ci = 0; ci < kC1Cin; ++ci, ++it) {
        const int st = (int)(it & 1u);
        mbar_wait(&a_empty[st], ((it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t t_st = t_lane + (uint32_t)st * kC1StageCols;
        const uint32_t row0 = slab + (uint32_t)(ci * kC1InRows) * pitch;
#pragma unroll
        for (int kh = 0; kh < ((p.dbg & 2) ? 0 : kC1KH); ++kh) {
          const uint32_t a = row0 + (uint32_t)kh * pitch;
          const float4 v0 = lds128(a), v1 = lds128(a + 16), v2 = lds128(a + 32);
          const float v3 = lds32(a + 48);
          // taps kw = 0..10 are words 2..12 of the four loads; the 12th column of the row is zero
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols, rtf32(v0.z), rtf32(v0.w), rtf32(v1.x), rtf32(v1.y));
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols + 4u, rtf32(v1.z), rtf32(v1.w), rtf32(v2.x), rtf32(v2.y));
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols + 8u, rtf32(v2.z), rtf32(v2.w), rtf32(v3), 0u);
        }
        if (!(p.dbg & 2)) tmem_st_x4(t_st + (uint32_t)(kC1KH * kC1RowCols), 0u, 0u, 0u, 0u);
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&a_full[st]);
      }
      mbar_arrive(&in_empty[s]);                                       // this thread is done with the slab
    }
  } else {
    // ---------------- epilogue: TMEM -> + bias (ReLU) -> y, coalesced along the output row ----------------
    const int q = warp & 3;
    const int pix = q * 32 + lane;
    const int r_local = pix >= p.OW ? 1 : 0;
    const int ow = pix - r_local * p.OW;
    const long long ohw = (long long)p.OH * p.OW;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int n = tile / p.tiles_per_image, t = tile - n * p.tiles_per_image;
      const int oh = t * kC1TileRows + r_local;
      const bool valid = pix < kC1TileRows * p.OW && oh < p.OH;
      const int acc = i & 1;
      mbar_wait(&d_full[acc], ((uint32_t)i >> 1) & 1u);
      tc_fence_after();
      float* yb = p.y + (long long)n * kC1BN * ohw + (long long)oh * p.OW + ow;
#pragma unroll 1
      for (int c = 0; c < kC1BN; c += 32) {
        uint32_t r[32] = {};
        if (!(p.dbg & 4)) tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * kC1BN + c), r);
        if (valid) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float v = __uint_as_float(r[j]);
            if (p.bias) v += __ldg(p.bias + c + j);
            if (p.relu) v = v > 0.f ? v : 0.f;
            yb[(long long)(c + j) * ohw] = v;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&d_empty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kC1TmemCols);
}

}  // namespace
}  // namespace b200

namespace b200 {

// Host side of k_order 2 of b200ddp_conv2d_fwd (see include/b200ddp.h).
int conv1_fwd_launch(const float* x, const float* wmat, long long ldw, const float* bias, float* y, int Nb, int Cin, int H,
                     int W, int Cout, int KH, int KW, int stride, int pad, int relu, cudaStream_t stream) {
  B200_CHECK_ARG(Cin == kC1Cin && KH == kC1KH && KW == kC1KW && stride == kC1Stride && Cout == kC1BN,
                 "conv2d_fwd k_order 2 is the first-layer kernel: Cin 3, 11x11, stride 4, Cout 64 (got Cin %d, %dx%d, stride %d, Cout %d)",
                 Cin, KH, KW, stride, Cout);
  B200_CHECK_ARG(pad == kC1XOff - kC1Shift, "conv2d_fwd k_order 2: padding must be %d (got %d)", kC1XOff - kC1Shift, pad);
  const int OWq = (W + 2 * pad - KW) / stride + 1;
  const int in_cols = (kC1Stride * (OWq - 1) + kC1Shift + kC1KW + 3) / 4 * 4;      // last pixel's window, rounded to 16 bytes
  B200_CHECK_ARG(in_cols <= 256 && OWq > 0, "conv2d_fwd k_order 2: input rows of at most ~250 pixels (got %d)", W);
  B200_CHECK_ARG(ldw >= kC1KBlocks * 32 && (ldw & 3) == 0, "conv2d_fwd k_order 2: weight matrix needs ld >= %d (weight_prep mode 5)",
                 kC1KBlocks * 32);
  B200_CHECK_ARG(((uintptr_t)x & 15) == 0 && (W & 3) == 0, "conv2d_fwd k_order 2: x must be 16-byte aligned with W %% 4 == 0");
  Conv1Params p;
  p.bias = bias; p.y = y; p.Nb = Nb; p.H = H; p.W = W; p.pad = pad; p.relu = relu;
  p.OH = (H + 2 * pad - KH) / stride + 1;
  p.OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(p.OH > 0 && p.OW > 0 && kC1TileRows * p.OW <= 128, "conv2d_fwd k_order 2: output rows of at most 64 pixels (got %d)", p.OW);
  p.in_cols = in_cols;
  p.tiles_per_image = (p.OH + kC1TileRows - 1) / kC1TileRows;
  p.tiles = p.tiles_per_image * Nb;
  p.dbg = getenv("B200DDP_C1_DBG") ? atoi(getenv("B200DDP_C1_DBG")) : 0;
  CUtensorMap tmX, tmW;
  const long long gdim[3] = {W, H, (long long)Nb * Cin};
  const long long gstr[2] = {(long long)W * 4, (long long)W * H * 4};
  const int box[3] = {in_cols, kC1InRows, kC1Cin};
  int rc = tma_map_nd(&tmX, x, 3, gdim, gstr, box);
  if (rc) return rc;
  rc = tma_map_2d(&tmW, wmat, ldw, Cout, ldw, 32, kC1BN, 0);
  if (rc) return rc;
  const uint32_t in_bytes = (uint32_t)(kC1Cin * kC1InRows * in_cols * 4);
  const size_t smem = 1024 + (size_t)kC1KBlocks * kC1BN * 128 + 2 * (size_t)((in_bytes + 1023u) & ~1023u) + 13 * 8 + 16;
  static std::atomic<unsigned long long> attr_devices{0};
  cudaError_t e = ensure_max_dyn_smem(conv1_fwd_kernel, attr_devices);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(conv1_fwd_kernel) -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  int grid = sm_budget();
  if (grid > p.tiles) grid = p.tiles;
  conv1_fwd_kernel<<<grid, kC1Threads, smem, stream>>>(tmX, tmW, p);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv1_fwd_kernel launch (grid %d smem %zu) -> %s", grid, smem, cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

}  // namespace b200
