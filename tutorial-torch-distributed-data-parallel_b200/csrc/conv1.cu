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
constexpr int kC1GatherWarps = 8;                  // two groups of 4: each stage's kernel rows are split between them
constexpr int kC1Threads = 32 * (2 + kC1GatherWarps + 4);   // TMA, MMA, gather warps, 4 epilogue warps
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
      mbar_init(&in_empty[s], 32 * kC1GatherWarps);
      mbar_init(&a_full[s], 32 * kC1GatherWarps);
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
          for (int k = 0; k < kC1StageCols / 8; ++k) {
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
  } else if (warp < 2 + kC1GatherWarps) {
    // ---------------- gather: thread <-> output pixel of the tile; smem slab -> TMEM ----------------
    const int q = warp & 3;                                            // TMEM lane quarter of this warp
    const int grp = (warp - 2) >> 2;                                   // group 0: kernel rows 0..5, group 1: rows 6..10 + padding
    const int kh0 = grp * 6, kh1 = grp ? kC1KH : 6;
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
        for (int r6 = 0; r6 < 6; ++r6) {
          const int kh = kh0 + r6;
          if (kh >= kh1) break;
          const uint32_t a = row0 + (uint32_t)kh * pitch;
          const float4 v0 = lds128(a), v1 = lds128(a + 16), v2 = lds128(a + 32);
          const float v3 = lds32(a + 48);
          // taps kw = 0..10 are words 2..12 of the four loads; the 12th column of the row is zero
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols, rtf32(v0.z), rtf32(v0.w), rtf32(v1.x), rtf32(v1.y));
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols + 4u, rtf32(v1.z), rtf32(v1.w), rtf32(v2.x), rtf32(v2.y));
          tmem_st_x4(t_st + (uint32_t)kh * kC1RowCols + 8u, rtf32(v2.z), rtf32(v2.w), rtf32(v3), 0u);
        }
        if (grp) tmem_st_x4(t_st + (uint32_t)(kC1KH * kC1RowCols), 0u, 0u, 0u, 0u);
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
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * kC1BN + c), r);
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


// ---------------------------------------------------------------------------------------------------------------
// Weight gradient of the same layer:  dW[co][ci][kh][kw] = sum over (n, oh, ow) of dY[n][co][oh][ow] * x[n][ci][4oh-2+kh][4ow-2+kw]
// (reference multi-GPU-training-torch.py:125, loss.backward()). As a GEMM:  D[k'][co] = sum_pixel A[k'][pixel] * B[pixel][co].
//   * same persistent CTAs, same TMA-staged image slab (two output rows per tile) as the forward kernel;
//   * A = the im2col matrix TRANSPOSED, built in tensor memory: TMEM lane = reduction-index row k' (4 blocks of 128 lanes
//     cover K' = 408), TMEM column = pixel of the tile (2 x 56: 55 pixels + one zero per output row). A thread owns one
//     k' = (ci, kh, kw) and walks the pixels: consecutive lanes are consecutive kw, i.e. consecutive words of the slab;
//   * B = the dY tile [64 channels][112 pixel columns], K-major SWIZZLE_128B in shared memory. dY cannot come through
//     TMA (its row and plane strides, 220 and 12,100 bytes, are not multiples of 16), so four loader warps read it with
//     coalesced asynchronous 4-byte copies (lane = pixel) straight into the swizzled layout (the tensor core truncates
//     this operand to TF32);
//   * the four 128 x 64 accumulators stay in TMEM across ALL tiles of a CTA; each CTA writes one partial [408][64] at the
//     end and conv1_wgrad_finalize_kernel adds the partials in a fixed order (deterministic) into torch's layout.
constexpr int kWgCols = 112;                       // pixel columns per tile: 2 rows x (55 + 1 zero)
constexpr int kWgRowCols = 56;
constexpr int kWgMTiles = 4;                       // 4 x 128 lanes >= 408 reduction rows
constexpr uint32_t kWgACol0 = 256;                 // accumulators [0, 256), two A stages of 112 columns from 256
constexpr int kWgThreads = 32 * 14;                // TMA, MMA, 8 gather warps (2 groups), 4 dY loader warps
constexpr uint32_t kWgBBytes = 4 * kC1BN * 128;    // dY tile: 4 k-blocks of 32 pixel columns x 64 channels

struct Conv1WgradParams {
  const float* dy;
  float* partial;       // [gridDim.x][408][64]
  int Nb, H, W, OH, OW, pad, in_cols, tiles_per_image, tiles;
};

__global__ void __launch_bounds__(kWgThreads, 1)
conv1_wgrad_kernel(const __grid_constant__ CUtensorMap tmX, Conv1WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t in_bytes = (uint32_t)(kC1Cin * kC1InRows * p.in_cols * 4);
  const uint32_t in_stride = (in_bytes + 1023u) & ~1023u;
  uint8_t* s_b = smem;                                // [2][kWgBBytes]
  uint8_t* s_in = smem + 2 * kWgBBytes;               // [2][in_stride]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_in + 2 * in_stride);
  uint64_t* in_full = bars;           // [2] TMA -> gather
  uint64_t* in_empty = bars + 2;      // [2] gather (256 threads) -> TMA
  uint64_t* b_full = bars + 4;        // [2] loaders (128 threads) -> MMA
  uint64_t* b_empty = bars + 6;       // [2] MMA -> loaders
  uint64_t* a_full = bars + 8;        // [2] gather group (128 threads) -> MMA
  uint64_t* a_empty = bars + 10;      // [2] MMA -> gather group
  uint64_t* d_done = bars + 12;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], 256);
      mbar_init(&b_full[s], 128);
      mbar_init(&b_empty[s], 1);
      mbar_init(&a_full[s], 128);
      mbar_init(&a_empty[s], 1);
    }
    mbar_init(d_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = ((int)blockIdx.x < p.tiles) ? (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < my_tiles; ++i) {
        const int tile = (int)blockIdx.x + i * (int)gridDim.x;
        const int n = tile / p.tiles_per_image, t = tile - n * p.tiles_per_image;
        const int s = i & 1;
        mbar_wait(&in_empty[s], (((uint32_t)i >> 1) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&in_full[s], in_bytes);
        tma_load_3d(s_in + s * in_stride, &tmX, &in_full[s], -kC1XOff, t * kC1TileRows * kC1Stride - p.pad, n * kC1Cin);
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: per tile, M-tile mt uses A stage mt & 1 (written by gather group mt & 1) ----------------
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kC1BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t b_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    uint32_t ia[2] = {0, 0};                                           // uses of each A stage so far
    for (int i = 0; i < my_tiles; ++i) {
      const int tb = i & 1;
      mbar_wait(&b_full[tb], ((uint32_t)i >> 1) & 1u);
      fence_proxy_async_smem();                                        // the tile was written by cp.async (generic proxy)
      const uint32_t b_saddr = smem_u32(s_b + tb * kWgBBytes);
      for (int mt = 0; mt < kWgMTiles; ++mt) {
        const int st = mt & 1;
        mbar_wait(&a_full[st], ia[st] & 1u);
        ++ia[st];
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int c8 = 0; c8 < kWgCols / 8; ++c8) {
            const uint32_t b_addr = b_saddr + (uint32_t)(c8 >> 2) * (kC1BN * 128) + (uint32_t)(c8 & 3) * 32u;
            const uint32_t b_lo = ((b_addr & 0x3FFFFu) >> 4) | ((16u >> 4) << 16);
            umma_tf32_ts(tmem_base + (uint32_t)mt * kC1BN, tmem_base + kWgACol0 + (uint32_t)st * kWgCols + (uint32_t)c8 * 8u, b_lo,
                         b_hi, idesc, (i | c8) ? 1u : 0u);
          }
          umma_commit(&a_empty[st]);
          if (mt == kWgMTiles - 1) umma_commit(&b_empty[tb]);
        }
        __syncwarp();
      }
    }
    if (elect_one()) umma_commit(d_done);
    __syncwarp();
  } else if (warp < 10) {
    // ---------------- A gather: thread <-> reduction row k' of M-tile mt; slab -> TMEM, 8 pixel columns per store ----------------
    const int q = warp & 3, grp = (warp - 2) >> 2;
    const uint32_t pitch = (uint32_t)p.in_cols * 4u;
    uint32_t ia = 0;
    for (int i = 0; i < my_tiles; ++i) {
      const int s = i & 1;
      mbar_wait(&in_full[s], ((uint32_t)i >> 1) & 1u);
      const uint32_t slab = smem_u32(s_in + s * in_stride);
      for (int half = 0; half < 2; ++half) {
        const int mt = grp + 2 * half;
        const int k = mt * 128 + q * 32 + lane;
        const int ci = k / kC1StageCols, r = k - ci * kC1StageCols, kh = r / kC1RowCols, kw = r - kh * kC1RowCols;
        const bool valid = k < kC1K && r < kC1KH * kC1RowCols && kw < kC1KW;
        // word of the slab that multiplies pixel (row 0, ow 0) for this k'
        const uint32_t src = slab + ((uint32_t)(ci * kC1InRows + kh) * (uint32_t)p.in_cols + (uint32_t)(kC1Shift + kw)) * 4u;
        mbar_wait(&a_empty[grp], (ia & 1u) ^ 1u);
        ++ia;
        tc_fence_after();
        const uint32_t t_st = tmem_base + ((uint32_t)(q * 32) << 16) + kWgACol0 + (uint32_t)grp * kWgCols;
#pragma unroll
        for (int row = 0; row < kC1TileRows; ++row) {
          const uint32_t rsrc = src + (uint32_t)(row * kC1Stride) * pitch;
#pragma unroll
          for (int c8 = 0; c8 < kWgRowCols / 8; ++c8) {
            uint32_t v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int ow = c8 * 8 + j;
              v[j] = (valid && ow < p.OW) ? rtf32(lds32(rsrc + (uint32_t)ow * (kC1Stride * 4u))) : 0u;
            }
            tmem_st_x8(t_st + (uint32_t)(row * kWgRowCols + c8 * 8), v);
          }
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&a_full[grp]);
      }
      mbar_arrive(&in_empty[s]);
    }
    // ---------------- epilogue: this CTA's partial sums, TMEM -> workspace [k'][co] ----------------
    mbar_wait(d_done, 0);
    tc_fence_after();
    float* part = p.partial + (size_t)blockIdx.x * kC1K * kC1BN;
    for (int half = 0; half < 2; ++half) {
      const int mt = grp + 2 * half;
      const int k = mt * 128 + q * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < kC1BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * kC1BN + c), r);
        if (k < kC1K && my_tiles > 0) {
          float4* dst = reinterpret_cast<float4*>(part + (size_t)k * kC1BN + c);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                 __uint_as_float(r[4 * j + 3]));
        }
      }
    }
  } else {
    // ---------------- dY loaders: coalesced global loads (lane = pixel column) -> swizzled K-major tile ----------------
    const int lw = warp - 10;                                          // 16 channels per warp
    const long long ohw = (long long)p.OH * p.OW;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int n = tile / p.tiles_per_image, t = tile - n * p.tiles_per_image;
      const int oh0 = t * kC1TileRows;
      const int tb = i & 1;
      mbar_wait(&b_empty[tb], (((uint32_t)i >> 1) & 1u) ^ 1u);
      const uint32_t bb = smem_u32(s_b + tb * kWgBBytes);
      const float* dyn = p.dy + (long long)n * kC1BN * ohw;
      // asynchronous 4-byte copies global -> shared (dY rows are only 4-byte aligned), zero-filled where the column is
      // padding or beyond the image; the barrier is signalled by the copy unit when this thread's copies have landed,
      // so the loader goes straight on to the next tile and ~8 KB per warp are in flight without holding registers
#pragma unroll 4
      for (int u = 0; u < 16; ++u) {
        const int co = lw * 16 + u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = j * 32 + lane;                                 // pixel column of the tile
          const int row = c >= kWgRowCols ? 1 : 0, ow = c - row * kWgRowCols;
          const bool ok = c < kWgCols && ow < p.OW && oh0 + row < p.OH;
          const float* src = ok ? dyn + (long long)co * ohw + (long long)(oh0 + row) * p.OW + ow : dyn;
          const uint32_t a = bb + (uint32_t)j * (kC1BN * 128) + (uint32_t)co * 128u +
                             ((((uint32_t)lane >> 2) ^ ((uint32_t)co & 7u)) << 4) + ((uint32_t)lane & 3u) * 4u;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(a), "l"(src), "r"(ok ? 4 : 0) : "memory");
        }
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&b_full[tb])) : "memory");

    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// dW[co][ci][kh][kw] = sum over CTAs of partial[cta][ci*136 + kh*12 + kw][co], in a fixed order: a warp owns 8
// consecutive output channels of one (ci, kh, kw); its four lane groups add every fourth partial, then the four sums are
// combined by two shuffle steps (deterministic).
__global__ void __launch_bounds__(256) conv1_wgrad_finalize_kernel(const float* __restrict__ partial, int nparts,
                                                                   float* __restrict__ dw) {
  const int total8 = kC1Cin * kC1KH * kC1KW * (kC1BN / 8);
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= total8) return;
  const int kr = warp / (kC1BN / 8), co = (warp - kr * (kC1BN / 8)) * 8 + (lane & 7), zp = lane >> 3;
  const int ci = kr / (kC1KH * kC1KW), rem = kr - ci * (kC1KH * kC1KW), kh = rem / kC1KW, kw = rem - kh * kC1KW;
  const float* src = partial + (size_t)(ci * kC1StageCols + kh * kC1RowCols + kw) * kC1BN + co;
  float acc = 0.f;
  for (int z = zp; z < nparts; z += 4) acc += __ldg(src + (size_t)z * kC1K * kC1BN);
  acc += __shfl_xor_sync(0xffffffffu, acc, 8);
  acc += __shfl_xor_sync(0xffffffffu, acc, 16);
  if (zp == 0) dw[(size_t)co * (kC1Cin * kC1KH * kC1KW) + kr] = acc;
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

namespace b200 {

long long conv1_wgrad_workspace_bytes(int Nb, int H, int W, int pad) {
  const int OH = (H + 2 * pad - kC1KH) / kC1Stride + 1;
  long long tiles = (long long)((OH + kC1TileRows - 1) / kC1TileRows) * Nb;
  long long grid = sm_budget();
  if (grid > tiles) grid = tiles;
  return grid * kC1K * kC1BN * 4;
}

// Host side of k_order 2 of b200ddp_conv2d_wgrad.
int conv1_wgrad_launch(const float* x, const float* dy, float* dw, int Nb, int Cin, int H, int W, int Cout, int KH, int KW,
                       int stride, int pad, void* ws, long long ws_bytes, cudaStream_t stream) {
  B200_CHECK_ARG(Cin == kC1Cin && KH == kC1KH && KW == kC1KW && stride == kC1Stride && Cout == kC1BN && pad == kC1XOff - kC1Shift,
                 "conv2d_wgrad k_order 2 is the first-layer kernel: Cin 3, 11x11, stride 4, pad 2, Cout 64");
  B200_CHECK_ARG(((uintptr_t)x & 15) == 0 && (W & 3) == 0, "conv2d_wgrad k_order 2: x must be 16-byte aligned with W %% 4 == 0");
  Conv1WgradParams p;
  p.dy = dy; p.Nb = Nb; p.H = H; p.W = W; p.pad = pad;
  p.OH = (H + 2 * pad - KH) / stride + 1;
  p.OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(p.OH > 0 && p.OW > 0 && p.OW < kWgRowCols, "conv2d_wgrad k_order 2: output rows of at most %d pixels (got %d)",
                 kWgRowCols - 1, p.OW);
  p.in_cols = (kC1Stride * (p.OW - 1) + kC1Shift + kC1KW + 3) / 4 * 4;
  // A gather warp reads, per pixel, 11 + 11 + 8 consecutive words of three consecutive slab rows (lanes = 12-padded kernel
  // rows): with a row pitch of 12 (mod 32) words the three runs fall on disjoint banks; any other pitch costs a second
  // wavefront per load (ncu: the shared-memory pipe is the busiest unit of this kernel)
  while ((p.in_cols & 31) != 12) p.in_cols += 4;
  B200_CHECK_ARG(p.in_cols <= 256, "conv2d_wgrad k_order 2: input rows too wide");
  p.tiles_per_image = (p.OH + kC1TileRows - 1) / kC1TileRows;
  p.tiles = p.tiles_per_image * Nb;
  int grid = sm_budget();
  if (grid > p.tiles) grid = p.tiles;
  const long long need = (long long)grid * kC1K * kC1BN * 4;
  B200_CHECK_ARG(ws && ws_bytes >= need && ((uintptr_t)ws & 15) == 0, "conv2d_wgrad k_order 2: workspace of %lld bytes needed (got %lld)",
                 need, ws_bytes);
  p.partial = reinterpret_cast<float*>(ws);
  CUtensorMap tmX;
  const long long gdim[3] = {W, H, (long long)Nb * Cin};
  const long long gstr[2] = {(long long)W * 4, (long long)W * H * 4};
  const int box[3] = {p.in_cols, kC1InRows, kC1Cin};
  int rc = tma_map_nd(&tmX, x, 3, gdim, gstr, box);
  if (rc) return rc;
  const uint32_t in_bytes = (uint32_t)(kC1Cin * kC1InRows * p.in_cols * 4);
  const size_t smem = 1024 + 2 * (size_t)kWgBBytes + 2 * (size_t)((in_bytes + 1023u) & ~1023u) + 13 * 8 + 16;
  static std::atomic<unsigned long long> attr_devices{0};
  cudaError_t e = ensure_max_dyn_smem(conv1_wgrad_kernel, attr_devices);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(conv1_wgrad_kernel) -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  conv1_wgrad_kernel<<<grid, kWgThreads, smem, stream>>>(tmX, p);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv1_wgrad_kernel launch (grid %d smem %zu) -> %s", grid, smem, cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  const int total = kC1Cin * kC1KH * kC1KW * kC1BN;
  conv1_wgrad_finalize_kernel<<<(total * 4 + 255) / 256, 256, 0, stream>>>(p.partial, grid, dw);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv1_wgrad_finalize_kernel launch -> %s", cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

}  // namespace b200
