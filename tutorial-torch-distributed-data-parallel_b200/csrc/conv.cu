// C1-C5 — the five convolutions of the feature extractor (torchvision alexnet.py:22,25,28,30,32, reached
// from reference multi-GPU-training-torch.py:121/:125) as implicit GEMMs on the 5th-generation tensor
// cores: TF32 operands read straight from the fp32 NCHW tensors, fp32 accumulation in TMEM.
// (The reference runs these through cuDNN, also in TF32 — torch's default for convolutions.)
//
// No im2col matrix is ever materialised and no NCHW<->NHWC conversion pass runs (the stock path spends
// ~1 ms per step in cudnn::nchwToNhwcKernel / nhwcToNchwKernel).
//
//   forward / dgrad   D[m = pixel][n = co] = sum_k  col(pixel, k) * W[co][k]          (conv_fwd_kernel)
//       A = im2col tile: producer warps gather it (thread <-> output pixel, 32 k values in registers), round to
//           TF32 and write their tile row into TENSOR MEMORY with one tcgen05.st; the MMA is the TS form
//           tcgen05.mma [d_tmem], [a_tmem], b_desc. B = W rows, K-major, TMA SWIZZLE_128B into a shared-memory
//           ring filled by a dedicated TMA warp. One warp issues the MMAs (warp-uniform loop, elect.sync).
//       Epilogue (the producer warps): tcgen05.ld, + bias[co], ReLU, coalesced stores y[n][co][pix].
//       dgrad of a stride-1 convolution is the same kernel run on dY with the flipped / transposed
//       weights produced by conv_weight_prep (pad' = KH-1-pad).
//   wgrad             D[m = k][n = co]     = sum_pixel col(pixel, k) * dY[co][pixel]   (conv_wgrad_kernel)
//       both operands K-major (32 consecutive pixels per smem row, SWIZZLE_128B), gathered by the producer
//       warps into shared memory (the TMEM form would need lane = k-row, i.e. uncoalesced loads); R row
//       sub-tiles per CTA share one gathered dY tile (R accumulators in TMEM); the pixel range is split over
//       CTAs (deterministic partials + finalize) because only ceil(K/(128 R)) x ceil(Cout/BN) tiles exist.
//
// Two orderings of the reduction index k (the weight matrix is permuted to match by conv_weight_prep):
//   k_order 0  k = (ci, kh, kw)   torch's own order; per-k offsets + window bits come from a device table. Any Cin.
//   k_order 1  k = (ci/32, kh, kw, ci%32)   "channel-block-major", needs Cin % 32 == 0: a whole 32-row k-block
//              shares one (kh, kw) - one bounds test per k-block, loads `base + r * plane` with the plane size a
//              template constant for the toy model's shapes - and the 32-channel slab of the tile's rows is
//              re-read by every tap, so it stays in L1.
//
// What bounds these kernels, as measured (profiles/r01_summary.md): the MMA-issuing warp's barrier round, then
// the instruction count of the gather, then the unified L1/shared-memory data pipe (hence A in TMEM), then the
// bytes gathered per MMA (hence R sub-tiles in wgrad), then the per-tile fixed cost for short reductions
// (hence the two-CTAs-per-SM shape for conv1).
#include "gemm.h"
#include "tc.cuh"
#include "../../include/b200ddp.h"

#include <mutex>

namespace b200 {

namespace {

using namespace tc;

constexpr int CBM = 128;            // UMMA_M
constexpr int CBK = 32;             // fp32 per k-block (128 bytes)
constexpr int CUK = 8;              // K per tcgen05.mma kind::tf32
constexpr int kWgGroups = 3;        // wgrad: 3 producer groups, batches of 32 loads software-pipelined in registers
constexpr int kWgProdWarps = 4 * kWgGroups;
constexpr int kWgThreads = 32 * (kWgProdWarps + 1);

// fp32 -> tf32 with rounding (half away from zero on the magnitude): the tensor core drops the low 13
// mantissa bits, so adding half an ulp of tf32 first turns its truncation into rounding.
__device__ __forceinline__ float round_tf32(float v) { return __uint_as_float(__float_as_uint(v) + 0x1000u); }

__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// 32 loads base + r*stride (r < nvalid, zeros beyond). PLANE > 0: the stride is PLANE*4 bytes and becomes the
// instruction's immediate offset; otherwise a 64-bit pointer walks by `stride4` (2 integer instructions per load).
template <int PLANE>
__device__ __forceinline__ void load_strided(const char* base, uint32_t stride4, int nvalid, float (&v)[32]) {
  if (PLANE > 0) {
    const float* q = reinterpret_cast<const float*>(base);
    if (nvalid >= 32) {
#pragma unroll
      for (int r = 0; r < 32; ++r) v[r] = __ldg(q + r * PLANE);
    } else {
#pragma unroll
      for (int r = 0; r < 32; ++r) {
        v[r] = 0.f;
        if (r < nvalid) v[r] = __ldg(q + r * PLANE);
      }
    }
  } else {
    const char* q = base;
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      v[r] = 0.f;
      if (r < nvalid) v[r] = __ldg(reinterpret_cast<const float*>(q));
      q += stride4;
    }
  }
}

template <int BN>
constexpr uint32_t tmem_cols() { return BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256; }

struct ConvFwdParams {
  const float* x;
  float* y;
  const float* bias;
  const int2* tab;     // k_order 0: [kblocks*32] {ci*H*W + kh*W + kw, 1 << kh | 1 << (16 + kw)}; {0, 1 << 31} beyond K
  int Nb, Cin, H, W, Cout, OH, OW, KH, KW, stride, pad;
  int kblocks, M, OHW, relu, stages;
  long long* trace;    // B200_CONV_TRACE builds only: [kblocks][8] clock64 stamps of CTA (0,0); null otherwise
};

// PL: -1 = table path (k_order 0); 0 = channel-inner path with a runtime channel stride; > 0 = channel-inner path
// with the plane size H*W known at compile time (the toy model's 27x27 and 13x13 planes), so that the 32 loads
// of a k-block are `LDG [base + immediate]` with no address arithmetic at all.
// Two shapes of CTA. Default: 3 producer groups, one CTA per SM, all 512 TMEM columns (accumulator + 8 A stages).
// TWO: 2 producer groups and 256 TMEM columns (64-column accumulator + 6 A stages) so that two CTAs share an SM:
// with short tiles the prologue / pipeline fill / drain / epilogue of a tile (~5 us) overlaps the other CTA's main loop.
template <bool TWO>
struct FwdShape {
  static constexpr int kGroupsT = TWO ? 2 : 3;
  static constexpr int kProdWarpsT = 4 * kGroupsT;
  static constexpr int kThreads = 32 * (kProdWarpsT + 2);   // + MMA warp (TMEM alloc) + weight-TMA warp
  static constexpr int kCtasPerSm = TWO ? 2 : 1;
  static constexpr uint32_t kTmemColsT = TWO ? 256 : 512;
  static constexpr uint32_t kATmemCol0T = TWO ? 64 : 256;
  static constexpr int kMaxStages = TWO ? 6 : 8;
};

// TWO (only with BN == 64): the two-CTAs-per-SM shape, chosen by the host for short reduction loops (conv1 forward:
// 18 k-blocks per tile); long loops (conv2 dgrad: 150) are faster with one CTA of three producer groups.
// KS = 32-wide k-blocks per ring stage (1 or 2). With a 64-channel tile the four MMAs of a k-block take 128 cycles but
// the MMA warp's barrier round (wait, fence, issue, commit) ~540: KS = 2 halves the rounds per reduction step.
template <int BN, int PL, bool TWO = false, int KS = 1>
__global__ void __launch_bounds__(FwdShape<TWO>::kThreads, FwdShape<TWO>::kCtasPerSm)
conv_fwd_kernel(const __grid_constant__ CUtensorMap tmW, ConvFwdParams p) {
  static_assert(!TWO || BN == 64, "the two-CTA shape has room for a 64-column accumulator only");
  static_assert(KS == 1 || KS == 2, "one or two k-blocks per stage");
  constexpr bool CI = PL >= 0;
  constexpr int kGroups = FwdShape<TWO>::kGroupsT, kProdWarps = FwdShape<TWO>::kProdWarpsT;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // The im2col tile (A) goes registers -> TENSOR MEMORY (tcgen05.st), not shared memory: ncu showed the unified
  // L1/shared-memory data pipe as the limiter (gather LDGs + 128 STS + TMA writes + the MMA's own reads of A and B
  // per k-block); with A in TMEM the STS and the MMA's A reads leave that pipe. Shared memory holds only the
  // TMA-fed weight tiles. TMEM: accumulator in columns [0, BN), A stage s in [kATmemCol0 + 32 s, +32).
  constexpr uint32_t B_BYTES = BN * CBK * 4;
  constexpr uint32_t STAGE_BYTES = KS * B_BYTES;
  constexpr uint32_t kATmemCol0 = FwdShape<TWO>::kATmemCol0T, kTmemCols = FwdShape<TWO>::kTmemColsT;
  const int stages = p.stages;                        // <= FwdShape<TWO>::kMaxStages / KS (32*KS A columns each)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * CBM, n0 = blockIdx.y * BN;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 128 + 1); // the 128 threads of one producer group + the TMA warp's arrive.expect_tx
      mbar_init(&empty_bar[s], 1);      // tcgen05.commit
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
  }
  if (warp == kProdWarps) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kProdWarps) {
    // ---------------- im2col gather producers: thread <-> output pixel, group <-> every kGroups-th k-block ----------------
    const int group = warp >> 2, gwarp = warp & 3, gtid = tid & 127;
    const int m = m0 + gtid;
    const bool valid_m = m < p.M;
    const int mm = valid_m ? m : 0;
    const int n_img = mm / p.OHW, pix = mm - n_img * p.OHW;
    const int oh = pix / p.OW, ow = pix - oh * p.OW;
    const int ih0 = oh * p.stride - p.pad, iw0 = ow * p.stride - p.pad;
    const long long hw = (long long)p.H * p.W;
    const uint32_t hw4 = (uint32_t)hw * 4u;                 // byte stride between channels
    const char* xb = reinterpret_cast<const char*>(p.x + ((long long)n_img * p.Cin * hw + (long long)ih0 * p.W + iw0));
    // TMEM address of this warp's lane quarter (a warp may only touch lanes 32*(warp%4) .. +31): lane = tile row
    const uint32_t t_lane = tmem_base + ((uint32_t)(gwarp * 32) << 16);
    // The producer loops are bound by instruction issue (ncu: 13 instructions per gathered element before this
    // was restructured), so everything that can be hoisted is: the reduction cursor (kh, kw, channel block) and the
    // smem ring cursor (stage, phase) advance by additions, the per-element address is one IMAD.WIDE
    // (base + r * channel stride), and in the table path the bounds test is one AND + compare against a per-thread
    // mask of valid window rows / columns.
    int cur_cblk = 0, cur_kh = 0, cur_kw = 0;  // CI: position of this group's next k-block to LOAD
    if (CI) {       // k-block kb = cblk * KH*KW + kh*KW + kw  (channel-block-major: the 32 channel planes of a
      const int khw = p.KH * p.KW;       // block are swept by all taps back to back, so they stay in L1)
      cur_cblk = (group * KS) / khw;
      const int khkw = group * KS - cur_cblk * khw;
      cur_kh = khkw / p.KW;
      cur_kw = khkw - cur_kh * p.KW;
    }
    uint32_t win_mask = 0;                     // table path: bit kh = window row inside the image, bit 16+kw = column
    if (!CI && valid_m) {
      auto span = [](int i0, int n, int limit) -> uint32_t {
        const int lo = i0 < 0 ? -i0 : 0;
        int hi = limit - i0;
        hi = hi > n ? n : hi;
        return hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
      };
      win_mask = span(ih0, p.KH, p.H) | (span(iw0, p.KW, p.W) << 16);
    }
    // Software pipeline in registers: the 32 loads of this group's NEXT k-block are issued before the current
    // one is stored, and before its smem slot is even free - loads never wait for the MMA, stores never wait
    // for DRAM/L2 (the data arrived while the previous block was being stored).
#ifdef B200_CONV_TRACE
    const bool tr = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && gtid == 0;
#define B200_TRACE(kb, slot) if (tr) p.trace[(kb) * 8 + (slot)] = clock64()
#else
#define B200_TRACE(kb, slot)
#endif
    auto advance = [&](int steps) {           // move the (cblk, kh, kw) cursor forward by `steps` k-blocks
      for (int t = 0; t < steps; ++t)
        if (++cur_kw == p.KW) {
          cur_kw = 0;
          if (++cur_kh == p.KH) { cur_kh = 0; ++cur_cblk; }
        }
    };
    auto issue = [&](int kb, float (&v)[CBK]) {
      B200_TRACE(kb, 0);
      if (CI) {
        const bool ok = valid_m && (unsigned)(ih0 + cur_kh) < (unsigned)p.H && (unsigned)(iw0 + cur_kw) < (unsigned)p.W;
        const char* q = xb + (((long long)(cur_cblk << 5) * hw + cur_kh * p.W + cur_kw) << 2);
        load_strided<PL>(q, hw4, ok ? CBK : 0, v);     // constant stride H*W between channels
      } else {
        const int2* tab = p.tab + kb * CBK;
#pragma unroll
        for (int r = 0; r < CBK; ++r) {
          const int2 t = __ldg(tab + r);      // uniform address: one broadcast transaction per warp
          v[r] = 0.f;
          if ((win_mask & (uint32_t)t.y) == (uint32_t)t.y)
            v[r] = __ldg(reinterpret_cast<const float*>(xb + ((unsigned long long)(uint32_t)t.x << 2)));
        }
      }
    };
    int c_stage = group;                       // ring cursor of this group's next stage to STORE (stages >= kGroups)
    uint32_t c_phase = 0;
    auto commit = [&](int kb, int sub, const float (&v)[CBK]) {
      B200_TRACE(kb, 6);
      if (sub == 0) mbar_wait(&empty_bar[c_stage], c_phase ^ 1u);
      B200_TRACE(kb, 1);
      uint32_t rr[CBK];
#pragma unroll
      for (int r = 0; r < CBK; ++r) rr[r] = __float_as_uint(round_tf32(v[r]));
      // row = this thread's pixel, 32 K columns of sub-block `sub` of the stage
      tmem_st_32x32(t_lane + kATmemCol0 + (uint32_t)(c_stage * KS + sub) * CBK, rr);
      if (sub == KS - 1) {
        tmem_wait_st();
        B200_TRACE(kb, 7);
        tc_fence_before();
        mbar_arrive(&full_bar[c_stage]);
        B200_TRACE(kb, 2);
        c_stage += kGroups;
        if (c_stage >= stages) { c_stage -= stages; c_phase ^= 1u; }
      }
    };
    {
      // this group's items, in order: stage j = group, group + G, ...; within a stage the sub-blocks 0..KS-1;
      // item i is k-block (group + (i / KS) * G) * KS + i % KS. Loads of item i+1 are in flight while item i is stored.
      const int nstage = p.kblocks / KS;                      // host guarantees kblocks % KS == 0
      const int my_stages = group < nstage ? (nstage - group + kGroups - 1) / kGroups : 0;
      const int nitems = my_stages * KS;
      auto kb_of = [&](int i) { return (group + (i / KS) * kGroups) * KS + (i % KS); };
      auto step_after = [&](int i) { return (i % KS) == KS - 1 ? kGroups * KS - (KS - 1) : 1; };
      float va[CBK], vb[CBK];
      int i = 0;
      if (nitems > 0) { issue(kb_of(0), va); if (CI) advance(step_after(0)); }
      while (i < nitems) {
        if (i + 1 < nitems) { issue(kb_of(i + 1), vb); if (CI) advance(step_after(i + 1)); }
        commit(kb_of(i), i % KS, va);
        if (++i >= nitems) break;
        if (i + 1 < nitems) { issue(kb_of(i + 1), va); if (CI) advance(step_after(i + 1)); }
        commit(kb_of(i), i % KS, vb);
        ++i;
      }
    }
    // ---------------- epilogue: TMEM lane = pixel, columns = output channels; 32-column chunks round-robin over groups ----------------
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    float* yb = p.y + (long long)n_img * p.Cout * p.OHW + pix;
#pragma unroll 1
    for (int c = group * 32; c < BN; c += 32 * kGroups) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(gwarp * 32) << 16) + (uint32_t)c, r);
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
  } else if (warp == kProdWarps) {
    // ---------------- MMA issuer ----------------
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(CBM >> 4) << 24);
    mma_issue_loop_ts<CBK / CUK, KS>(tmem_base, tmem_base + kATmemCol0, CBK * KS, idesc, smem_u32(smem), STAGE_BYTES,
                                     operand_k_major(), smem_u32(&full_bar[0]), smem_u32(&empty_bar[0]),
                                     smem_u32(tmem_full_bar), stages, p.kblocks / KS, B_BYTES);
  } else if (lane == 0) {
    // ---------------- weight-tile TMA producer: runs the whole ring ahead of the gather groups, so the B tile of
    // a stage is requested the moment the stage frees up instead of when a gather group gets round to it ----------------
    int s = 0;
    uint32_t ph = 0;
    for (int kb = 0; kb < p.kblocks; kb += KS) {
      mbar_wait(&empty_bar[s], ph ^ 1u);
      mbar_arrive_expect_tx(&full_bar[s], STAGE_BYTES);
#pragma unroll
      for (int sub = 0; sub < KS; ++sub)      // box {32 k, BN co} each
        tma_load_2d(smem + (size_t)s * STAGE_BYTES + sub * B_BYTES, &tmW, &full_bar[s], (kb + sub) * CBK, n0);
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kProdWarps) tmem_dealloc(tmem_base, kTmemCols);
}

struct ConvWgradParams {
  const float* x;
  const float* dy;
  float* out;          // gradient matrix [Cout][K] in GEMM k order (splits == 1, k_order 0) or partials [z][Cout][K]
  float* dw_direct;    // k_order 1 and splits == 1: dW in torch order, written with the permuted index
  const int2* tab;     // k_order 0: [m_tiles*128] {input offset, 1 << kh | 1 << (16 + kw)}; {0, 1 << 31} beyond K
  int Nb, Cin, H, W, Cout, OH, OW, KH, KW, stride, pad;
  int K, M, OHW, kblocks_total, kblocks_per_split, splits, stages;
};

constexpr uint32_t tmem_cols_for(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : n <= 256 ? 256 : 512; }

// PL as in conv_fwd_kernel; PL > 0 additionally means OH*OW == H*W == PL.
// R = 128-row sub-tiles of the gradient matrix per CTA (R*BN <= 512 TMEM columns). The kernel is bound by the bytes
// it gathers (12 warps x 32 loads in flight against ~1900 cycles of L2 latency under load): a CTA that owns R*128
// k-rows gathers the dY tile once for R accumulators instead of once per accumulator (-30 % gathered bytes at R = 2).
template <int BN, int PL, int R>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(ConvWgradParams p) {
  constexpr bool CI = PL >= 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_SUB_BYTES = CBM * CBK * 4;
  constexpr uint32_t A_BYTES = R * A_SUB_BYTES;
  constexpr uint32_t B_BYTES = BN * CBK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = tmem_cols_for(R * BN);
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tmem_full_bar = empty_bar + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  int2* s_tab = reinterpret_cast<int2*>(tmem_slot + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * (CBM * R), n0 = blockIdx.y * BN, z = blockIdx.z;
  const int kb_begin = z * p.kblocks_per_split;
  int kb_end = kb_begin + p.kblocks_per_split;
  if (kb_end > p.kblocks_total) kb_end = p.kblocks_total;
  const int nkb = kb_end - kb_begin;

  if (!CI)
    for (int t = tid; t < CBM * R; t += blockDim.x) s_tab[t] = __ldg(p.tab + m0 + t);   // table is padded to the tile grid
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 128);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWgProdWarps) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kWgProdWarps) {
    // ---------------- producers: lane <-> pixel of the k-block, warp <-> quarter of each sub-tile's rows,
    //                  group <-> every kWgGroups-th k-block ----------------
    const int group = warp >> 2, gwarp = warp & 3;
    constexpr int B_ROWS = BN / 4;
    const long long hw = (long long)p.H * p.W;
    const uint32_t hw4 = (uint32_t)hw * 4u, ohw4 = (uint32_t)p.OHW * 4u;      // byte strides between channels
    const float inv_ohw = 1.0f / (float)p.OHW, inv_ow = 1.0f / (float)p.OW;
    // K-major SWIZZLE_128B: row i at i*128 B, 16-byte chunk c stored at chunk c ^ (i & 7)
    const uint32_t c16 = (uint32_t)lane >> 2, e16 = ((uint32_t)lane & 3u) * 4u;
    uint32_t s_chunk[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) s_chunk[c] = ((c16 ^ (uint32_t)c) << 4) + e16;
    // channel-block-major order: a warp's 32 rows of a sub-tile are one (channel block, kh, kw) = 32 channels of a
    // tap; the warps of a tile read (mostly) the same 32 channel planes at neighbouring taps -> L1 hits
    const int khw_n = p.KH * p.KW;
    const float inv_khw = 1.0f / (float)khw_n, inv_kw = 1.0f / (float)p.KW;
    // Work of one k-block for a thread = batches of <= 32 loads: [A rows of sub-tile 0] .. [A rows of sub-tile R-1]
    // [B rows 0..31] [B rows 32..]. Software pipeline in registers: the loads of the NEXT batch are issued before the
    // current batch is stored (and before the current k-block's smem slot is known to be free), so a batch of loads
    // is always in flight. Cursors advance by additions, index splits use float reciprocals (exact: operands < 2^23,
    // quotient error << half a step), addresses are `base + immediate` (PL > 0) or a walking 64-bit pointer.
    constexpr int NB = R + (B_ROWS + 31) / 32;
    int li = group, lt = 0;            // load cursor: (k-block, batch)
    int l_pxl = kb_begin * CBK + group * CBK + lane;   // pixel of this lane in k-block li
    bool l_valid = false;
    int l_ih0 = 0, l_iw0 = 0;
    uint32_t l_mask = 0;
    const char* l_xb = reinterpret_cast<const char*>(p.x);
    const char* l_dyb = reinterpret_cast<const char*>(p.dy);
    auto load = [&](float (&v)[32]) {
      if (lt == 0) {
        l_valid = l_pxl < p.M;
        const int pp = l_valid ? l_pxl : 0;
        int n_img = (int)(((float)pp + 0.5f) * inv_ohw), pix = pp - n_img * p.OHW;
        if (pix < 0) { --n_img; pix += p.OHW; } else if (pix >= p.OHW) { ++n_img; pix -= p.OHW; }   // never taken at these sizes
        int oh = (int)(((float)pix + 0.5f) * inv_ow), ow = pix - oh * p.OW;
        if (ow < 0) { --oh; ow += p.OW; } else if (ow >= p.OW) { ++oh; ow -= p.OW; }
        l_ih0 = oh * p.stride - p.pad;
        l_iw0 = ow * p.stride - p.pad;
        l_xb = reinterpret_cast<const char*>(p.x + ((long long)n_img * p.Cin * hw + (long long)l_ih0 * p.W + l_iw0));
        l_dyb = reinterpret_cast<const char*>(p.dy + ((long long)n_img * p.Cout * p.OHW + pix));
        l_pxl += kWgGroups * CBK;
        if (!CI) {
          l_mask = 0;
          if (l_valid) {
            auto span = [](int i0, int n, int limit) -> uint32_t {
              const int lo = i0 < 0 ? -i0 : 0;
              int hi = limit - i0;
              hi = hi > n ? n : hi;
              return hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
            };
            l_mask = span(l_ih0, p.KH, p.H) | (span(l_iw0, p.KW, p.W) << 16);
          }
        }
      }
      if (lt < R) {
        const int k0 = m0 + lt * CBM + gwarp * 32;         // first of this warp's 32 rows in sub-tile lt
        if (CI) {
          const bool rows_valid = k0 < p.K;
          const int slab = rows_valid ? (k0 >> 5) : 0;     // = cblk * KH*KW + kh*KW + kw
          int cblk = (int)(((float)slab + 0.5f) * inv_khw), khkw = slab - cblk * khw_n;
          if (khkw < 0) { --cblk; khkw += khw_n; } else if (khkw >= khw_n) { ++cblk; khkw -= khw_n; }
          int kh = (int)(((float)khkw + 0.5f) * inv_kw), kw = khkw - kh * p.KW;
          if (kw < 0) { --kh; kw += p.KW; } else if (kw >= p.KW) { ++kh; kw -= p.KW; }
          const bool ok = l_valid && rows_valid && (unsigned)(l_ih0 + kh) < (unsigned)p.H && (unsigned)(l_iw0 + kw) < (unsigned)p.W;
          const char* q = l_xb + (((long long)(cblk << 5) * hw + kh * p.W + kw) << 2);
          load_strided<PL>(q, hw4, ok ? 32 : 0, v);
        } else {
          const int2* tb = s_tab + lt * CBM + gwarp * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int2 t = tb[j];
            v[j] = 0.f;
            if ((l_mask & (uint32_t)t.y) == (uint32_t)t.y)
              v[j] = __ldg(reinterpret_cast<const float*>(l_xb + ((unsigned long long)(uint32_t)t.x << 2)));
          }
        }
      } else {
        const int r0 = (lt - R) * 32;
        const int co0 = n0 + gwarp * B_ROWS + r0;
        int nrows = B_ROWS - r0;                       // rows of this batch that exist in the tile and the tensor
        if (nrows > p.Cout - co0) nrows = p.Cout - co0;
        if (!l_valid) nrows = 0;
        const char* q = l_dyb + (unsigned long long)ohw4 * (unsigned)co0;
        load_strided<PL>(q, ohw4, nrows, v);
      }
      if (++lt == NB) { lt = 0; li += kWgGroups; }
    };
    int si = group, st_ = 0;           // store cursor: (k-block, batch)
    int c_stage = group;               // its ring slot and phase (stages >= kWgGroups)
    uint32_t c_phase = 0;
    auto store = [&](const float (&v)[32]) {
      const uint32_t sa = smem_u32(smem + (size_t)c_stage * STAGE_BYTES);
      if (st_ == 0) mbar_wait(&empty_bar[c_stage], c_phase ^ 1u);
      if (st_ < R) {
        uint32_t sd[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) sd[c] = sa + (uint32_t)st_ * A_SUB_BYTES + (uint32_t)(gwarp * 32) * 128u + s_chunk[c];
#pragma unroll
        for (int j = 0; j < 32; ++j)      // (gwarp*32 + j) & 7 == j & 7
          sts_f32(sd[j & 7] + (uint32_t)j * 128u, round_tf32(v[j]));
      } else {
        const int r0 = (st_ - R) * 32;
        const int nrows = B_ROWS - r0;
        uint32_t sd[8];
#pragma unroll
        for (int c = 0; c < 8; ++c)       // B_ROWS and r0 are multiples of 8 -> row & 7 == j & 7
          sd[c] = sa + A_BYTES + (uint32_t)(gwarp * B_ROWS + r0) * 128u + s_chunk[c];
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < nrows) sts_f32(sd[j & 7] + (uint32_t)j * 128u, round_tf32(v[j]));
      }
      if (++st_ == NB) {
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[c_stage]);
        st_ = 0;
        si += kWgGroups;
        c_stage += kWgGroups;
        if (c_stage >= stages) { c_stage -= stages; c_phase ^= 1u; }
      }
    };
    {
      float ra[32], rb[32];
      if (li < nkb) load(ra);
      while (si < nkb) {
        if (li < nkb) load(rb);
        store(ra);
        if (si >= nkb) break;
        if (li < nkb) load(ra);
        store(rb);
      }
    }
    // ---------------- epilogue: TMEM lane = k index, columns [sub*BN, sub*BN + BN) = output channels of sub-tile `sub` ----------------
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
#pragma unroll 1
    for (int sub = 0; sub < R; ++sub) {
      const int m = m0 + sub * CBM + gwarp * 32 + lane;
      float* ob;
      const long long ld = p.K;
      if (CI && p.splits == 1) {       // un-permute (cblk,kh,kw,ci%32) -> torch's (ci,kh,kw) while storing
        const int mm = m < p.K ? m : 0;
        const int slab_e = mm >> 5, cblk_e = slab_e / khw_n, khkw = slab_e - cblk_e * khw_n;
        const int ci = (cblk_e << 5) + (mm & 31);
        ob = p.dw_direct + (long long)ci * khw_n + khkw;
      } else {
        ob = p.out + (long long)z * p.Cout * p.K + m;
      }
#pragma unroll 1
      for (int c = group * 32; c < BN; c += 32 * kWgGroups) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(gwarp * 32) << 16) + (uint32_t)(sub * BN + c), r);
        if (m < p.K) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int co = n0 + c + j;
            if (co < p.Cout) ob[(long long)co * ld] = __uint_as_float(r[j]);
          }
        }
      }
    }
  } else {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(CBM >> 4) << 24);
    mma_issue_loop<CBK / CUK, R>(tmem_base, idesc, smem_u32(smem), STAGE_BYTES, A_BYTES, operand_k_major(),
                                 operand_k_major(), smem_u32(&full_bar[0]), smem_u32(&empty_bar[0]),
                                 smem_u32(tmem_full_bar), stages, nkb, nullptr, A_SUB_BYTES, BN);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kWgProdWarps) tmem_dealloc(tmem_base, TMEM_COLS);
}

// dW[co][ci][kh][kw] = sum_z ws[z][co][((ci/32)*KH*KW + kh*KW + kw)*32 + ci%32]   (k_order 1 partials -> torch order)
__global__ void __launch_bounds__(256) wgrad_finalize_ci_kernel(const float* __restrict__ ws, int splits, int Cout, int Cin,
                                                                int KHW, float* __restrict__ dw) {
  // one thread per 4 consecutive GEMM-order k of one output channel (K % 32 == 0: the four share (cblk, khkw) and are
  // four consecutive channels): 16-byte loads of the partials, four slices in flight, 32-bit index arithmetic.
  // Partials are added in slice order (deterministic).
  const int K = Cin * KHW, K4 = K >> 2;
  const int total4 = Cout * K4;                       // host guarantees Cout*K < 2^31
  const long long slice4 = ((long long)Cout * K) >> 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const int co = i / K4, k = (i - co * K4) << 2;
    const float4* src = reinterpret_cast<const float4*>(ws) + i;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int z = 0;
    for (; z + 4 <= splits; z += 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(src + (long long)(z + u) * slice4);
#pragma unroll
      for (int u = 0; u < 4; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    }
    for (; z < splits; ++z) {
      const float4 v = __ldg(src + (long long)z * slice4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    const int slab = k >> 5, cblk = slab / KHW, khkw = slab - cblk * KHW;
    const int ci = (cblk << 5) + (k & 31);
    float* o = dw + (long long)co * K + (long long)ci * KHW + khkw;      // torch order: channels KHW apart
    o[0] = acc.x; o[KHW] = acc.y; o[2 * KHW] = acc.z; o[3 * KHW] = acc.w;
  }
}

// mode 0: out[co][k] = w[co][k] for k < K, zero padding up to ld_out (row stride multiple of 4 for TMA)
// mode 1: dgrad weights, k' = (co, kh', kw'):  out[ci][co*KH*KW + (KH-1-kh)*KW + (KW-1-kw)] = w[co][ci][kh][kw]
// mode 2: forward weights, channel-block-major:  out[co][((ci/32)*KHW + kh*KW + kw)*32 + ci%32]               = w[co][ci][kh][kw]
// mode 3: dgrad weights, channel-block-major:    out[ci][((co/32)*KHW + (KH-1-kh)*KW + (KW-1-kw))*32 + co%32] = w[co][ci][kh][kw]
// mode 4 | S << 8: forward weights of the space-to-depth form of a stride-S convolution (see s2d_pad_kernel): the
//         equivalent stride-1 kernel has KH' = ceil(KH/S) taps over Cp = 64 channels c' = (ci*S + a)*S + b with
//         w'[co][c'][q][r] = w[co][ci][S*q + a][S*r + b] (0 outside the KHxKW window and for c' >= Cin*S*S), stored
//         channel-block-major: out[co][((c'/32)*KH'*KW' + q*KW' + r)*32 + c'%32]
__device__ __forceinline__ void weight_prep_body(const float* __restrict__ w, float* __restrict__ out, int Cout, int Cin,
                                                 int KH, int KW, long long ld_out, int mode, long long first,
                                                 long long stride) {
  const int KHW = KH * KW;
  if ((mode & 0xff) == 4) {
    const int S = mode >> 8, KHp = (KH + S - 1) / S, KWp = (KW + S - 1) / S, Cp = 64;
    const int Kp = Cp * KHp * KWp;
    const long long total = (long long)Cout * Kp;
    for (long long i = first; i < total; i += stride) {
      // i enumerates the OUTPUT row-major over [co][Kp]
      const int co = (int)(i / Kp), k = (int)(i - (long long)co * Kp);
      const int slab = k >> 5, cblk = slab / (KHp * KWp), qr = slab - cblk * (KHp * KWp);
      const int q = qr / KWp, r = qr - q * KWp, cp = (cblk << 5) + (k & 31);
      const int b = cp % S, a = (cp / S) % S, ci = cp / (S * S);
      const int kh = S * q + a, kw = S * r + b;
      float v = 0.f;
      if (ci < Cin && kh < KH && kw < KW) v = __ldg(w + (((long long)co * Cin + ci) * KH + kh) * KW + kw);
      out[(long long)co * ld_out + k] = v;
    }
    return;
  }
  if (mode == 5) {
    // first-layer slab order (conv1.cu): k' = ci * SC + kh * 12 + kw, SC = KH * 12 rounded up to a multiple of 8; zero elsewhere
    const int SC = (KH * 12 + 7) / 8 * 8, ld = (int)ld_out;
    const int total = Cout * ld;
    for (int i = (int)first; i < total; i += (int)stride) {
      const int co = i / ld, k = i - co * ld;
      const int ci = k / SC, r = k - ci * SC, kh = r / 12, kw = r - kh * 12;
      float v = 0.f;
      if (ci < Cin && kh < KH && kw < KW) v = __ldg(w + (((long long)co * Cin + ci) * KH + kh) * KW + kw);
      out[i] = v;
    }
    return;
  }
  if (mode == 0) {
    const int K = Cin * KHW;
    const int total = Cout * (int)ld_out, ld = (int)ld_out;       // host guarantees < 2^31
    for (int i = (int)first; i < total; i += (int)stride) {
      const int co = i / ld, k = i - co * ld;
      out[i] = k < K ? __ldg(w + (long long)co * K + k) : 0.f;
    }
  } else if (mode == 2 || mode == 3) {
    // i enumerates the OUTPUT (coalesced stores; the strided reads of the 10 MB of weights are served by L2)
    const int total = Cout * Cin * KHW;                           // host guarantees < 2^31
    if (mode == 2) {
      const int K = Cin * KHW;
      for (int i = (int)first; i < total; i += (int)stride) {
        const int co = i / K, k = i - co * K;
        const int slab = k >> 5, cblk = slab / KHW, khkw = slab - cblk * KHW, ci = (cblk << 5) + (k & 31);
        out[(long long)co * ld_out + k] = __ldg(w + ((long long)co * Cin + ci) * KHW + khkw);
      }
    } else {
      const int Kp = Cout * KHW;
      for (int i = (int)first; i < total; i += (int)stride) {
        const int ci = i / Kp, k = i - ci * Kp;
        const int slab = k >> 5, cob = slab / KHW, khkwp = slab - cob * KHW, co = (cob << 5) + (k & 31);
        out[(long long)ci * ld_out + k] = __ldg(w + ((long long)co * Cin + ci) * KHW + (KHW - 1 - khkwp));
      }
    }
  } else {
    const int total = Cout * Cin * KHW;
    for (int i = (int)first; i < total; i += (int)stride) {
      // i enumerates the OUTPUT: [ci][co][kh'][kw']
      const int kwp = i % KW;
      int t = i / KW;
      const int khp = t % KH;
      t /= KH;
      const int co = t % Cout, ci = t / Cout;
      const long long src = (((long long)co * Cin + ci) * KH + (KH - 1 - khp)) * KW + (KW - 1 - kwp);
      out[(long long)ci * ld_out + (long long)co * KHW + khp * KW + kwp] = __ldg(w + src);
    }
  }
}

__global__ void __launch_bounds__(256) conv_weight_prep_kernel(const float* __restrict__ w, float* __restrict__ out,
                                                               int Cout, int Cin, int KH, int KW, long long ld_out,
                                                               int mode) {
  weight_prep_body(w, out, Cout, Cin, KH, KW, ld_out, mode, (long long)blockIdx.x * blockDim.x + threadIdx.x,
                   (long long)gridDim.x * blockDim.x);
}

// All weight copies of a step in ONE launch (the toy model needs 9: five forward, four dgrad): job j owns the
// CTAs [first_block[j], first_block[j+1]).
constexpr int kMaxPrepJobs = 16;
struct PrepJobs {
  b200ddp_weight_prep_job job[kMaxPrepJobs];
  int first_block[kMaxPrepJobs + 1];
  int n;
};
__global__ void __launch_bounds__(256) conv_weight_prep_multi_kernel(const __grid_constant__ PrepJobs J) {
  int j = 0;
  while (j + 1 < J.n && (int)blockIdx.x >= J.first_block[j + 1]) ++j;
  const b200ddp_weight_prep_job& q = J.job[j];
  const int nblocks = J.first_block[j + 1] - J.first_block[j];
  weight_prep_body(q.w, q.out, q.Cout, q.Cin, q.KH, q.KW, q.ld_out, q.mode,
                   (long long)((int)blockIdx.x - J.first_block[j]) * blockDim.x + threadIdx.x,
                   (long long)nblocks * blockDim.x);
}

// Space-to-depth with padding for a stride-S convolution over few channels (conv1: 3 channels, 11x11, stride 4):
//   out[n][(c*S + a)*S + b][i][j] = x[n][c][S*i + a - pad][S*j + b - pad]
// turns it into a stride-1 ceil(KH/S) x ceil(KW/S) convolution over C*S*S channels, whose im2col gather reads
// CONSECUTIVE pixels (the strided original touches four 128-byte lines per warp load). One thread per input element;
// `out` has Cp >= C*S*S channel planes of Hs x Ws and must be zero where no input lands (zeroed once by the caller:
// the border and the channels beyond C*S*S are never written).
__global__ void __launch_bounds__(256) s2d_pad_kernel(const float* __restrict__ x, float* __restrict__ out, int C, int H,
                                                      int W, int S, int pad, int Cp, int Hs, int Ws) {
  // grid: (chunks of a plane, n*C + c). One thread per (row h, group k of S consecutive columns): it reads x[h][S*k + t],
  // t < S, and stores element t to channel (c, a, (pad+t) % S) at column k + (pad+t) / S. Consecutive lanes are
  // consecutive k, so every load and every one of the S stores of a warp is one contiguous run.
  const int plane = blockIdx.y;
  const int n = plane / C, c = plane - n * C;
  const int Wk = (W + S - 1) / S;
  const int per_plane = H * Wk;
  float* op = out + (long long)n * Cp * Hs * Ws;
  const long long cs = (long long)Hs * Ws;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < per_plane; e += gridDim.x * blockDim.x) {
    const int h = e / Wk, k = e - h * Wk;
    const int hp = h + pad, i = hp / S, a = hp - i * S;
    if (i >= Hs) continue;
    const float* xr = x + ((long long)plane * H + h) * W + (long long)S * k;
    float* orow = op + ((long long)((c * S + a) * S) * cs) + (long long)i * Ws;      // channel (c,a,0), row i
    for (int t = 0; t < S; ++t) {
      if (S * k + t >= W) break;
      const int q = pad + t, dj = q / S, b = q - dj * S, j = k + dj;
      if (j < Ws) orow[b * cs + j] = __ldg(xr + t);
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

template <int BN, int PL, bool TWO = false, int KS = 1>
int launch_fwd(const CUtensorMap& tmW, const ConvFwdParams& p, cudaStream_t stream) {
  const size_t stage_bytes = (size_t)BN * CBK * 4 * KS;  // weight tiles only: the im2col tile lives in TMEM
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 16;
  static std::atomic<unsigned long long> attr_devices{0};
  {
    cudaError_t e = ensure_max_dyn_smem(conv_fwd_kernel<BN, PL, TWO, KS>, attr_devices);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv_fwd_kernel<%d>) -> %s", BN, cudaGetErrorString(e));
      return kCudaErrBase + (int)e;
    }
  }
  dim3 grid((p.M + CBM - 1) / CBM, (p.Cout + BN - 1) / BN, 1);
  conv_fwd_kernel<BN, PL, TWO, KS><<<grid, FwdShape<TWO>::kThreads, smem, stream>>>(tmW, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv_fwd_kernel<%d> launch (grid %d,%d smem %zu) -> %s", BN, grid.x, grid.y, smem, cudaGetErrorString(e));
    return kCudaErrBase + (int)e;
  }
  count_launch();
  return 0;
}

template <int BN, int PL, int R>
int launch_wgrad(const ConvWgradParams& p, cudaStream_t stream) {
  const size_t stage_bytes = (size_t)(CBM * R + BN) * CBK * 4;
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + (2 * p.stages + 1) * 8 + 32 + (size_t)CBM * R * 8;
  static std::atomic<unsigned long long> attr_devices{0};
  {
    cudaError_t e = ensure_max_dyn_smem(conv_wgrad_kernel<BN, PL, R>, attr_devices);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(conv_wgrad_kernel<%d,%d,%d>) -> %s", BN, PL, R, cudaGetErrorString(e));
      return kCudaErrBase + (int)e;
    }
  }
  dim3 grid((p.K + CBM * R - 1) / (CBM * R), (p.Cout + BN - 1) / BN, p.splits);
  conv_wgrad_kernel<BN, PL, R><<<grid, kWgThreads, smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("conv_wgrad_kernel<%d,%d,%d> launch (grid %d,%d,%d smem %zu) -> %s", BN, PL, R, grid.x, grid.y, grid.z, smem,
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
  int bn, r, m_tiles, n_tiles, kblocks_total, splits, per;
};

inline WgradPlan plan_wgrad(int Nb, int Cin, int OH, int OW, int Cout, int KH, int KW) {
  WgradPlan w;
  const int K = Cin * KH * KW;
  w.bn = pick_bn(Cout);
  // row sub-tiles per CTA: 2 for the channel-block-major layers (2*BN <= 512 TMEM columns, 3 stages <= 192 KB),
  // 3 for the table path with a 64-channel tile (conv1: K = 363 = one CTA row), else 1
  const bool ci = (Cin % 32) == 0;
  w.r = 1;
  if (ci && K > CBM) w.r = 2;
  if (!ci && w.bn == 64 && K > 2 * CBM) w.r = 3;
  w.m_tiles = (K + CBM * w.r - 1) / (CBM * w.r);
  w.n_tiles = (Cout + w.bn - 1) / w.bn;
  const long long M = (long long)Nb * OH * OW;
  w.kblocks_total = (int)((M + CBK - 1) / CBK);
  int splits = sm_budget() / (w.m_tiles * w.n_tiles);
  if (splits < 1) splits = 1;
  if (splits > sm_budget()) splits = sm_budget();
  if (splits > w.kblocks_total) splits = w.kblocks_total;
  w.per = (w.kblocks_total + splits - 1) / splits;
  w.splits = (w.kblocks_total + w.per - 1) / w.per;
  return w;
}

}  // namespace

}  // namespace b200

using namespace b200;

#ifdef B200_CONV_TRACE
static long long* g_conv_trace = nullptr;
extern "C" int b200ddp_debug_conv_trace(long long* dev_buf) { g_conv_trace = dev_buf; return 0; }
#endif

extern "C" {

int b200ddp_conv2d_fwd(const float* x, const float* wmat, long long ldw, const float* bias, float* y, const int* ktab,
                       int k_order, int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride, int pad, int relu,
                       int device, void* stream) {
  B200_CHECK_ARG(x && wmat && y, "conv2d_fwd: null argument");
  if (k_order == 2) {       // the first layer's own kernel (conv1.cu): image slab staged in shared memory by TMA
    B200_CHECK_ARG(Nb > 0 && H > 0 && W > 0, "conv2d_fwd: bad shape");
    B200_DEVICE(device);
    return conv1_fwd_launch(x, wmat, ldw, bias, y, Nb, Cin, H, W, Cout, KH, KW, stride, pad, relu, as_stream(stream));
  }
  B200_CHECK_ARG(k_order == 0 || k_order == 1, "conv2d_fwd: k_order must be 0 (ci,kh,kw), 1 (ci/32,kh,kw,ci%%32) or 2 (first layer)");
  B200_CHECK_ARG(k_order == 1 || (ktab && ((uintptr_t)ktab & 7) == 0), "conv2d_fwd: k_order 0 needs an 8-byte aligned ktab");
  B200_CHECK_ARG(k_order == 0 || (Cin % 32) == 0, "conv2d_fwd: k_order 1 needs Cin %% 32 == 0 (Cin = %d)", Cin);
  B200_CHECK_ARG(k_order == 1 || (KH <= 15 && KW <= 15), "conv2d_fwd: k_order 0 supports kernel sizes up to 15 (got %dx%d)", KH, KW);
  B200_CHECK_ARG(Nb > 0 && Cin > 0 && Cout > 0 && KH > 0 && KW > 0 && stride > 0 && pad >= 0, "conv2d_fwd: bad shape");
  B200_CHECK_ARG((ldw & 3) == 0 && ((uintptr_t)wmat & 15) == 0, "conv2d_fwd: weight rows must be 16-byte aligned (ldw %% 4 == 0)");
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(OH > 0 && OW > 0, "conv2d_fwd: empty output");
  B200_CHECK_ARG((long long)Nb * OH * OW < (1ll << 31) && (long long)Cin * H * W < (1ll << 31), "conv2d_fwd: tensor too large");
  B200_DEVICE(device);
  const int K = Cin * KH * KW;
  ConvFwdParams p;
  p.x = x; p.y = y; p.bias = bias; p.tab = reinterpret_cast<const int2*>(ktab);
  p.Nb = Nb; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.OH = OH; p.OW = OW; p.KH = KH; p.KW = KW; p.stride = stride; p.pad = pad;
  p.kblocks = (K + CBK - 1) / CBK;
  p.M = Nb * OH * OW; p.OHW = OH * OW; p.relu = relu;
  p.trace = nullptr;
#ifdef B200_CONV_TRACE
  p.trace = g_conv_trace;
#endif
  const int bn = pick_bn(Cout);     // (128-wide tiles for the 256-channel layers were measured: 99 vs 95 us, no)
  // ring depth: 8 A stages fit the TMEM columns left of the accumulator; the weight-tile ring is <= 8 x 32 KB,
  // which leaves >= ~90 KB of the SM array as L1 for the channel-block-major reuse
  int stages = (int)((160 * 1024) / ((size_t)bn * CBK * 4));      // refined below once KS is known
  const bool two = bn == 64 && p.kblocks <= 32;      // short tiles: two CTAs per SM
  // 64-channel tiles with a long, even reduction (conv2 dgrad): two k-blocks per stage
  const int ks = (!two && bn <= 192 && k_order == 1 && p.kblocks >= 16 && p.kblocks % 2 == 0) ? 2 : 1;
  const int max_stages = (two ? FwdShape<true>::kMaxStages : FwdShape<false>::kMaxStages) / ks;
  const int groups = two ? FwdShape<true>::kGroupsT : FwdShape<false>::kGroupsT;
  stages /= ks;
  if (stages > max_stages) stages = max_stages;
  if (stages > p.kblocks / ks) stages = p.kblocks / ks;
  if (stages < groups) stages = groups;
  p.stages = stages;
  CUtensorMap tmW;
  int rc = tma_map_2d(&tmW, wmat, ldw, Cout, ldw, CBK, bn, 0);
  if (rc) return rc;
  cudaStream_t s = as_stream(stream);
#define B200_FWD_BN(PLV)                                         \
  switch (bn) {                                                  \
    case 64: return launch_fwd<64, PLV>(tmW, p, s);              \
    case 128: return launch_fwd<128, PLV>(tmW, p, s);            \
    case 192: return launch_fwd<192, PLV>(tmW, p, s);            \
    default: return launch_fwd<256, PLV>(tmW, p, s);             \
  }
  if (k_order == 1) {
    const int plane = H * W;
    if (two) {                                                         // conv1 after space-to-depth, small test shapes
      if (plane == 57 * 57) return launch_fwd<64, 57 * 57, true>(tmW, p, s);
      return launch_fwd<64, 0, true>(tmW, p, s);
    }
    if (ks == 2) {                                                     // bn <= 192
#define B200_FWD_KS2(PLV)                                                       \
      switch (bn) {                                                             \
        case 64: return launch_fwd<64, PLV, false, 2>(tmW, p, s);               \
        case 128: return launch_fwd<128, PLV, false, 2>(tmW, p, s);             \
        default: return launch_fwd<192, PLV, false, 2>(tmW, p, s);              \
      }
      if (plane == 27 * 27) { B200_FWD_KS2(27 * 27) }
      if (plane == 13 * 13) { B200_FWD_KS2(13 * 13) }
      B200_FWD_KS2(0)
#undef B200_FWD_KS2
    }
    if (plane == 27 * 27) { B200_FWD_BN(27 * 27) }
    if (plane == 13 * 13) { B200_FWD_BN(13 * 13) }
    B200_FWD_BN(0)
  }
  if (two) return launch_fwd<64, -1, true>(tmW, p, s);
  B200_FWD_BN(-1)
#undef B200_FWD_BN
}

long long b200ddp_conv2d_wgrad_workspace_bytes(int Nb, int Cin, int H, int W, int Cout, int KH, int KW, int stride,
                                               int pad) {
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  const WgradPlan w = plan_wgrad(Nb, Cin, OH, OW, Cout, KH, KW);
  long long need = w.splits > 1 ? (long long)w.splits * Cout * Cin * KH * KW * 4 : 0;
  if (Cin == 3 && KH == 11 && KW == 11 && stride == 4 && pad == 2 && Cout == 64) {     // also enough for k_order 2 (conv1.cu)
    const long long slab = conv1_wgrad_workspace_bytes(Nb, H, W, pad);
    if (slab > need) need = slab;
  }
  return need;
}

int b200ddp_conv2d_wgrad(const float* x, const float* dy, float* dw, const int* ktab, int k_order, int Nb, int Cin, int H,
                         int W, int Cout, int KH, int KW, int stride, int pad, void* ws, long long ws_bytes, int device,
                         void* stream) {
  B200_CHECK_ARG(x && dy && dw, "conv2d_wgrad: null argument");
  if (k_order == 2) {       // the first layer's own kernel (conv1.cu)
    B200_CHECK_ARG(Nb > 0 && H > 0 && W > 0, "conv2d_wgrad: bad shape");
    B200_DEVICE(device);
    return conv1_wgrad_launch(x, dy, dw, Nb, Cin, H, W, Cout, KH, KW, stride, pad, ws, ws_bytes, as_stream(stream));
  }
  B200_CHECK_ARG(k_order == 0 || k_order == 1, "conv2d_wgrad: k_order must be 0, 1 or 2 (first layer)");
  B200_CHECK_ARG(k_order == 1 || (ktab && ((uintptr_t)ktab & 7) == 0), "conv2d_wgrad: k_order 0 needs an 8-byte aligned ktab");
  B200_CHECK_ARG(k_order == 0 || (Cin % 32) == 0, "conv2d_wgrad: k_order 1 needs Cin %% 32 == 0 (Cin = %d)", Cin);
  B200_CHECK_ARG(k_order == 1 || (KH <= 15 && KW <= 15), "conv2d_wgrad: k_order 0 supports kernel sizes up to 15 (got %dx%d)", KH, KW);
  const int OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  B200_CHECK_ARG(OH > 0 && OW > 0 && Nb > 0, "conv2d_wgrad: bad shape");
  B200_CHECK_ARG((long long)Nb * OH * OW < (1ll << 31) && (long long)Cin * H * W < (1ll << 31), "conv2d_wgrad: tensor too large");
  B200_DEVICE(device);
  const WgradPlan w = plan_wgrad(Nb, Cin, OH, OW, Cout, KH, KW);
  const int K = Cin * KH * KW;
  B200_CHECK_ARG(w.splits == 1 || (ws && ws_bytes >= (long long)w.splits * Cout * K * 4), "conv2d_wgrad: workspace too small");
  ConvWgradParams p;
  p.x = x; p.dy = dy; p.tab = reinterpret_cast<const int2*>(ktab);
  p.out = w.splits > 1 ? reinterpret_cast<float*>(ws) : dw;
  p.dw_direct = dw;
  p.Nb = Nb; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.OH = OH; p.OW = OW; p.KH = KH; p.KW = KW;
  p.stride = stride; p.pad = pad;
  p.K = K; p.M = Nb * OH * OW; p.OHW = OH * OW;
  p.kblocks_total = w.kblocks_total; p.kblocks_per_split = w.per; p.splits = w.splits;
  const size_t stage_bytes = (size_t)(CBM * w.r + w.bn) * CBK * 4;
  int stages = (int)((200 * 1024) / stage_bytes);
  if (stages > 6) stages = 6;
  if (stages > w.per) stages = w.per;
  if (stages < kWgGroups) stages = kWgGroups;
  p.stages = stages;
  cudaStream_t s = as_stream(stream);
  int rc;
#define B200_WG_BN(PLV, RV)                                      \
  switch (w.bn) {                                                \
    case 64: rc = launch_wgrad<64, PLV, RV>(p, s); break;        \
    case 128: rc = launch_wgrad<128, PLV, RV>(p, s); break;      \
    case 192: rc = launch_wgrad<192, PLV, RV>(p, s); break;      \
    default: rc = launch_wgrad<256, PLV, RV>(p, s); break;       \
  }
  if (k_order == 1) {
    const int plane = (H * W == OH * OW) ? H * W : 0;            // the compile-time stride serves x and dy
    if (w.r == 2) {
      if (plane == 27 * 27) { B200_WG_BN(27 * 27, 2) }
      else if (plane == 13 * 13) { B200_WG_BN(13 * 13, 2) }
      else { B200_WG_BN(0, 2) }
    } else {
      B200_WG_BN(0, 1)
    }
  } else if (w.r == 3) {
    rc = launch_wgrad<64, -1, 3>(p, s);
  } else {
    B200_WG_BN(-1, 1)
  }
#undef B200_WG_BN
  if (rc) return rc;
  if (w.splits > 1) {
    if (k_order == 1) {
      const long long total = (long long)Cout * K;
      B200_CHECK_ARG(total < (1LL << 31), "conv2d_wgrad: weight tensor too large");
      long long want = (total / 4 + 255) / 256;
      const int grid = (int)(want < num_sms() * 16 ? want : num_sms() * 16);
      wgrad_finalize_ci_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(ws), w.splits, Cout, Cin, KH * KW, dw);
      B200_LAUNCH_OK();
      return 0;
    }
    Epilogue ep;
    return splitk_finalize_launch(reinterpret_cast<const float*>(ws), w.splits, Cout, K, dw, K, ep, s);
  }
  return 0;
}

static int check_prep_mode(int Cout, int Cin, int KH, int KW, long long ld_out, int mode, const char* who) {
  const int m = mode & 0xff, S = mode >> 8;
  B200_CHECK_ARG(m >= 0 && m <= 5 && (m == 4 || S == 0), "%s: mode must be 0..3, 4 | S << 8 or 5 (got %d)", who, mode);
  if (m == 5) {
    B200_CHECK_ARG(KW <= 12 && ld_out >= (long long)Cin * ((KH * 12 + 7) / 8 * 8), "%s: mode 5 needs KW <= 12 and ld_out >= Cin * roundup8(12 KH)", who);
    return 0;
  }
  if (m == 4) {
    B200_CHECK_ARG(S >= 2 && S <= 8 && Cin * S * S <= 64, "%s: space-to-depth needs 2 <= S <= 8 and Cin*S*S <= 64", who);
    B200_CHECK_ARG(ld_out >= 64LL * ((KH + S - 1) / S) * ((KW + S - 1) / S), "%s: ld_out too small", who);
    return 0;
  }
  B200_CHECK_ARG((m == 0 || m == 2) ? ld_out >= (long long)Cin * KH * KW : ld_out >= (long long)Cout * KH * KW,
                 "%s: ld_out too small", who);
  return 0;
}
static long long prep_total(int Cout, int Cin, int KH, int KW, long long ld_out, int mode);
static int check_prep_size(int Cout, int Cin, int KH, int KW, long long ld_out, int mode, const char* who) {
  B200_CHECK_ARG(prep_total(Cout, Cin, KH, KW, ld_out, mode) < (1LL << 31) && (long long)Cout * ld_out < (1LL << 31),
                 "%s: weight tensor too large", who);
  return 0;
}
static long long prep_total(int Cout, int Cin, int KH, int KW, long long ld_out, int mode) {
  if ((mode & 0xff) == 4) {
    const int S = mode >> 8;
    return (long long)Cout * 64 * ((KH + S - 1) / S) * ((KW + S - 1) / S);
  }
  return (mode == 0 || mode == 5) ? (long long)Cout * ld_out : (long long)Cout * Cin * KH * KW;
}

int b200ddp_conv2d_weight_prep(const float* w, float* out, int Cout, int Cin, int KH, int KW, long long ld_out, int mode,
                               int device, void* stream) {
  B200_CHECK_ARG(w && out && Cout > 0 && Cin > 0 && KH > 0 && KW > 0, "conv2d_weight_prep: bad argument");
  int rc_mode = check_prep_mode(Cout, Cin, KH, KW, ld_out, mode, "conv2d_weight_prep");
  if (!rc_mode) rc_mode = check_prep_size(Cout, Cin, KH, KW, ld_out, mode, "conv2d_weight_prep");
  if (rc_mode) return rc_mode;
  B200_DEVICE(device);
  const long long total = prep_total(Cout, Cin, KH, KW, ld_out, mode);
  long long want = (total + 255) / 256;
  const int grid = (int)(want < num_sms() * 8 ? (want < 1 ? 1 : want) : num_sms() * 8);
  conv_weight_prep_kernel<<<grid, 256, 0, as_stream(stream)>>>(w, out, Cout, Cin, KH, KW, ld_out, mode);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_conv2d_weight_prep_multi(const b200ddp_weight_prep_job* jobs, int njobs, int device, void* stream) {
  B200_CHECK_ARG(jobs && njobs > 0 && njobs <= kMaxPrepJobs, "conv2d_weight_prep_multi: 1..%d jobs (got %d)", kMaxPrepJobs, njobs);
  PrepJobs J;
  J.n = njobs;
  int blocks = 0;
  for (int j = 0; j < njobs; ++j) {
    const b200ddp_weight_prep_job& q = jobs[j];
    B200_CHECK_ARG(q.w && q.out && q.Cout > 0 && q.Cin > 0 && q.KH > 0 && q.KW > 0, "conv2d_weight_prep_multi: job %d: bad argument", j);
    int rc_mode = check_prep_mode(q.Cout, q.Cin, q.KH, q.KW, q.ld_out, q.mode, "conv2d_weight_prep_multi");
    if (!rc_mode) rc_mode = check_prep_size(q.Cout, q.Cin, q.KH, q.KW, q.ld_out, q.mode, "conv2d_weight_prep_multi");
    if (rc_mode) return rc_mode;
    const long long total = prep_total(q.Cout, q.Cin, q.KH, q.KW, q.ld_out, q.mode);
    long long want = (total + 1023) / 1024;                      // ~4 elements per thread
    if (want < 1) want = 1;
    if (want > num_sms() * 4) want = num_sms() * 4;
    J.job[j] = q;
    J.first_block[j] = blocks;
    blocks += (int)want;
  }
  J.first_block[njobs] = blocks;
  B200_DEVICE(device);
  conv_weight_prep_multi_kernel<<<blocks, 256, 0, as_stream(stream)>>>(J);
  B200_LAUNCH_OK();
  return 0;
}

int b200ddp_space_to_depth(const float* x, float* out, int Nb, int C, int H, int W, int S, int pad, int Cp, int Hs, int Ws,
                           int device, void* stream) {
  B200_CHECK_ARG(x && out && Nb > 0 && C > 0 && H > 0 && W > 0, "space_to_depth: bad argument");
  B200_CHECK_ARG(S >= 2 && S <= 8 && pad >= 0 && Cp >= C * S * S && Hs > 0 && Ws > 0, "space_to_depth: bad geometry");
  B200_DEVICE(device);
  B200_CHECK_ARG((long long)Nb * C <= 65535 && (long long)H * W < (1LL << 30), "space_to_depth: tensor too large");
  const int per_plane = H * ((W + S - 1) / S);
  s2d_pad_kernel<<<dim3((per_plane + 1023) / 1024, Nb * C), 256, 0, as_stream(stream)>>>(x, out, C, H, W, S, pad, Cp, Hs, Ws);
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

namespace b200 {

// ---------------------------------------------------------------------------------------------------------------
// Channel sum for callers that have no device memory to lend as scratch.
//
// b200ddp_channel_sum() needs a caller-provided scratch buffer (per-CTA partial sums + one ticket per channel).
// Some callers cannot produce one: code that runs where every byte of GPU memory already belongs to someone
// else's pool and no allocator is reachable (a plugin called from inside another framework's step), or code on a
// path where cudaMalloc is off limits because it synchronises the whole device and can fail under a memory cap.
// For them the same kernel runs on a fixed scratch area that lives inside this library's device image.
//
// Cost and limits, so nobody is surprised:
//   * the area is kStaticScratchBytes (258 KB) and every GPU that loads libb200ddp.so keeps it until the process
//     exits, whether or not this entry point is ever called;
//   * it fits at most kStaticMaxChannels (512) channels; larger C is refused with an error, not truncated;
//   * all calls on a device share the one area and it keeps whatever the previous call left there. Each call
//     re-zeroes the tickets it is about to use, but two calls that are in flight on the GPU at the same time
//     (different streams, or different host threads feeding different streams) would trample each other's
//     partial sums: issue them on one stream, or order them yourself;
//   * the first call on a device looks the area's device address up (cudaGetSymbolAddress — the array name
//     itself is only a host-side handle, not something a kernel may dereference) and remembers it. That lookup
//     is a set-up step: make the first call on each device before you start capturing any stream.
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kStaticMaxChannels = 512;
constexpr int kStaticMaxSplits = 64;                                   // cs_splits() never exceeds this
constexpr long long kStaticTicketBytes = (long long)kStaticMaxChannels * 4;
constexpr long long kStaticScratchBytes = kStaticTicketBytes + (long long)kStaticMaxChannels * kStaticMaxSplits * 8;
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) unsigned char g_channel_sum_scratch[kStaticScratchBytes];

std::mutex g_static_mu;
void* g_static_base[kStaticMaxDevices] = {nullptr};    // per device: device address of g_channel_sum_scratch
}  // namespace

}  // namespace b200

extern "C" int b200ddp_channel_sum_static(const float* dy, float* out, int Nb, int C, int HW, int device, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(dy && out && Nb > 0 && C > 0 && HW > 0, "channel_sum_static: bad argument");
  B200_CHECK_ARG(C <= kStaticMaxChannels,
                 "channel_sum_static: C = %d channels do not fit the built-in scratch area (at most %d); "
                 "use b200ddp_channel_sum with your own scratch buffer", C, kStaticMaxChannels);
  B200_CHECK_ARG(device >= 0 && device < kStaticMaxDevices, "channel_sum_static: device ordinal %d out of range", device);
  B200_DEVICE(device);                                   // `device` is current from here on
  void* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_base[device]) {                        // first call on this device: one-time address lookup
      void* ptr = nullptr;
      B200_CUDA(cudaGetSymbolAddress(&ptr, g_channel_sum_scratch));
      g_static_base[device] = ptr;
    }
    base = g_static_base[device];
  }
  const int S = cs_splits(Nb, C);                        // <= kStaticMaxSplits
  unsigned int* tickets = reinterpret_cast<unsigned int*>(base);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(base) + kStaticTicketBytes);
  // the area still holds the previous call's state: the tickets of the channels used now must start at zero
  B200_CUDA(cudaMemsetAsync(tickets, 0, (size_t)C * sizeof(unsigned int), as_stream(stream)));
  channel_sum_kernel<<<dim3(S, C), kCsThreads, 0, as_stream(stream)>>>(dy, Nb, C, HW, partials, tickets, out);
  B200_LAUNCH_OK();
  return 0;
}
