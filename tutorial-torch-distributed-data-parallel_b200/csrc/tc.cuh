// tcgen05 / TMEM / TMA / mbarrier PTX wrappers shared by the tensor-core kernels (gemm_tf32.cu, conv.cu).
// sm_100a only. Every wait is bounded: a pipeline bug traps instead of hanging the GPU box.
#pragma once
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>

namespace b200 {
namespace tc {

constexpr unsigned long long kWaitLimitNs = 4000000000ull;  // bounded waits: trap after 4 s instead of hanging the box

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0, spins = 0;
  unsigned long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) break;
    if ((++spins & 0x3ffu) == 0) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kWaitLimitNs) {
        printf("b200ddp: mbarrier wait timed out (block %d,%d,%d thread %d parity %u)\n", blockIdx.x, blockIdx.y,
               blockIdx.z, threadIdx.x, parity);
        __trap();
      }
    }
  }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns of TMEM -> 32 registers per thread.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor (sm_100 format, version 1).
//   K-major tile  [rows][32 fp32], SWIZZLE_128B (layout type 2): 8-row groups are 1024 B apart (SBO);
//                  LBO unused (1).
//   MN-major tile, 32-bit operands: the only legal layout is SWIZZLE_128B_BASE32B (layout type 1; TMA
//                  swizzle 128B_ATOM_32B): rows of 32 MN-contiguous elements (128 B) whose 32-byte chunks
//                  are XOR-ed with (row % 4); a k-group is 4 rows = 512 B (SBO); the next 32-element MN
//                  atom is +4096 B (LBO) because each TMA box holds 32 k-rows x 128 B.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
  const uint32_t lo = ((saddr & 0x3FFFFu) >> 4) | ((lbo_bytes >> 4) << 16);
  const uint32_t hi = (sbo_bytes >> 4) | (1u << 14) /* version */ | (layout_type << 29);
  return (uint64_t)lo | ((uint64_t)hi << 32);
}


// plain arrive (count 1) and transaction-count registration without arrival
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// make generic-proxy shared-memory writes (st.shared) visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// named barrier among `count` threads of the CTA (ids 1..15; id 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// ---- the MMA issuer's main loop -------------------------------------------------------------------------------
// One thread walks the smem ring: wait full[s] -> UMMAS x tcgen05.mma -> tcgen05.commit(empty[s]).
// Everything this thread executes is one dependent chain, so its instruction count per k-block bounds the
// tile rate (measured with clock64 on B200: ~85 cycles per MMA when each descriptor is rebuilt from the stage
// index, plus two runtime integer divisions for stage/phase - more than the tensor core needs for the MMA
// itself at N <= 192). Hence: descriptors are advanced by adding to their low word (the 14-bit start-address
// field, units of 16 B; smem addresses are < 256 KB so the sum never carries into the LBO field), stage and
// phase are counters, and barrier addresses are running 32-bit shared addresses.
struct OperandDesc {
  uint32_t lbo, sbo, layout_type, k_advance;   // bytes, bytes, desc layout type, bytes between consecutive UMMA k-steps
};
__device__ __forceinline__ constexpr OperandDesc operand_k_major() { return {16u, 1024u, 2u, 32u}; }
__device__ __forceinline__ constexpr OperandDesc operand_mn_major() { return {4096u, 512u, 1u, 1024u}; }

__device__ __forceinline__ void umma_tf32_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_addr(uint32_t bar_saddr) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_saddr) : "memory");
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t addr, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  unsigned long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) break;
    if ((++spins & 0x3ffu) == 0) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kWaitLimitNs) {
        printf("b200ddp: mbarrier wait timed out (block %d,%d,%d thread %d parity %u)\n", blockIdx.x, blockIdx.y,
               blockIdx.z, threadIdx.x, parity);
        __trap();
      }
    }
  }
}

// UMMAS = MMAs per k-block (BK / 8 for tf32). Stage layout: [A tile | B tile], stage_bytes apart.
// Called by ALL 32 lanes of the MMA warp with warp-uniform arguments: the loop state then lives in uniform
// registers and only the tcgen05 instructions sit behind elect.sync (a `lane == 0` branch around the whole loop
// makes the compiler rebuild every descriptor through R2UR + an ELECT loop, ~85 cycles per MMA with 12 producer
// warps competing for issue slots).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// RSUB > 1: the stage holds RSUB A sub-tiles (a_sub_bytes apart) that share the B tile; sub-tile r accumulates into
// TMEM columns [r * d_sub_cols, ...).
template <int UMMAS, int RSUB = 1>
__device__ __forceinline__ void mma_issue_loop(uint32_t tmem_d, uint32_t idesc, uint32_t smem_saddr, uint32_t stage_bytes,
                                               uint32_t a_bytes, OperandDesc a, OperandDesc b, uint32_t full_saddr,
                                               uint32_t empty_saddr, uint32_t tmem_full_saddr, int stages, int nkb,
                                               long long* trace = nullptr, uint32_t a_sub_bytes = 0,
                                               uint32_t d_sub_cols = 0) {
  const uint32_t a_lo0 = ((smem_saddr & 0x3FFFFu) >> 4) | ((a.lbo >> 4) << 16);
  const uint32_t b_lo0 = (((smem_saddr + a_bytes) & 0x3FFFFu) >> 4) | ((b.lbo >> 4) << 16);
  const uint32_t a_hi = (a.sbo >> 4) | (1u << 14) | (a.layout_type << 29);
  const uint32_t b_hi = (b.sbo >> 4) | (1u << 14) | (b.layout_type << 29);
  const uint32_t a_step = a.k_advance >> 4, b_step = b.k_advance >> 4, stage_step = stage_bytes >> 4;
  uint32_t a_lo = a_lo0, b_lo = b_lo0, fb = full_saddr, eb = empty_saddr, ph = 0, acc = 0;
  int s = 0;
#pragma unroll 1
  for (int i = 0; i < nkb; ++i) {
#ifdef B200_CONV_TRACE
    if (trace && (threadIdx.x & 31) == 0) trace[i * 8 + 3] = clock64();
#endif
    mbar_wait_addr(fb, ph);
#ifdef B200_CONV_TRACE
    if (trace && (threadIdx.x & 31) == 0) trace[i * 8 + 4] = clock64();
#endif
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int r = 0; r < RSUB; ++r)
#pragma unroll
        for (int k = 0; k < UMMAS; ++k)
          umma_tf32_lohi(tmem_d + r * d_sub_cols, a_lo + r * (a_sub_bytes >> 4) + k * a_step, a_hi, b_lo + k * b_step, b_hi,
                         idesc, k ? 1u : acc);
      umma_commit_addr(eb);               // frees this smem stage when the MMAs above retire
    }
    acc = 1u;
#ifdef B200_CONV_TRACE
    if (trace && (threadIdx.x & 31) == 0) trace[i * 8 + 5] = clock64();
#endif
    if (++s == stages) {
      s = 0; ph ^= 1u; a_lo = a_lo0; b_lo = b_lo0; fb = full_saddr; eb = empty_saddr;
    } else {
      a_lo += stage_step; b_lo += stage_step; fb += 8u; eb += 8u;
    }
  }
  if (elect_one()) umma_commit_addr(tmem_full_saddr);      // accumulator complete -> epilogue may read TMEM
  __syncwarp();
}

// ---- A operand in tensor memory ("TS" form): tcgen05.mma [d], [a_tmem], b_desc --------------------------------
// Lane = row of the 128-row tile, one 32-bit column per tf32 K element: a thread that owns one tile row and holds
// 32 K-values in registers writes them with ONE tcgen05.st (32x32b.x32) - no shared-memory store, no swizzle, and
// the MMA does not read A through the shared-memory port at all.
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], db, %4, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// MMA loop for the TS form: A stage s = TMEM columns [tmem_a0 + s*a_stage_cols, +UMMAS*8), B stage s in shared
// memory at smem_saddr + s*stage_bytes. Same warp-uniform structure as mma_issue_loop.
// KSUB k-blocks per stage: their A columns are contiguous (32 each), their B tiles b_sub_bytes apart.
template <int UMMAS, int KSUB = 1>
__device__ __forceinline__ void mma_issue_loop_ts(uint32_t tmem_d, uint32_t tmem_a0, uint32_t a_stage_cols, uint32_t idesc,
                                                  uint32_t smem_saddr, uint32_t stage_bytes, OperandDesc b,
                                                  uint32_t full_saddr, uint32_t empty_saddr, uint32_t tmem_full_saddr,
                                                  int stages, int nkb, uint32_t b_sub_bytes = 0) {
  const uint32_t b_lo0 = ((smem_saddr & 0x3FFFFu) >> 4) | ((b.lbo >> 4) << 16);
  const uint32_t b_hi = (b.sbo >> 4) | (1u << 14) | (b.layout_type << 29);
  const uint32_t b_step = b.k_advance >> 4, stage_step = stage_bytes >> 4;
  uint32_t ta = tmem_a0, b_lo = b_lo0, fb = full_saddr, eb = empty_saddr, ph = 0, acc = 0;
  int s = 0;
#pragma unroll 1
  for (int i = 0; i < nkb; ++i) {
    mbar_wait_addr(fb, ph);
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int sub = 0; sub < KSUB; ++sub)
#pragma unroll
        for (int k = 0; k < UMMAS; ++k)
          umma_tf32_ts(tmem_d, ta + (uint32_t)(sub * UMMAS + k) * 8u, b_lo + sub * (b_sub_bytes >> 4) + k * b_step, b_hi, idesc,
                       (sub | k) ? 1u : acc);
      umma_commit_addr(eb);               // frees the TMEM A stage and the smem B stage when these MMAs retire
    }
    acc = 1u;
    if (++s == stages) {
      s = 0; ph ^= 1u; ta = tmem_a0; b_lo = b_lo0; fb = full_saddr; eb = empty_saddr;
    } else {
      ta += a_stage_cols; b_lo += stage_step; fb += 8u; eb += 8u;
    }
  }
  if (elect_one()) umma_commit_addr(tmem_full_saddr);
  __syncwarp();
}

}  // namespace tc
}  // namespace b200
