// R4 — the gradient all-reduce as OUR kernel over NVLink 5 / NVSwitch peer memory.
//
// Replaces c10d::Reducer's per-bucket ncclAllReduce (torch reducer.hpp:276-286 via _AllReduceBySumCommHook,
// reached from reference multi-GPU-training-torch.py:245 / :125) including the divide by world size that stock DDP
// does in a separate pass (default_comm_hooks.hpp:36-40).
//
// The gradient arena is a SYMMETRIC buffer: every rank allocates it with ncclMemAlloc and registers it as an NCCL
// window (comm.cu), which maps all ranks' copies into this process — each at a unicast address, and all of them
// together behind one MULTICAST address served by the NVSwitch. NCCL is used for that bootstrap only; the data
// path is the kernel below:
//
//   * the element range of a bucket is cut into `nranks` contiguous shards; rank r owns shard r;
//   * barrier 1 (flags in a second symmetric buffer): every rank's bucket is complete;
//   * owner of an element:  v = multimem.ld_reduce.add [mc + i]   — the switch reads the element from all ranks
//                                                                   and returns the sum (in-switch reduction),
//                           v *= 1/world                            (the averaging rides along)
//                           multimem.st [mc + i], v                — the switch writes it into all ranks' buffers;
//     without a multicast mapping the same owner sums unicast peer loads and stores to every peer (two-shot P2P);
//   * barrier 2: every rank's stores have landed.
//
// Every element is read and written by exactly one rank, so the reduction is in place, replicas receive bit-identical
// values, and the result does not depend on timing. Per GPU the links carry S bytes in each direction (S = bucket
// bytes) against 2 S (N-1)/N for a ring, and the SMs of a rank touch only S/N bytes.
//
// Only ordinary and multimem load/store instructions touch peer memory here (no TMA).
// Waits are bounded (trap after kWaitNs) so that a missing peer cannot hang the GPU.
#include "common.cuh"
#include "peer.h"

namespace b200 {
namespace {

// CTA shape chosen to CO-RESIDE with a running convolution CTA instead of displacing it: the conv kernels hold
// 14 warps x 128 registers (57,344 of an SM's 65,536 registers) and all of its shared / tensor memory, so what is left
// is exactly 256 threads x 32 registers and no shared memory. The multimem kernel is built to that budget
// (__launch_bounds__(256, 8)); many small CTAs then ride along on busy SMs and the convolution grids keep all 148 SMs.
// A second shape does the opposite: few CTAs of 1024 threads x 4 loads in flight each (64 KB per SM), which fill the
// register file and therefore own their SM like NCCL's channels do; the grid planners of the compute kernels then
// leave those SMs out (b200ddp_set_sm_budget). Which one is faster is a measurement.
constexpr int kArThreadsSmall = 256, kUnrollSmall = 2;
constexpr int kArThreadsBig = 1024, kUnrollBig = 4;
constexpr unsigned long long kWaitNs = 8000000000ull;

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer (and own) gradient words as they are in memory NOW: system-scope relaxed load, never an L1 hit from an earlier step
__device__ __forceinline__ float4 ld_sys_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_sys_f4_scalar(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 mm_ld_reduce(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// All ranks' CTA `blockIdx.x` meet: thread t < nranks publishes `epoch` in rank t's inbox[cta][my rank] and waits
// for rank t's epoch in the local inbox[cta][t].
__device__ __forceinline__ void peer_barrier(const PeerArgs& a, unsigned int epoch) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < a.nranks && t != a.rank) {
    const size_t slot = (size_t)blockIdx.x * kPeerMaxRanks;
    st_release_sys(reinterpret_cast<unsigned int*>(a.flags.peer[t]) + slot + a.rank, epoch);
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(a.flags.peer[a.rank]) + slot + t;
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {          // wrap-safe: epochs only grow
      if ((++spins & 0xfffu) == 0) {
        const unsigned long long now = gtimer();
        if (t0 == 0) t0 = now;
        else if (now - t0 > kWaitNs) {
          printf("b200ddp: all-reduce barrier timed out (rank %d waits for rank %d, cta %d, epoch %u)\n", a.rank, t,
                 blockIdx.x, epoch);
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

template <bool MC, int kArThreads, int kUnroll>
__global__ void __launch_bounds__(kArThreads, kArThreads == 256 ? (MC ? 8 : 4) : 1)
peer_allreduce_kernel(PeerArgs a, long long off_bytes, long long n4, float scale) {
  peer_barrier(a, a.epoch + 1);                  // everybody's bucket is final (their producers ran before this kernel)
  const long long shard = (n4 + a.nranks - 1) / a.nranks;
  const long long lo = shard * a.rank, hi = lo + shard < n4 ? lo + shard : n4;
  const long long stride = (long long)gridDim.x * kArThreads;
  if (MC) {
    float4* mc = reinterpret_cast<float4*>(a.data.mc + off_bytes);
    long long i = lo + (long long)blockIdx.x * kArThreads + threadIdx.x;
    for (; i + (kUnroll - 1) * stride < hi; i += kUnroll * stride) {
      float4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) v[u] = mm_ld_reduce(reinterpret_cast<const float*>(mc + i + u * stride));
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
        mm_st(reinterpret_cast<float*>(mc + i + u * stride), v[u]);
      }
    }
    for (; i < hi; i += stride) {
      float4 v = mm_ld_reduce(reinterpret_cast<const float*>(mc + i));
      v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
      mm_st(reinterpret_cast<float*>(mc + i), v);
    }
  } else {
    for (long long i = lo + (long long)blockIdx.x * kArThreads + threadIdx.x; i < hi; i += stride) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 v[kPeerMaxRanks];
#pragma unroll
      for (int p = 0; p < kPeerMaxRanks; ++p)
        if (p < a.nranks) v[p] = ld_sys_f4(reinterpret_cast<const float4*>(a.data.peer[p] + off_bytes) + i);
#pragma unroll
      for (int p = 0; p < kPeerMaxRanks; ++p)       // fixed order: rank 0, 1, ... (deterministic)
        if (p < a.nranks) { acc.x += v[p].x; acc.y += v[p].y; acc.z += v[p].z; acc.w += v[p].w; }
      acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
#pragma unroll
      for (int p = 0; p < kPeerMaxRanks; ++p)
        if (p < a.nranks) *(reinterpret_cast<float4*>(a.data.peer[p] + off_bytes) + i) = acc;
    }
  }
  __threadfence_system();                        // this thread's stores are ordered before the flag that follows
  peer_barrier(a, a.epoch + 2);                  // everybody's shard has landed everywhere
}

}  // namespace

// One CTA. Flags: the last CTA slot of the flag buffer (the bucket all-reduce never uses kPeerMaxCtas CTAs).
__global__ void __launch_bounds__(256) peer_small_collective_kernel(PeerArgs a, const float* __restrict__ send,
                                                                    float* __restrict__ out, int n, int sum_mode, float scale) {
  const unsigned parity = a.epoch & 1u;
  const size_t slot = ((size_t)parity * kPeerMaxRanks + (size_t)a.rank) * kSmallMaxFloats;
  if (a.data.mc) {
    // one store reaches every rank; 16 bytes per packet (the slot is 16-byte aligned and padded, the tail is zero-filled)
    float* dst = reinterpret_cast<float*>(a.data.mc) + slot;
    for (int i = threadIdx.x * 4; i < n; i += blockDim.x * 4) {
      float4 v;
      v.x = send[i];
      v.y = i + 1 < n ? send[i + 1] : 0.f;
      v.z = i + 2 < n ? send[i + 2] : 0.f;
      v.w = i + 3 < n ? send[i + 3] : 0.f;
      mm_st(dst + i, v);
    }
  } else {
    for (int p = 0; p < a.nranks; ++p) {
      float* dst = reinterpret_cast<float*>(a.data.peer[p]) + slot;
      for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = send[i];
    }
  }
  __threadfence_system();
  // flag round on slot kPeerMaxCtas - 1 (peer_barrier indexes by blockIdx.x == 0, so address it by hand)
  __syncthreads();
  const int t = threadIdx.x;
  if (t < a.nranks && t != a.rank) {
    const size_t fs = (size_t)(kPeerMaxCtas - 1) * kPeerMaxRanks;
    st_release_sys(reinterpret_cast<unsigned int*>(a.flags.peer[t]) + fs + a.rank, a.epoch);
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(a.flags.peer[a.rank]) + fs + t;
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while ((int)(ld_acquire_sys(mine) - a.epoch) < 0) {
      if ((++spins & 0xfffu) == 0) {
        const unsigned long long now = gtimer();
        if (t0 == 0) t0 = now;
        else if (now - t0 > kWaitNs) {
          printf("b200ddp: small collective timed out (rank %d waits for rank %d, epoch %u)\n", a.rank, t, a.epoch);
          __trap();
        }
      }
    }
  }
  __syncthreads();
  const float* mine_base = reinterpret_cast<const float*>(a.data.peer[a.rank]) + (size_t)parity * kPeerMaxRanks * kSmallMaxFloats;
  if (sum_mode) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float acc = 0.f;
      for (int r = 0; r < a.nranks; ++r) acc += ld_sys_f4_scalar(mine_base + (size_t)r * kSmallMaxFloats + i);
      out[i] = acc * scale;
    }
  } else {
    for (int r = 0; r < a.nranks; ++r)
      for (int i = threadIdx.x; i < n; i += blockDim.x) out[(size_t)r * n + i] = ld_sys_f4_scalar(mine_base + (size_t)r * kSmallMaxFloats + i);
  }
}

int peer_small_collective_launch(const PeerArgs& a, const float* send, float* out, int n, int sum_mode, float scale,
                                 cudaStream_t stream) {
  B200_CHECK_ARG(send && out && n > 0 && n <= kSmallMaxFloats, "small collective: 1..%d floats (got %d)", kSmallMaxFloats, n);
  B200_CHECK_ARG(a.nranks >= 2 && a.nranks <= kPeerMaxRanks, "small collective: 2..%d ranks", kPeerMaxRanks);
  peer_small_collective_kernel<<<1, 256, 0, stream>>>(a, send, out, n, sum_mode, scale);
  B200_LAUNCH_OK();
  return 0;
}

int peer_allreduce_launch(const PeerArgs& a, long long off_bytes, long long n, float scale, int ctas, int big_ctas,
                          cudaStream_t stream) {
  B200_CHECK_ARG(n >= 0 && (n & 3) == 0 && (off_bytes & 15) == 0, "peer_allreduce: range must be 16-byte aligned (n = %lld)", n);
  B200_CHECK_ARG(a.nranks >= 2 && a.nranks <= kPeerMaxRanks, "peer_allreduce: 2..%d ranks", kPeerMaxRanks);
  if (n == 0) return 0;
  if (ctas < 1) ctas = 1;
  if (ctas > kPeerMaxCtas - 1) ctas = kPeerMaxCtas - 1;       // the last flag slot belongs to the small collectives
  const long long n4 = n >> 2;
  if (big_ctas) {
    if (a.data.mc) peer_allreduce_kernel<true, kArThreadsBig, kUnrollBig><<<ctas, kArThreadsBig, 0, stream>>>(a, off_bytes, n4, scale);
    else peer_allreduce_kernel<false, kArThreadsBig, 1><<<ctas, kArThreadsBig, 0, stream>>>(a, off_bytes, n4, scale);
  } else {
    if (a.data.mc) peer_allreduce_kernel<true, kArThreadsSmall, kUnrollSmall><<<ctas, kArThreadsSmall, 0, stream>>>(a, off_bytes, n4, scale);
    else peer_allreduce_kernel<false, kArThreadsSmall, 1><<<ctas, kArThreadsSmall, 0, stream>>>(a, off_bytes, n4, scale);
  }
  B200_LAUNCH_OK();
  return 0;
}

}  // namespace b200
