// Shared helpers for libb200ddp.so (sm_100a only).
//
// Conventions of the C ABI (see include/b200ddp.h):
//   * every entry point returns 0 on success, non-zero on failure, and stores a
//     human readable message retrievable through b200ddp_last_error();
//   * all buffers are borrowed device pointers, all work is stream ordered,
//     nothing synchronises the device behind the caller's back;
//   * every kernel launch of this library bumps a process-wide counter so the
//     bench can report `gpu_launches` from evidence instead of by hand.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

namespace b200 {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs. Sizes fixed-size scratch layouts only; launch grids use num_sms()
int num_sms();                // SM count of the CURRENT device (cudaDevAttrMultiProcessorCount, cached per device)

// Error codes. CUDA runtime errors are offset so they stay disjoint from NCCL ones.
enum : int {
  kOk = 0,
  kInvalidArg = 1,
  kNotInitialised = 2,
  kCudaErrBase = 1000,
  kNcclErrBase = 2000,
  kDriverErrBase = 3000,
};

void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;
// SMs the grid planners may count on (b200ddp_set_sm_budget): all 148 unless a resident collective owns some.
extern std::atomic<int> g_sm_budget;
inline int sm_budget() {
  const int b = g_sm_budget.load(std::memory_order_relaxed), n = num_sms();
  return (b <= 0 || b > n) ? n : b;       // 0 = every SM of the device
}

inline void count_launch(int n = 1) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

#define B200_CHECK_ARG(cond, ...)                \
  do {                                           \
    if (!(cond)) {                               \
      ::b200::set_error(__VA_ARGS__);            \
      return ::b200::kInvalidArg;                \
    }                                            \
  } while (0)

#define B200_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return ::b200::kCudaErrBase + (int)_e;                                              \
    }                                                                                     \
  } while (0)

// Launch-status check after a <<<>>> launch (does not synchronise).
#define B200_LAUNCH_OK()                                                                  \
  do {                                                                                    \
    cudaError_t _e = cudaGetLastError();                                                  \
    if (_e != cudaSuccess) {                                                              \
      ::b200::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return ::b200::kCudaErrBase + (int)_e;                                              \
    }                                                                                     \
    ::b200::count_launch();                                                               \
  } while (0)

// RAII: make `device` current for the duration of a call (hooks run on autograd
// worker threads, so no thread-local CUDA state may be assumed).
struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) {
      err = cudaSetDevice(device);
      changed = (err == cudaSuccess);
    }
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};

#define B200_DEVICE(device)                                                    \
  ::b200::DeviceGuard _guard(device);                                          \
  if (_guard.err != cudaSuccess) {                                             \
    ::b200::set_error("cudaSetDevice(%d) -> %s", device, cudaGetErrorString(_guard.err)); \
    return ::b200::kCudaErrBase + (int)_guard.err;                             \
  }

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute of a kernel: remember, per device ordinal,
// whether it has been raised (one bit per device in `done_mask`, one mask per kernel instantiation).
template <typename F>
inline cudaError_t ensure_max_dyn_smem(F* func, std::atomic<unsigned long long>& done_mask, int bytes = 227 * 1024) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const unsigned long long bit = 1ull << (dev & 63);
  if (done_mask.load(std::memory_order_acquire) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done_mask.fetch_or(bit, std::memory_order_release);
  return e;
}

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32). `red` holds >= 32 floats.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) red[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  T r = (threadIdx.x < nw) ? red[threadIdx.x] : T(0);
  if (wid == 0) r = warp_sum(r);
  if (threadIdx.x == 0) red[0] = r;
  __syncthreads();
  r = red[0];
  __syncthreads();
  return r;
}

// Streaming 128-bit accesses: data touched once per step should not pollute L1.
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld_f4(const float4* p) { return *p; }
__device__ __forceinline__ void st_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// Philox4x32-10 (Salmon et al. 2011). Counter = (idx_lo, idx_hi, 0, 0), key = seed.
// Geometry independent: element i of a tensor always uses counter i/4, lane i%4.
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)M0 * c0;
    const uint64_t p1 = (uint64_t)M1 * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// keep-decision for dropout element `idx`: uniform u = bits * 2^-32, keep iff u >= p.
__host__ __device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, uint32_t p_threshold) {
  uint32_t r[4];
  const uint64_t ctr = idx >> 2;
  philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), r);
  return r[idx & 3] >= p_threshold;
}

}  // namespace b200
