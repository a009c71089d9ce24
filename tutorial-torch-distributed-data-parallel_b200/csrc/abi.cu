// Error reporting, version and launch accounting for libb200ddp.so.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

#include "../../include/b200ddp.h"

namespace b200 {

static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launches{0};
std::atomic<int> g_sm_budget{0};          // 0 = all SMs of the current device

int num_sms() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return kNumSMs;
  int n = cache[dev & 63].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return kNumSMs;
  cache[dev & 63].store(n, std::memory_order_relaxed);
  return n;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace b200

extern "C" {

const char* b200ddp_last_error(void) { return b200::g_err; }

int b200ddp_version(void) { return 100; }

unsigned long long b200ddp_launch_count(void) { return b200::g_launches.load(); }

int b200ddp_set_sm_budget(int sms) {
  const int prev = b200::sm_budget();
  if (sms <= 0) sms = 0;                   // all SMs of whichever device a call runs on
  else if (sms < 16) sms = 16;
  b200::g_sm_budget.store(sms);
  return prev;
}

int b200ddp_device_info(int device, int* sm_count, int* cc_major, int* cc_minor) {
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, device));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  return 0;
}

}  // extern "C"
