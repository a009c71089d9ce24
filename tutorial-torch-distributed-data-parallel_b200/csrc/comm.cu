// R4 — gradient all-reduce per bucket over NCCL (NVLink 5 / NVSwitch), called directly
// from this library on a dedicated high-priority communication stream.
//
// Restates c10d::Reducer::all_reduce_bucket + _AllReduceBySumCommHook
// (torch/csrc/distributed/c10d/reducer.hpp:276-286, default_comm_hooks.hpp:9-51) with the
// 1/world_size averaging folded INTO the collective (ncclRedOpCreatePreMulSum, nccl.h:327-337)
// instead of a separate pre-divide pass over the bucket.
//
// NCCL is not linked: the library dlopen()s the libnccl.so.2 that torch already loaded so
// both share one NCCL instance.
#include "common.cuh"
#include "../../include/b200ddp.h"

#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

#include <mutex>

namespace b200 {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*RedOpCreatePreMulSum)(ncclRedOp_t*, void*, ncclDataType_t, ncclScalarResidence_t, ncclComm_t) = nullptr;
  ncclResult_t (*RedOpDestroy)(ncclRedOp_t, ncclComm_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int load_nccl(const char* path) {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.lib) return kOk;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
  if (!h && path && path[0]) h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    set_error("cannot dlopen libnccl.so.2 (%s)", dlerror());
    return kNotInitialised;
  }
#define B200_SYM(field, name)                                     \
  *(void**)(&g_nccl.field) = dlsym(h, name);                      \
  if (!g_nccl.field) {                                            \
    set_error("libnccl.so.2 lacks symbol %s", name);              \
    return kNotInitialised;                                       \
  }
  B200_SYM(GetUniqueId, "ncclGetUniqueId");
  B200_SYM(CommInitRank, "ncclCommInitRank");
  B200_SYM(CommDestroy, "ncclCommDestroy");
  B200_SYM(CommAbort, "ncclCommAbort");
  B200_SYM(AllReduce, "ncclAllReduce");
  B200_SYM(AllGather, "ncclAllGather");
  B200_SYM(Broadcast, "ncclBroadcast");
  B200_SYM(RedOpCreatePreMulSum, "ncclRedOpCreatePreMulSum");
  B200_SYM(RedOpDestroy, "ncclRedOpDestroy");
  B200_SYM(GetErrorString, "ncclGetErrorString");
  B200_SYM(GetVersion, "ncclGetVersion");
#undef B200_SYM
  g_nccl.lib = h;
  return kOk;
}

#define B200_NCCL(expr)                                                                       \
  do {                                                                                        \
    ncclResult_t _r = (expr);                                                                 \
    if (_r != ncclSuccess) {                                                                  \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(_r)); \
      return ::b200::kNcclErrBase + (int)_r;                                                  \
    }                                                                                         \
  } while (0)

constexpr int kMaxInflight = 64;

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ready[kMaxInflight];
  cudaEvent_t done[kMaxInflight];
  int inflight = 0;
  ncclRedOp_t avg_op;        // PreMulSum(1/nranks), created once
  bool have_avg_op = false;
};

}  // namespace b200

using namespace b200;

extern "C" {

int b200ddp_nccl_load(const char* path) { return load_nccl(path); }

int b200ddp_nccl_version(int* version) {
  int rc = load_nccl(nullptr);
  if (rc) return rc;
  B200_NCCL(g_nccl.GetVersion(version));
  return 0;
}

int b200ddp_comm_unique_id(void* out, int nbytes) {
  B200_CHECK_ARG(out && nbytes >= (int)sizeof(ncclUniqueId), "comm_unique_id: need %d bytes",
                 (int)sizeof(ncclUniqueId));
  int rc = load_nccl(nullptr);
  if (rc) return rc;
  ncclUniqueId id;
  B200_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(out, &id, sizeof(id));
  return 0;
}

int b200ddp_comm_create(void** out, const void* id_bytes, int rank, int nranks, int device) {
  B200_CHECK_ARG(out && id_bytes, "comm_create: null argument");
  B200_CHECK_ARG(nranks >= 1 && rank >= 0 && rank < nranks, "comm_create: bad rank %d / %d", rank, nranks);
  int rc = load_nccl(nullptr);
  if (rc) return rc;
  B200_DEVICE(device);
  Comm* c = new Comm();
  c->rank = rank; c->nranks = nranks; c->device = device;
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank -> %s", g_nccl.GetErrorString(r));
    delete c;
    return kNcclErrBase + (int)r;
  }
  int lo = 0, hi = 0;
  B200_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  B200_CUDA(cudaStreamCreateWithPriority(&c->comm_stream, cudaStreamNonBlocking, hi));
  for (int i = 0; i < kMaxInflight; ++i) {
    B200_CUDA(cudaEventCreateWithFlags(&c->ready[i], cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming));
  }
  float inv = 1.0f / (float)nranks;
  r = g_nccl.RedOpCreatePreMulSum(&c->avg_op, &inv, ncclFloat32, ncclScalarHostImmediate, c->comm);
  c->have_avg_op = (r == ncclSuccess);
  *out = c;
  return 0;
}

int b200ddp_comm_destroy(void* h) {
  if (!h) return 0;
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  cudaStreamSynchronize(c->comm_stream);
  if (c->have_avg_op) g_nccl.RedOpDestroy(c->avg_op, c->comm);
  g_nccl.CommDestroy(c->comm);
  for (int i = 0; i < kMaxInflight; ++i) {
    cudaEventDestroy(c->ready[i]);
    cudaEventDestroy(c->done[i]);
  }
  cudaStreamDestroy(c->comm_stream);
  delete c;
  return 0;
}

int b200ddp_comm_rank(void* h, int* rank, int* nranks) {
  B200_CHECK_ARG(h, "comm_rank: null handle");
  Comm* c = reinterpret_cast<Comm*>(h);
  if (rank) *rank = c->rank;
  if (nranks) *nranks = c->nranks;
  return 0;
}

static int allreduce_on(Comm* c, float* buf, long long n, float premul, cudaStream_t s) {
  if (premul == 1.0f) {
    B200_NCCL(g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat32, ncclSum, c->comm, s));
  } else if (c->have_avg_op && premul == 1.0f / (float)c->nranks) {
    B200_NCCL(g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat32, c->avg_op, c->comm, s));
  } else {
    ncclRedOp_t op;
    B200_NCCL(g_nccl.RedOpCreatePreMulSum(&op, &premul, ncclFloat32, ncclScalarHostImmediate, c->comm));
    ncclResult_t r = g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat32, op, c->comm, s);
    g_nccl.RedOpDestroy(op, c->comm);
    B200_NCCL(r);
  }
  return 0;
}

// In-place sum (premul applied to every rank's contribution first) on the CALLER's stream.
int b200ddp_comm_allreduce(void* h, float* buf, long long n, float premul, void* stream) {
  B200_CHECK_ARG(h && buf && n >= 0, "comm_allreduce: bad argument");
  if (n == 0) return 0;
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  return allreduce_on(c, buf, n, premul, as_stream(stream));
}

int b200ddp_comm_allgather(void* h, const float* send, float* recv, long long n_per_rank, void* stream) {
  B200_CHECK_ARG(h && send && recv && n_per_rank >= 0, "comm_allgather: bad argument");
  if (n_per_rank == 0) return 0;
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  B200_NCCL(g_nccl.AllGather(send, recv, (size_t)n_per_rank, ncclFloat32, c->comm, as_stream(stream)));
  return 0;
}

int b200ddp_comm_broadcast(void* h, void* buf, long long nbytes, int root, void* stream) {
  B200_CHECK_ARG(h && buf && nbytes >= 0, "comm_broadcast: bad argument");
  if (nbytes == 0) return 0;
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  B200_NCCL(g_nccl.Broadcast(buf, buf, (size_t)nbytes, ncclChar, root, c->comm, as_stream(stream)));
  return 0;
}

// Bucket all-reduce on the communication stream: the bucket becomes valid at the current tail
// of `compute_stream`; the collective runs concurrently with whatever the caller enqueues next
// (the rest of backward). b200ddp_bucket_reduce_join() makes `compute_stream` wait for all of them.
int b200ddp_bucket_reduce_launch(void* h, float* buf, long long n, float premul, void* compute_stream) {
  B200_CHECK_ARG(h && buf && n >= 0, "bucket_reduce_launch: bad argument");
  if (n == 0) return 0;
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_CHECK_ARG(c->inflight < kMaxInflight, "bucket_reduce_launch: more than %d buckets in flight", kMaxInflight);
  B200_DEVICE(c->device);
  const int slot = c->inflight;
  B200_CUDA(cudaEventRecord(c->ready[slot], as_stream(compute_stream)));
  B200_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ready[slot], 0));
  int rc = allreduce_on(c, buf, n, premul, c->comm_stream);
  if (rc) return rc;
  B200_CUDA(cudaEventRecord(c->done[slot], c->comm_stream));
  c->inflight = slot + 1;
  return 0;
}

int b200ddp_bucket_reduce_join(void* h, void* compute_stream) {
  B200_CHECK_ARG(h, "bucket_reduce_join: null handle");
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  for (int i = 0; i < c->inflight; ++i)
    B200_CUDA(cudaStreamWaitEvent(as_stream(compute_stream), c->done[i], 0));
  c->inflight = 0;
  return 0;
}

}  // extern "C"
