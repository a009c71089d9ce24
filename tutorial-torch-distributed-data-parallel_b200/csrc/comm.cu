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
#include "peer.h"
#include "../../include/b200ddp.h"

#include <dlfcn.h>
#include <nccl.h>
#include <nccl_device.h>     // host-side types of NCCL 2.28's device API: ncclDevComm, ncclWindow_vidmem (no device code used)
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

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
  // optional (NCCL >= 2.27 / 2.28): user-buffer registration, symmetric windows, device communicators
  ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
  ncclResult_t (*MemFree)(void*) = nullptr;
  ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
  ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
  ncclResult_t (*CommWindowRegister)(ncclComm_t, void*, size_t, ncclWindow_t*, int) = nullptr;
  ncclResult_t (*CommWindowDeregister)(ncclComm_t, ncclWindow_t) = nullptr;
  ncclResult_t (*DevCommCreate)(ncclComm_t, ncclDevCommRequirements_t const*, ncclDevComm_t*) = nullptr;
  ncclResult_t (*DevCommDestroy)(ncclComm_t, ncclDevComm_t const*) = nullptr;
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
#define B200_OPT(field, name) *(void**)(&g_nccl.field) = dlsym(h, name);
  B200_OPT(MemAlloc, "ncclMemAlloc");
  B200_OPT(MemFree, "ncclMemFree");
  B200_OPT(CommRegister, "ncclCommRegister");
  B200_OPT(CommDeregister, "ncclCommDeregister");
  B200_OPT(CommWindowRegister, "ncclCommWindowRegister");
  B200_OPT(CommWindowDeregister, "ncclCommWindowDeregister");
  B200_OPT(DevCommCreate, "ncclDevCommCreate");
  B200_OPT(DevCommDestroy, "ncclDevCommDestroy");
#undef B200_OPT
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

// A buffer every rank allocated with ncclMemAlloc and registered collectively (arena_register of SURVEY 8(b)).
struct SymBuf {
  void* ptr = nullptr;
  size_t bytes = 0;
  ncclWindow_t win = nullptr;      // symmetric window (peer + multicast mappings), or null
  void* reg = nullptr;             // plain ncclCommRegister handle when no window could be made
  PeerMap map{};                   // valid when win != null
};

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ready[kMaxInflight];
  cudaEvent_t done[kMaxInflight];
  cudaEvent_t t0[kMaxInflight];    // timing pair around each bucket's collective (b200ddp_comm_bucket_times)
  long long last_n[kMaxInflight];
  int last_count = 0;
  bool timing = false;
  int inflight = 0;
  ncclRedOp_t avg_op;        // PreMulSum(1/nranks), created once
  bool have_avg_op = false;
  // peer-memory path (our own all-reduce kernel, peer_allreduce.cu)
  std::vector<SymBuf> bufs;
  SymBuf flags;
  SymBuf small;              // scratch of the small collectives (SyncBatchNorm statistics)
  unsigned int small_epoch = 0;
  bool peer_tried = false, peer_ok = false;
  ncclDevComm_t devcomm;
  bool have_devcomm = false;
  unsigned int epoch = 0;
  int ar_ctas = 64;
  int ar_big = 0;            // 1: 1024-thread CTAs that own their SM (B200DDP_AR_BIG)
  int algo = 0;              // 0 = NCCL PreMulSum (ring), 1 = own kernel over peer memory, 2 = NCCL ncclSum on the registered arena
};

}  // namespace b200

using namespace b200;

static size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Maps of a symmetric window: every LSA rank's copy (unicast) and the NVSwitch multicast address.
static int fill_map(Comm* c, SymBuf& b) {
  ncclWindow_vidmem w;
  B200_CUDA(cudaMemcpy(&w, b.win, sizeof(w), cudaMemcpyDefault));
  for (int r = 0; r < kPeerMaxRanks; ++r)
    b.map.peer[r] = r < c->nranks ? nccl::utility::add4G(w.lsaFlatBase, (int)(r * w.stride4G)) : nullptr;
  b.map.mc = nullptr;
  if (c->have_devcomm && c->devcomm.lsaMultimem.mcBasePtr)
    b.map.mc = reinterpret_cast<char*>(c->devcomm.lsaMultimem.mcBasePtr) + 4096ull * w.mcOffset4K;
  return 0;
}

// One-time set-up of the peer path: a small symmetric flag buffer + the device communicator (multicast handle).
// Collective: every rank runs it at the same point (first arena_alloc).
static int peer_setup(Comm* c) {
  c->peer_tried = true;
  if (!g_nccl.MemAlloc || !g_nccl.CommWindowRegister || !g_nccl.DevCommCreate) return 0;
  if (c->nranks < 2 || c->nranks > kPeerMaxRanks) return 0;
  const char* e = getenv("B200DDP_PEER");
  if (e && e[0] == '0') return 0;
  const size_t fbytes = round_up((size_t)kPeerMaxCtas * kPeerMaxRanks * sizeof(unsigned int), 2u << 20);
  B200_NCCL(g_nccl.MemAlloc(&c->flags.ptr, fbytes));
  c->flags.bytes = fbytes;
  B200_CUDA(cudaMemset(c->flags.ptr, 0, fbytes));
  B200_CUDA(cudaDeviceSynchronize());
  ncclResult_t r = g_nccl.CommWindowRegister(c->comm, c->flags.ptr, fbytes, &c->flags.win, NCCL_WIN_COLL_SYMMETRIC);
  if (r != ncclSuccess || !c->flags.win) {
    c->flags.win = nullptr;
    return 0;                                   // no symmetric memory on this system: NCCL collectives only
  }
  ncclDevCommRequirements_t reqs;
  memset(&reqs, 0, sizeof(reqs));
  const char* m = getenv("B200DDP_PEER_MULTICAST");
  reqs.lsaMultimem = !(m && m[0] == '0');
  r = g_nccl.DevCommCreate(c->comm, &reqs, &c->devcomm);
  if (r != ncclSuccess && reqs.lsaMultimem) {
    reqs.lsaMultimem = false;                   // no NVLS: peer loads / stores over NVLink instead
    r = g_nccl.DevCommCreate(c->comm, &reqs, &c->devcomm);
  }
  if (r != ncclSuccess) return 0;
  c->have_devcomm = true;
  if (c->devcomm.lsaSize != c->nranks || c->devcomm.lsaRank != c->rank) return 0;   // ranks outside one NVLink domain
  int rc = fill_map(c, c->flags);
  if (rc) return rc;
  // scratch of the small collectives: [2 parities][ranks][kSmallMaxFloats]
  const size_t sbytes = round_up(kSmallScratchBytes, 2u << 20);
  B200_NCCL(g_nccl.MemAlloc(&c->small.ptr, sbytes));
  c->small.bytes = sbytes;
  B200_CUDA(cudaMemset(c->small.ptr, 0, sbytes));
  B200_CUDA(cudaDeviceSynchronize());
  r = g_nccl.CommWindowRegister(c->comm, c->small.ptr, sbytes, &c->small.win, NCCL_WIN_COLL_SYMMETRIC);
  if (r != ncclSuccess || !c->small.win) {
    c->small.win = nullptr;
    return 0;
  }
  rc = fill_map(c, c->small);
  if (rc) return rc;
  // nobody may write a flag before everybody has zeroed theirs: one tiny collective as the rendezvous
  float* tmp = nullptr;
  B200_CUDA(cudaMalloc(&tmp, 256));
  B200_CUDA(cudaMemset(tmp, 0, 256));
  B200_NCCL(g_nccl.AllReduce(tmp, tmp, 1, ncclFloat32, ncclSum, c->comm, c->comm_stream));
  B200_CUDA(cudaStreamSynchronize(c->comm_stream));
  cudaFree(tmp);
  c->peer_ok = true;
  return 0;
}

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
    B200_CUDA(cudaEventCreate(&c->done[i]));       // timing-capable: pairs with t0[i] when timing is on
    B200_CUDA(cudaEventCreate(&c->t0[i]));
  }
  if (const char* e = getenv("B200DDP_AR_CTAS")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kPeerMaxCtas) c->ar_ctas = v;
  }
  if (const char* e = getenv("B200DDP_AR_BIG")) c->ar_big = atoi(e) != 0;
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
  if (c->have_devcomm && g_nccl.DevCommDestroy) g_nccl.DevCommDestroy(c->comm, &c->devcomm);
  c->bufs.push_back(c->flags);
  c->bufs.push_back(c->small);
  for (SymBuf& b : c->bufs) {
    if (b.win && g_nccl.CommWindowDeregister) g_nccl.CommWindowDeregister(c->comm, b.win);
    if (b.reg && g_nccl.CommDeregister) g_nccl.CommDeregister(c->comm, b.reg);
  }
  g_nccl.CommDestroy(c->comm);
  for (SymBuf& b : c->bufs)
    if (b.ptr && g_nccl.MemFree) g_nccl.MemFree(b.ptr);
  c->bufs.clear();
  for (int i = 0; i < kMaxInflight; ++i) {
    cudaEventDestroy(c->ready[i]);
    cudaEventDestroy(c->done[i]);
    cudaEventDestroy(c->t0[i]);
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

static const SymBuf* find_buf(const Comm* c, const float* buf, long long n) {
  const char* lo = reinterpret_cast<const char*>(buf);
  for (const SymBuf& b : c->bufs)
    if (b.win && lo >= reinterpret_cast<const char*>(b.ptr) && lo + 4 * n <= reinterpret_cast<const char*>(b.ptr) + b.bytes)
      return &b;
  return nullptr;
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
  if (c->timing) B200_CUDA(cudaEventRecord(c->t0[slot], c->comm_stream));
  int rc = 0;
  const SymBuf* sb = (c->algo == 1 && c->peer_ok) ? find_buf(c, buf, n) : nullptr;
  if (sb && (n & 3) == 0 && ((reinterpret_cast<char*>(buf) - reinterpret_cast<char*>(sb->ptr)) & 15) == 0) {
    // our kernel: in-switch (multimem) or peer-to-peer reduction of the owner's shard, averaging folded in
    PeerArgs a;
    a.data = sb->map;
    a.flags = c->flags.map;
    a.rank = c->rank;
    a.nranks = c->nranks;
    a.epoch = c->epoch;
    c->epoch += 2;
    rc = peer_allreduce_launch(a, reinterpret_cast<char*>(buf) - reinterpret_cast<char*>(sb->ptr), n, premul, c->ar_ctas,
                               c->ar_big, c->comm_stream);
  } else {
    rc = allreduce_on(c, buf, n, premul, c->comm_stream);
  }
  if (rc) return rc;
  B200_CUDA(cudaEventRecord(c->done[slot], c->comm_stream));
  c->last_n[slot] = n;
  c->inflight = slot + 1;
  return 0;
}

// Make `stream` wait for the most recently launched bucket collective (per-bucket optimizer launches).
int b200ddp_bucket_wait_last(void* h, void* stream) {
  B200_CHECK_ARG(h, "bucket_wait_last: null handle");
  Comm* c = reinterpret_cast<Comm*>(h);
  if (c->inflight == 0) return 0;
  B200_DEVICE(c->device);
  B200_CUDA(cudaStreamWaitEvent(as_stream(stream), c->done[c->inflight - 1], 0));
  return 0;
}

int b200ddp_bucket_reduce_join(void* h, void* compute_stream) {
  B200_CHECK_ARG(h, "bucket_reduce_join: null handle");
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_DEVICE(c->device);
  for (int i = 0; i < c->inflight; ++i)
    B200_CUDA(cudaStreamWaitEvent(as_stream(compute_stream), c->done[i], 0));
  c->last_count = c->timing ? c->inflight : 0;
  c->inflight = 0;
  return 0;
}

// Small latency-bound collective over peer memory (B1: SyncBatchNorm statistics): sum_mode 0 = all-gather, out
// [nranks][n]; 1 = sum over ranks in rank order times `scale`, out [n] (bit-identical on every rank). n <= 2048 floats,
// on the CALLER's stream. Returns kNotInitialised when the peer path does not exist (the caller then uses NCCL).
int b200ddp_comm_small_collective(void* h, const float* send, float* out, int n, int sum_mode, float scale, void* stream) {
  B200_CHECK_ARG(h && send && out, "comm_small_collective: null argument");
  Comm* c = reinterpret_cast<Comm*>(h);
  if (!c->peer_tried) {
    B200_DEVICE(c->device);
    int rc = peer_setup(c);
    if (rc) return rc;
  }
  if (!c->peer_ok || !c->small.win || n > kSmallMaxFloats) {
    set_error("comm_small_collective: peer path unavailable or message too large (%d floats)", n);
    return kNotInitialised;
  }
  B200_DEVICE(c->device);
  PeerArgs a;
  a.data = c->small.map;
  a.flags = c->flags.map;
  a.rank = c->rank;
  a.nranks = c->nranks;
  a.epoch = ++c->small_epoch;
  return peer_small_collective_launch(a, send, out, n, sum_mode, scale, as_stream(stream));
}

// ---- arena_register: the gradient arena as an NCCL-registered, peer-mapped buffer -------------------
// Collective over the communicator. *mode_out: 0 = device memory registered with NCCL (zero-copy collectives),
// 1 = symmetric window, peer-to-peer mappings, 2 = symmetric window with an NVSwitch multicast mapping as well.
int b200ddp_comm_arena_alloc(void* h, long long nbytes, void** ptr_out, int* mode_out) {
  B200_CHECK_ARG(h && ptr_out && nbytes > 0, "comm_arena_alloc: bad argument");
  Comm* c = reinterpret_cast<Comm*>(h);
  B200_CHECK_ARG(g_nccl.MemAlloc && g_nccl.MemFree, "comm_arena_alloc: this NCCL has no ncclMemAlloc");
  B200_DEVICE(c->device);
  if (!c->peer_tried) {
    int rc = peer_setup(c);
    if (rc) return rc;
  }
  SymBuf b;
  b.bytes = round_up((size_t)nbytes, 2u << 20);
  B200_NCCL(g_nccl.MemAlloc(&b.ptr, b.bytes));
  B200_CUDA(cudaMemset(b.ptr, 0, b.bytes));
  B200_CUDA(cudaDeviceSynchronize());
  int mode = 0;
  if (c->peer_ok) {
    ncclResult_t r = g_nccl.CommWindowRegister(c->comm, b.ptr, b.bytes, &b.win, NCCL_WIN_COLL_SYMMETRIC);
    if (r == ncclSuccess && b.win) {
      int rc = fill_map(c, b);
      if (rc) return rc;
      mode = b.map.mc ? 2 : 1;
    } else {
      b.win = nullptr;
    }
  }
  if (!b.win && g_nccl.CommRegister) {
    if (g_nccl.CommRegister(c->comm, b.ptr, b.bytes, &b.reg) != ncclSuccess) b.reg = nullptr;
  }
  c->bufs.push_back(b);
  *ptr_out = b.ptr;
  if (mode_out) *mode_out = mode;
  return 0;
}

// algo: 0 = NCCL all-reduce with PreMulSum (ring), 1 = our peer-memory kernel (needs an arena from comm_arena_alloc in
// mode >= 1; other buffers fall back to NCCL), 2 = as 0; the caller passes pre-scaled data with premul == 1 so that
// NCCL may pick its fastest algorithm for plain sums (NVLS) on the registered arena. Returns 1 when the peer path exists.
int b200ddp_comm_set_algo(void* h, int algo, int ctas) {
  if (!h || algo < 0 || algo > 2) return -1;
  Comm* c = reinterpret_cast<Comm*>(h);
  c->algo = algo;
  if (ctas >= 1 && ctas <= kPeerMaxCtas) c->ar_ctas = ctas;
  return c->peer_ok ? 1 : 0;
}

// CTA shape of the own all-reduce kernel: 0 = 256 threads (co-resident with compute CTAs), 1 = 1024 threads (owns its SM).
int b200ddp_comm_set_cta_shape(void* h, int big) {
  if (!h) return -1;
  reinterpret_cast<Comm*>(h)->ar_big = big != 0;
  return 0;
}

// Per-bucket device time of the collectives of the most recently JOINED backward pass (timing must be on and the
// caller must have synchronised the device). Returns the number of entries written.
int b200ddp_comm_set_timing(void* h, int on) {
  B200_CHECK_ARG(h, "comm_set_timing: null handle");
  reinterpret_cast<Comm*>(h)->timing = on != 0;
  return 0;
}
int b200ddp_comm_bucket_times(void* h, float* ms_out, long long* n_out, int max_entries) {
  if (!h || !ms_out || max_entries < 0) return -1;
  Comm* c = reinterpret_cast<Comm*>(h);
  int k = 0;
  for (; k < c->last_count && k < max_entries; ++k) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->t0[k], c->done[k]) != cudaSuccess) { (void)cudaGetLastError(); ms = -1.f; }
    ms_out[k] = ms;
    if (n_out) n_out[k] = c->last_n[k];
  }
  return k;
}

}  // extern "C"
