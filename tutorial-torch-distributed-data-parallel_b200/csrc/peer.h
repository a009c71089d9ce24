// Private interface between comm.cu (NCCL bootstrap: symmetric windows, multicast handle) and peer_allreduce.cu
// (the all-reduce kernels that run over those mappings).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

constexpr int kPeerMaxRanks = 8;        // one NVSwitch box
constexpr int kPeerMaxCtas = 256;       // CTAs of one all-reduce launch (flag slots per rank)

// Where every rank's copy of a symmetric buffer is mapped in THIS process.
struct PeerMap {
  char* peer[kPeerMaxRanks];    // unicast address of rank r's buffer base (peer[rank] is the local one)
  char* mc;                     // multicast address of the buffer base (NVSwitch multimem), null if none
};

// flags: a second, small symmetric buffer: inbox[cta][src_rank] of uint32 epochs, written by peers.
struct PeerArgs {
  PeerMap data;
  PeerMap flags;
  int rank, nranks;
  unsigned int epoch;           // the two barriers of this launch use epoch + 1 and epoch + 2
};

// In-place sum over ranks of buf[off_bytes .. off_bytes + 4 n) (n % 4 == 0, off_bytes % 16 == 0) of the symmetric
// buffer described by a.data; every element is multiplied by `scale` after the sum (1/world = the mean).
// big_ctas = 0: `ctas` CTAs of 256 threads that co-reside with compute CTAs; 1: CTAs of 1024 threads that own their SM.
int peer_allreduce_launch(const PeerArgs& a, long long off_bytes, long long n, float scale, int ctas, int big_ctas,
                          cudaStream_t stream);

// Small, latency-bound collectives (SyncBatchNorm statistics, B1) over the same mappings: every rank deposits its
// `n` floats (n <= kSmallMaxFloats) in slot [parity][rank] of every rank's scratch buffer, one flag round later every
// rank holds all contributions. out: [nranks][n] (sum_mode 0, all-gather) or [n] (sum_mode 1: sum in rank order, times
// `scale` - identical bits on every rank). One CTA; a.data = the scratch buffer's map, a.epoch = this call's epoch.
constexpr int kSmallMaxFloats = 2048;
constexpr size_t kSmallScratchBytes = 2ull * kPeerMaxRanks * kSmallMaxFloats * sizeof(float);
int peer_small_collective_launch(const PeerArgs& a, const float* send, float* out, int n, int sum_mode, float scale,
                                 cudaStream_t stream);

}  // namespace b200
