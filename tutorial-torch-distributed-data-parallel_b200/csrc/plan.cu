// R1 — gradient bucket planner (host code, no CUDA calls).
//
// Restates c10d::compute_bucket_assignment_by_size (declared at
// torch/csrc/distributed/c10d/reducer.hpp:592-598, bound as
// torch.distributed._compute_bucket_assignment_by_size and driven from
// torch/nn/parallel/distributed.py:1219-1244): walk the tensors in the given order, keep one
// open bucket per (dtype, device) key, close a bucket as soon as its byte size reaches the
// current cap of that key (caps advance 1 MiB -> 25 MiB -> stays), flush the leftovers, and
// order the buckets by their smallest tensor index.
#include "common.cuh"
#include "../../include/b200ddp.h"

#include <algorithm>
#include <map>
#include <vector>

extern "C" int b200ddp_plan_buckets(const long long* nbytes, const int* keys, int n, const long long* caps,
                                    int ncaps, int* bucket_of, int* nbuckets) {
  using namespace b200;
  B200_CHECK_ARG(n >= 0 && ncaps >= 1 && nbytes && caps && bucket_of && nbuckets, "plan_buckets: bad argument");
  struct Open {
    std::vector<int> idx;
    long long size = 0;
    int cap_i = 0;
  };
  std::map<int, Open> open;
  std::vector<std::vector<int>> result;
  for (int i = 0; i < n; ++i) {
    B200_CHECK_ARG(nbytes[i] >= 0, "plan_buckets: negative size at %d", i);
    Open& b = open[keys ? keys[i] : 0];
    b.idx.push_back(i);
    b.size += nbytes[i];
    if (b.size >= caps[b.cap_i]) {
      result.push_back(std::move(b.idx));
      b.idx.clear();
      b.size = 0;
      if (b.cap_i + 1 < ncaps) ++b.cap_i;
    }
  }
  for (auto& kv : open)
    if (!kv.second.idx.empty()) result.push_back(std::move(kv.second.idx));
  std::sort(result.begin(), result.end(), [](const std::vector<int>& a, const std::vector<int>& b) {
    return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
  });
  for (size_t b = 0; b < result.size(); ++b)
    for (int i : result[b]) bucket_of[i] = (int)b;
  *nbuckets = (int)result.size();
  return 0;
}
