// DDP bucket planner (native half of the Reducer, W5 in SURVEY.md section 8).
//
// The reference wraps its model in torch's DistributedDataParallel with default arguments
// (mnist_distributed.py:67); the bucketing that follows is done by torch's C++ Reducer
// (compute_bucket_assignment_by_size: first bucket capped at 1 MiB, the rest at 25 MiB,
// tensors grouped by dtype, a bucket is closed as soon as it reaches its cap, buckets
// rebuilt in gradient-arrival order after the first iteration).  This file re-implements
// that planning rule so the framework reproduces the reference's steady-state layout
// (720,000,040 B + 53,376 B for the sandbox CNN) and can plan arbitrary
// models; the Python side (parallel/ddp.py) owns the hooks and the flat buffers.
#include <cstdint>
#include <map>
#include <vector>

#define B200H_API extern "C" __attribute__((visibility("default")))

// sizes[i]: bytes of tensor i (in the order the caller wants them packed, i.e. expected
// gradient-arrival order); keys[i]: grouping key (dtype/device code); limits: per-key
// sequence of caps (the last one repeats).  bucket_of[i] receives the bucket id; ids are
// assigned in closing order.  Returns the number of buckets.
B200H_API int b200h_plan_buckets(const int64_t* sizes, const int32_t* keys, int n, const int64_t* limits,
                                 int nlimits, int32_t* bucket_of) {
  struct Acc { std::vector<int> idx; int64_t bytes = 0; int limit_pos = 0; };
  std::map<int32_t, Acc> open;
  int next_id = 0;
  auto close_bucket = [&](Acc& a) {
    for (int i : a.idx) bucket_of[i] = next_id;
    ++next_id;
    a.idx.clear();
    a.bytes = 0;
  };
  for (int i = 0; i < n; ++i) {
    Acc& a = open[keys ? keys[i] : 0];
    a.idx.push_back(i);
    a.bytes += sizes[i];
    const int64_t cap = nlimits > 0 ? limits[a.limit_pos < nlimits ? a.limit_pos : nlimits - 1] : INT64_MAX;
    if (a.bytes >= cap) {
      close_bucket(a);
      if (a.limit_pos + 1 < nlimits) ++a.limit_pos;
    }
  }
  for (auto& kv : open)
    if (!kv.second.idx.empty()) close_bucket(kv.second);
  return next_id;
}

// Flat-buffer layout of one bucket: offsets (in elements) with every tensor start aligned
// to `align_elems` (16 B for float4 / multimem.v4 access).  Returns the padded total.
B200H_API int64_t b200h_bucket_offsets(const int64_t* numels, int n, int64_t align_elems, int64_t* offsets) {
  int64_t off = 0;
  for (int i = 0; i < n; ++i) {
    offsets[i] = off;
    off += numels[i];
    if (align_elems > 1) off = (off + align_elems - 1) / align_elems * align_elems;
  }
  return off;
}
