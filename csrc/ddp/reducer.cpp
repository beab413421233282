// Host-side core of the gradient bucket reducer: bucket planning and per-iteration ready tracking.
//
// Behavioural model: the library Reducer that `loss.backward()` drives under
// DistributedDataParallel in the reference (ref:main.py:62,122): parameters are packed into flat
// buckets in gradient-ready order with a small first bucket (1 MiB) and a cap (25 MiB) for the rest;
// a bucket is handed to the collective as soon as all of its gradients have landed, and buckets are
// launched strictly in order so every rank issues the same sequence.  The device work (fused
// scale + cast + P2P all-reduce on a side stream) lives in csrc/coll; this file is pure bookkeeping
// so it can be tested on a CPU-only box.
#include <algorithm>
#include <cstdint>
#include <map>
#include <vector>

extern "C" {

// Plan buckets.  `nbytes[i]` / `group_key[i]` describe tensor i in the order given (dtype+device folded
// into group_key by the caller).  `limits` is the per-group sequence of size limits; the last entry
// repeats.  A bucket is closed once its size reaches the current limit.  Buckets are numbered by
// the position of their first tensor.  Returns the number of buckets; out_bucket_of[i] = bucket of i.
int dcp_bucket_assign(const int64_t* nbytes, const int32_t* group_key, int n, const int64_t* limits, int nlimits,
                      int32_t* out_bucket_of) {
  struct Open { std::vector<int> idx; int64_t size = 0; int limit_pos = 0; };
  std::map<int32_t, Open> open;
  std::vector<std::vector<int>> closed;
  for (int i = 0; i < n; ++i) {
    Open& b = open[group_key[i]];
    b.idx.push_back(i);
    b.size += nbytes[i];
    int64_t lim = nlimits > 0 ? limits[std::min(b.limit_pos, nlimits - 1)] : INT64_MAX;
    if (b.size >= lim) {
      closed.push_back(std::move(b.idx));
      b.idx.clear(); b.size = 0;
      if (b.limit_pos + 1 < nlimits) ++b.limit_pos;
    }
  }
  for (auto& kv : open) if (!kv.second.idx.empty()) closed.push_back(std::move(kv.second.idx));
  std::sort(closed.begin(), closed.end(), [](const std::vector<int>& a, const std::vector<int>& b) {
    return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
  });
  for (size_t b = 0; b < closed.size(); ++b) for (int i : closed[b]) out_bucket_of[i] = (int32_t)b;
  return (int)closed.size();
}

struct Reducer {
  std::vector<int32_t> bucket_of;
  std::vector<int32_t> bucket_total, bucket_pending;
  std::vector<uint8_t> ready;
  std::vector<int32_t> order;  // parameter indices in the order their gradients arrived
  int next_bucket = 0;
};

void* dcp_reducer_create(const int32_t* bucket_of, int n, int nbuckets) {
  auto* r = new Reducer;
  r->bucket_of.assign(bucket_of, bucket_of + n);
  r->bucket_total.assign(nbuckets, 0);
  for (int i = 0; i < n; ++i) {
    if (bucket_of[i] < 0 || bucket_of[i] >= nbuckets) { delete r; return nullptr; }
    r->bucket_total[bucket_of[i]]++;
  }
  r->bucket_pending = r->bucket_total;
  r->ready.assign(n, 0);
  return r;
}
void dcp_reducer_destroy(void* h) { delete static_cast<Reducer*>(h); }

// Start a new backward pass.
void dcp_reducer_reset(void* h) {
  auto* r = static_cast<Reducer*>(h);
  r->bucket_pending = r->bucket_total;
  std::fill(r->ready.begin(), r->ready.end(), 0);
  r->order.clear();
  r->next_bucket = 0;
}

// Gradient of parameter `idx` has landed.  Returns how many buckets became launchable (in order)
// and writes the first of them to *first; -1 if idx was already marked this iteration.
int dcp_reducer_mark_ready(void* h, int idx, int* first) {
  auto* r = static_cast<Reducer*>(h);
  if (idx < 0 || idx >= (int)r->ready.size() || r->ready[idx]) return -1;
  r->ready[idx] = 1;
  r->order.push_back(idx);
  r->bucket_pending[r->bucket_of[idx]]--;
  int start = r->next_bucket, k = 0;
  while (r->next_bucket < (int)r->bucket_total.size() && r->bucket_pending[r->next_bucket] == 0) { ++r->next_bucket; ++k; }
  if (first) *first = start;
  return k;
}

// Number of buckets not yet launched this iteration.
int dcp_reducer_unlaunched(void* h) {
  auto* r = static_cast<Reducer*>(h);
  return (int)r->bucket_total.size() - r->next_bucket;
}

int dcp_reducer_ready_order(void* h, int32_t* out, int cap) {
  auto* r = static_cast<Reducer*>(h);
  int n = (int)std::min<size_t>(r->order.size(), (size_t)cap);
  std::copy(r->order.begin(), r->order.begin() + n, out);
  return (int)r->order.size();
}

}  // extern "C"
