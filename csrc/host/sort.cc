// Parallel sort + unique of 64-bit keys (sample sort over the host thread pool).
//
// Role in the reference: the end of a feed pass — the key agent's per-thread key sets are merged and
// de-duplicated before the pass working set is built (PSAgentBase / BoxPSBase::EndFeedPass in the
// closed library; open twin PSGPUWrapper::PreBuildTask, fleet/ps_gpu_wrapper.cc:114-335, which
// shards the keys, sorts and uniques every shard on a thread pool).
#include <algorithm>
#include <random>

#include "util.h"

namespace pbh {
static ThreadPool* sort_pool() {
  static ThreadPool pool((int)std::max(2u, std::min(32u, std::thread::hardware_concurrency())));
  return &pool;
}
}  // namespace pbh

extern "C" {

// Sorts keys[0, n) ascending in place and removes duplicates; returns the number of unique keys
// (they occupy keys[0, ret)).
int64_t pbh_sort_unique_u64(uint64_t* keys, int64_t n) {
  using namespace pbh;
  if (n <= 0) return 0;
  ThreadPool* pool = sort_pool();
  const int W = pool->size();
  if (n < (1 << 16) || W < 2) {
    std::sort(keys, keys + n);
    return (int64_t)(std::unique(keys, keys + n) - keys);
  }
  // splitters from a sorted sample: P buckets of roughly equal size, bucket b < bucket b+1 key-wise
  const int P = W * 4;
  std::vector<uint64_t> sample;
  std::mt19937_64 rng(0x5eed);
  const int S = P * 64;
  sample.reserve(S);
  for (int i = 0; i < S; ++i) sample.push_back(keys[rng() % (uint64_t)n]);
  std::sort(sample.begin(), sample.end());
  std::vector<uint64_t> split;                 // bucket b holds keys in (split[b-1], split[b]]
  for (int b = 1; b < P; ++b) {
    uint64_t v = sample[(size_t)b * S / P];
    if (split.empty() || v != split.back()) split.push_back(v);
  }
  const int B = (int)split.size() + 1;
  auto bucket_of = [&](uint64_t k) { return (int)(std::lower_bound(split.begin(), split.end(), k) - split.begin()); };
  std::vector<std::vector<int64_t>> cnt(W, std::vector<int64_t>(B, 0));
  pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
    auto& c = cnt[w];
    for (size_t i = b; i < e; ++i) ++c[bucket_of(keys[i])];
  });
  std::vector<int64_t> start(B + 1, 0);
  std::vector<std::vector<int64_t>> at(W, std::vector<int64_t>(B, 0));
  for (int q = 0; q < B; ++q) {
    int64_t t = start[q];
    for (int w = 0; w < W; ++w) { at[w][q] = t; t += cnt[w][q]; }
    start[q + 1] = t;
  }
  std::vector<uint64_t> tmp((size_t)n);
  pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
    auto& a = at[w];
    for (size_t i = b; i < e; ++i) tmp[a[bucket_of(keys[i])]++] = keys[i];
  });
  std::vector<int64_t> uniq(B, 0);
  pool->parallel_range((size_t)B, [&](size_t qb, size_t qe, int) {
    for (size_t q = qb; q < qe; ++q) {
      uint64_t* lo = tmp.data() + start[q];
      uint64_t* hi = tmp.data() + start[q + 1];
      std::sort(lo, hi);
      uniq[q] = (int64_t)(std::unique(lo, hi) - lo);
    }
  });
  std::vector<int64_t> dst(B + 1, 0);
  for (int q = 0; q < B; ++q) dst[q + 1] = dst[q] + uniq[q];
  pool->parallel_range((size_t)B, [&](size_t qb, size_t qe, int) {
    for (size_t q = qb; q < qe; ++q) memcpy(keys + dst[q], tmp.data() + start[q], (size_t)uniq[q] * 8);
  });
  return dst[B];
}

// Groups keys by owning rank, owner = (mix64(key) >> 40) % nranks (pbx_owner_of, csrc/cuda/common.cuh):
// out[0, n) holds rank 0's keys, then rank 1's, ...; counts[r] = how many rank r owns.  Order inside a
// rank follows the input order.
int pbh_partition_by_owner(const uint64_t* keys, int64_t n, int nranks, uint64_t* out, int64_t* counts) {
  using namespace pbh;
  PBH_REQUIRE(nranks >= 1 && nranks <= 65536, "nranks %d out of range", nranks);
  ThreadPool* pool = sort_pool();
  const int W = pool->size();
  auto owner = [nranks](uint64_t k) { return (int)((mix64(k) >> 40) % (uint64_t)nranks); };
  std::vector<std::vector<int64_t>> cnt(W, std::vector<int64_t>(nranks, 0));
  pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
    auto& c = cnt[w];
    for (size_t i = b; i < e; ++i) ++c[owner(keys[i])];
  });
  std::vector<std::vector<int64_t>> at(W, std::vector<int64_t>(nranks, 0));
  int64_t t = 0;
  for (int r = 0; r < nranks; ++r) {
    int64_t first = t;
    for (int w = 0; w < W; ++w) { at[w][r] = t; t += cnt[w][r]; }
    counts[r] = t - first;
  }
  pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
    auto& a = at[w];
    for (size_t i = b; i < e; ++i) out[a[owner(keys[i])]++] = keys[i];
  });
  return 0;
}

}  // extern "C"
