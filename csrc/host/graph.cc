// Graph store for graph-embedding training (DeepWalk / metapath2vec style): typed edge lists in CSR,
// uniform neighbour sampling and (metapath) random walks.
//
// Role in the reference: the GPU graph engine — GpuPsGraphTable / GraphGpuWrapper (load_edge_file,
// load_node_file, graph_neighbor_sample, get_all_id; heter_ps/graph_gpu_wrapper.h:34-99,
// graph_gpu_ps_table_inl.cu) and the walk generator GraphDataGenerator (data_feed.h:892-975,
// data_feed.cu:200-1206).  Here the store is built on the host (this file) and either sampled on the
// host with the thread pool, or exported as CSR arrays that paddlebox_b200/graph.py uploads to HBM
// and samples there; node embeddings live in the same sparse table as every other feasign.
//
// Node ids are arbitrary uint64 values; inside an edge type they are mapped to dense indices in
// sorted-id order, so `idx = lower_bound(node_ids, id)` works on the host and as a searchsorted on
// the device.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <unordered_map>

#include "util.h"

namespace pbh {

struct EdgeType {
  std::vector<uint64_t> src, dst;        // raw edge list until build()
  std::vector<uint64_t> node_ids;        // sorted unique ids (sources and destinations)
  std::vector<int64_t> row_ptr;          // [n_nodes + 1]
  std::vector<int64_t> col;              // neighbour node INDEX per edge
  bool built = false;

  int64_t index_of(uint64_t id) const {
    auto it = std::lower_bound(node_ids.begin(), node_ids.end(), id);
    return (it != node_ids.end() && *it == id) ? (int64_t)(it - node_ids.begin()) : -1;
  }

  void build() {
    node_ids = src;
    node_ids.insert(node_ids.end(), dst.begin(), dst.end());
    std::sort(node_ids.begin(), node_ids.end());
    node_ids.erase(std::unique(node_ids.begin(), node_ids.end()), node_ids.end());
    const size_t n = node_ids.size(), m = src.size();
    row_ptr.assign(n + 1, 0);
    std::vector<int64_t> si(m), di(m);
    for (size_t e = 0; e < m; ++e) {
      si[e] = index_of(src[e]);
      di[e] = index_of(dst[e]);
      ++row_ptr[si[e] + 1];
    }
    for (size_t i = 0; i < n; ++i) row_ptr[i + 1] += row_ptr[i];
    col.assign(m, 0);
    std::vector<int64_t> fill(row_ptr.begin(), row_ptr.end() - 1);
    for (size_t e = 0; e < m; ++e) col[fill[si[e]]++] = di[e];
    for (size_t i = 0; i < n; ++i) std::sort(col.begin() + row_ptr[i], col.begin() + row_ptr[i + 1]);
    src.clear(); src.shrink_to_fit();
    dst.clear(); dst.shrink_to_fit();
    built = true;
  }
};

struct GraphStore {
  std::vector<EdgeType> types;
  EdgeType& type(int t) {
    if ((size_t)t >= types.size()) types.resize(t + 1);
    return types[t];
  }
};

static inline uint64_t rng_next(uint64_t& s) {      // splitmix64
  s += 0x9E3779B97F4A7C15ull;
  return mix64(s);
}

static ThreadPool* graph_pool() {
  static ThreadPool pool((int)std::max(2u, std::min(16u, std::thread::hardware_concurrency())));
  return &pool;
}

}  // namespace pbh

using namespace pbh;

extern "C" {

int pbh_graph_create(void** out) { *out = new GraphStore(); return 0; }
int pbh_graph_destroy(void* h) { delete (GraphStore*)h; return 0; }

int pbh_graph_add_edges(void* h, int type, const uint64_t* src, const uint64_t* dst, int64_t n, int reverse) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(!t.built, "edge type %d is already built", type);
  const uint64_t* a = reverse ? dst : src;
  const uint64_t* b = reverse ? src : dst;
  t.src.insert(t.src.end(), a, a + n);
  t.dst.insert(t.dst.end(), b, b + n);
  return 0;
}

// text file, one edge per line: "<src>\t<dst>[\t<anything>]" (load_edge_file, graph_gpu_wrapper.h:49)
int pbh_graph_load_edge_file(void* h, int type, const char* path, int reverse, int64_t* n_loaded) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(!t.built, "edge type %d is already built", type);
  FILE* fp = fopen(path, "r");
  PBH_REQUIRE(fp != nullptr, "cannot open %s", path);
  char* line = nullptr;
  size_t cap = 0;
  int64_t cnt = 0;
  while (getline(&line, &cap, fp) > 0) {
    char* p = line;
    char* e = nullptr;
    uint64_t a = strtoull(p, &e, 10);
    if (e == p) continue;
    p = e;
    uint64_t b = strtoull(p, &e, 10);
    if (e == p) continue;
    t.src.push_back(reverse ? b : a);
    t.dst.push_back(reverse ? a : b);
    ++cnt;
  }
  free(line);
  fclose(fp);
  if (n_loaded) *n_loaded = cnt;
  return 0;
}

int pbh_graph_build(void* h, int type) {
  ((GraphStore*)h)->type(type).build();
  return 0;
}

int64_t pbh_graph_num_nodes(void* h, int type) { return (int64_t)((GraphStore*)h)->type(type).node_ids.size(); }
int64_t pbh_graph_num_edges(void* h, int type) { return (int64_t)((GraphStore*)h)->type(type).col.size(); }
int pbh_graph_num_types(void* h) { return (int)((GraphStore*)h)->types.size(); }

// copies the CSR arrays out (for the upload to HBM): node_ids [n], row_ptr [n + 1], col [m]
int pbh_graph_export_csr(void* h, int type, uint64_t* node_ids, int64_t* row_ptr, int64_t* col) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(t.built, "edge type %d is not built", type);
  if (node_ids) memcpy(node_ids, t.node_ids.data(), t.node_ids.size() * 8);
  if (row_ptr) memcpy(row_ptr, t.row_ptr.data(), t.row_ptr.size() * 8);
  if (col) memcpy(col, t.col.data(), t.col.size() * 8);
  return 0;
}

// Uniform neighbour sampling without replacement (graph_neighbor_sample): for node i up to k
// neighbour ids into out[i*k ..], their number into out_cnt[i]; unused entries are 0.
int pbh_graph_sample_neighbors(void* h, int type, const uint64_t* nodes, int64_t n, int k, uint64_t seed,
                               uint64_t* out, int32_t* out_cnt) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(t.built, "edge type %d is not built", type);
  graph_pool()->parallel_range((size_t)n, [&](size_t b, size_t e, int) {
    std::vector<int64_t> pick;
    for (size_t i = b; i < e; ++i) {
      uint64_t* o = out + i * (size_t)k;
      std::fill(o, o + k, 0ull);
      int64_t idx = t.index_of(nodes[i]);
      int64_t deg = idx < 0 ? 0 : t.row_ptr[idx + 1] - t.row_ptr[idx];
      const int64_t* nb = idx < 0 ? nullptr : t.col.data() + t.row_ptr[idx];
      uint64_t s = mix64(seed ^ (nodes[i] * 0x9E3779B97F4A7C15ull) ^ (uint64_t)i);
      int cnt = (int)std::min<int64_t>(deg, k);
      if (deg <= k) {
        for (int j = 0; j < cnt; ++j) o[j] = t.node_ids[nb[j]];
      } else {                                         // Floyd's algorithm: k distinct positions of deg
        pick.clear();
        for (int64_t j = deg - k; j < deg; ++j) {
          int64_t r = (int64_t)(rng_next(s) % (uint64_t)(j + 1));
          if (std::find(pick.begin(), pick.end(), r) != pick.end()) r = j;
          pick.push_back(r);
        }
        for (int j = 0; j < k; ++j) o[j] = t.node_ids[nb[pick[j]]];
      }
      out_cnt[i] = cnt;
    }
  });
  return 0;
}

// Random walks: walk i starts at starts[i]; step s follows edge type metapath[s % n_meta].  A walk
// that reaches a node without out-edges (or unknown to the step's edge type) ends there: the rest of
// its row is 0 and out_len[i] holds the number of valid nodes (>= 1).  out is [n, walk_len].
int pbh_graph_random_walk(void* h, const int* metapath, int n_meta, const uint64_t* starts, int64_t n,
                          int walk_len, uint64_t seed, uint64_t* out, int32_t* out_len) {
  GraphStore* g = (GraphStore*)h;
  for (int m = 0; m < n_meta; ++m)
    PBH_REQUIRE((size_t)metapath[m] < g->types.size() && g->types[metapath[m]].built,
                "metapath edge type %d is not built", metapath[m]);
  graph_pool()->parallel_range((size_t)n, [&](size_t b, size_t e, int) {
    for (size_t i = b; i < e; ++i) {
      uint64_t* o = out + i * (size_t)walk_len;
      std::fill(o, o + walk_len, 0ull);
      uint64_t cur = starts[i];
      uint64_t s = mix64(seed + 0x632be59bd9b4e019ull * (uint64_t)(i + 1));
      o[0] = cur;
      int len = 1;
      for (; len < walk_len; ++len) {
        const EdgeType& t = g->types[metapath[(len - 1) % n_meta]];
        int64_t idx = t.index_of(cur);
        if (idx < 0) break;
        int64_t deg = t.row_ptr[idx + 1] - t.row_ptr[idx];
        if (deg == 0) break;
        cur = t.node_ids[t.col[t.row_ptr[idx] + (int64_t)(rng_next(s) % (uint64_t)deg)]];
        o[len] = cur;
      }
      if (out_len) out_len[i] = len;
    }
  });
  return 0;
}

}  // extern "C"
