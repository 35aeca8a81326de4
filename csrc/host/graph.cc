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
#include <cmath>
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <unordered_map>

#include "util.h"

namespace pbh {

struct EdgeType {
  std::vector<uint64_t> src, dst;        // raw edge list until build()
  std::vector<float> raw_w;              // optional edge weights (same length as src, or empty)
  std::vector<float> wgt;                // weight per CSR edge (empty: unweighted)
  // partitions of the node index range for paging a graph larger than HBM (make_partitions)
  std::vector<int64_t> part_begin;       // [n_parts + 1] node-index boundaries
  int next_part = 0;
  std::vector<uint64_t> node_ids;        // sorted unique ids (sources and destinations)
  std::vector<int64_t> row_ptr;          // [n_nodes + 1]
  std::vector<int64_t> col;              // neighbour node INDEX per edge
  bool built = false;

  int64_t index_of(uint64_t id) const {
    auto it = std::lower_bound(node_ids.begin(), node_ids.end(), id);
    return (it != node_ids.end() && *it == id) ? (int64_t)(it - node_ids.begin()) : -1;
  }

  void build(ThreadPool* pool = nullptr) {
    node_ids = src;
    node_ids.insert(node_ids.end(), dst.begin(), dst.end());
    radix_sort_u64(node_ids);
    node_ids.erase(std::unique(node_ids.begin(), node_ids.end()), node_ids.end());
    const size_t n = node_ids.size(), m = src.size();
    row_ptr.assign(n + 1, 0);
    std::vector<int64_t> si(m), di(m);
    auto par = [&](size_t count, const std::function<void(size_t, size_t)>& fn) {
      if (pool && count > 65536) pool->parallel_range(count, [&](size_t b, size_t e, int) { fn(b, e); });
      else fn(0, count);
    };
    par(m, [&](size_t b, size_t e) {                       // id -> dense index, both ends of every edge
      for (size_t k = b; k < e; ++k) { si[k] = index_of(src[k]); di[k] = index_of(dst[k]); }
    });
    for (size_t e = 0; e < m; ++e) ++row_ptr[si[e] + 1];
    for (size_t i = 0; i < n; ++i) row_ptr[i + 1] += row_ptr[i];
    col.assign(m, 0);
    const bool weighted = raw_w.size() == m && m > 0;
    if (weighted) wgt.assign(m, 0.f);
    std::vector<int64_t> fill(row_ptr.begin(), row_ptr.end() - 1);
    for (size_t e = 0; e < m; ++e) {
      const int64_t at = fill[si[e]]++;
      col[at] = di[e];
      if (weighted) wgt[at] = raw_w[e];
    }
    par(n, [&](size_t nb, size_t ne) {                     // neighbour lists sorted by index, rows in parallel
      std::vector<std::pair<int64_t, float>> tmp;
      for (size_t i = nb; i < ne; ++i) {
        const int64_t b = row_ptr[i], e2 = row_ptr[i + 1];
        if (!weighted) { std::sort(col.begin() + b, col.begin() + e2); continue; }
        tmp.resize((size_t)(e2 - b));
        for (int64_t j = b; j < e2; ++j) tmp[(size_t)(j - b)] = {col[j], wgt[j]};
        std::sort(tmp.begin(), tmp.end());
        for (int64_t j = b; j < e2; ++j) { col[j] = tmp[(size_t)(j - b)].first; wgt[j] = tmp[(size_t)(j - b)].second; }
      }
    });
    src.clear(); src.shrink_to_fit();
    dst.clear(); dst.shrink_to_fit();
    raw_w.clear(); raw_w.shrink_to_fit();
    part_begin.clear();
    next_part = 0;
    built = true;
  }
};

// Nodes of one type with their feature slots (load_node_file / get_feature_of_nodes,
// heter_ps/graph_gpu_wrapper.h:50, :99): per node a list of (slot id, feasign) pairs.
struct NodeType {
  std::vector<uint64_t> ids;             // as added, then sorted by finalize()
  std::vector<int64_t> off;              // [n + 1] into slot / val
  std::vector<int32_t> slot;
  std::vector<uint64_t> val;
  std::vector<float> weight;             // optional node weight (load_node_weight), 1 by default
  std::vector<int64_t> order;            // sorted position -> insertion position
  bool sorted = false;

  void finalize() {
    if (sorted) return;
    const size_t n = ids.size();
    order.resize(n);
    for (size_t i = 0; i < n; ++i) order[i] = (int64_t)i;
    std::sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return ids[a] < ids[b]; });
    sorted = true;
  }
  int64_t find(uint64_t id) {            // insertion position of `id`, -1 if absent
    finalize();
    size_t lo = 0, hi = order.size();
    while (lo < hi) {
      size_t mid = (lo + hi) / 2;
      if (ids[order[mid]] < id) lo = mid + 1; else hi = mid;
    }
    return (lo < order.size() && ids[order[lo]] == id) ? order[lo] : -1;
  }
};

struct GraphStore {
  std::vector<EdgeType> types;
  std::vector<NodeType> node_types;
  NodeType& ntype(int t) {
    if ((size_t)t >= node_types.size()) node_types.resize(t + 1);
    return node_types[t];
  }
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
  ((GraphStore*)h)->type(type).build(graph_pool());
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

// ---- weighted edges ------------------------------------------------------------------------
int pbh_graph_add_edges_weighted(void* h, int type, const uint64_t* src, const uint64_t* dst, const float* w,
                                 int64_t n, int reverse) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(!t.built, "edge type %d is already built", type);
  PBH_REQUIRE(t.raw_w.size() == t.src.size(), "edge type %d mixes weighted and unweighted edges", type);
  const uint64_t* a = reverse ? dst : src;
  const uint64_t* b = reverse ? src : dst;
  t.src.insert(t.src.end(), a, a + n);
  t.dst.insert(t.dst.end(), b, b + n);
  t.raw_w.insert(t.raw_w.end(), w, w + n);
  return 0;
}

// "<src>\t<dst>\t<weight>" lines; a missing weight counts as 1
int pbh_graph_load_weighted_edge_file(void* h, int type, const char* path, int reverse, int64_t* n_loaded) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(!t.built, "edge type %d is already built", type);
  PBH_REQUIRE(t.raw_w.size() == t.src.size(), "edge type %d mixes weighted and unweighted edges", type);
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
    p = e;
    float w = strtof(p, &e);
    if (e == p) w = 1.f;
    t.src.push_back(reverse ? b : a);
    t.dst.push_back(reverse ? a : b);
    t.raw_w.push_back(w);
    ++cnt;
  }
  free(line);
  fclose(fp);
  if (n_loaded) *n_loaded = cnt;
  return 0;
}

int pbh_graph_is_weighted(void* h, int type) { return ((GraphStore*)h)->type(type).wgt.empty() ? 0 : 1; }

int pbh_graph_export_weights(void* h, int type, float* out) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(t.built && !t.wgt.empty(), "edge type %d has no weights", type);
  memcpy(out, t.wgt.data(), t.wgt.size() * sizeof(float));
  return 0;
}

// Weighted neighbour sampling without replacement: k neighbours with probability proportional to the edge
// weight (Efraimidis-Spirakis keys u^(1/w), the k largest win) — the reference's weighted sample of the
// GPU graph table (graph_gpu_ps_table_inl.cu, weighted_sample kernels).  Unweighted types fall back to the
// uniform sampler.
int pbh_graph_sample_neighbors_weighted(void* h, int type, const uint64_t* nodes, int64_t n, int k, uint64_t seed,
                                        uint64_t* out, int32_t* out_cnt) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(t.built, "edge type %d is not built", type);
  if (t.wgt.empty()) return pbh_graph_sample_neighbors(h, type, nodes, n, k, seed, out, out_cnt);
  graph_pool()->parallel_range((size_t)n, [&](size_t b, size_t e, int) {
    std::vector<std::pair<double, int64_t>> keyed;
    for (size_t i = b; i < e; ++i) {
      uint64_t* o = out + i * (size_t)k;
      std::fill(o, o + k, 0ull);
      int64_t idx = t.index_of(nodes[i]);
      int64_t deg = idx < 0 ? 0 : t.row_ptr[idx + 1] - t.row_ptr[idx];
      const int64_t base = idx < 0 ? 0 : t.row_ptr[idx];
      uint64_t s = mix64(seed ^ (nodes[i] * 0x9E3779B97F4A7C15ull) ^ (uint64_t)i);
      int cnt = (int)std::min<int64_t>(deg, k);
      if (deg <= k) {
        for (int j = 0; j < cnt; ++j) o[j] = t.node_ids[t.col[base + j]];
      } else {
        keyed.clear();
        for (int64_t j = 0; j < deg; ++j) {
          const double u = ((double)(rng_next(s) >> 11) + 1.0) * (1.0 / 9007199254740993.0);   // (0, 1)
          const double w = t.wgt[base + j] > 0.f ? (double)t.wgt[base + j] : 1e-30;
          keyed.emplace_back(std::log(u) / w, j);                    // log of u^(1/w): same order
        }
        std::partial_sort(keyed.begin(), keyed.begin() + k, keyed.end(),
                          [](const std::pair<double, int64_t>& x, const std::pair<double, int64_t>& y) { return x.first > y.first; });
        for (int j = 0; j < k; ++j) o[j] = t.node_ids[t.col[base + keyed[j].second]];
      }
      out_cnt[i] = cnt;
    }
  });
  return 0;
}

// ---- node types with feature slots ---------------------------------------------------------
// ids [n]; features of node i are (slots[j], values[j]) for j in [offsets[i], offsets[i+1])
int pbh_graph_add_nodes(void* h, int ntype, const uint64_t* ids, int64_t n, const int64_t* offsets,
                        const int32_t* slots, const uint64_t* values) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  if (t.off.empty()) t.off.push_back(0);
  for (int64_t i = 0; i < n; ++i) {
    t.ids.push_back(ids[i]);
    if (offsets) {
      for (int64_t j = offsets[i]; j < offsets[i + 1]; ++j) { t.slot.push_back(slots[j]); t.val.push_back(values[j]); }
    }
    t.off.push_back((int64_t)t.val.size());
  }
  t.sorted = false;
  return 0;
}

// "<node type name>\t<id>[\t<slot>:<v>[<sep><v>...]]..." — the node file of load_node_file; `type_name` NULL or
// "" accepts every line, else only lines whose first field equals it.  `sep` separates several values of a slot.
int pbh_graph_load_node_file(void* h, int ntype, const char* path, const char* type_name, char sep, int64_t* n_loaded) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  if (t.off.empty()) t.off.push_back(0);
  FILE* fp = fopen(path, "r");
  PBH_REQUIRE(fp != nullptr, "cannot open %s", path);
  char* line = nullptr;
  size_t cap = 0;
  int64_t cnt = 0;
  const size_t tn = type_name ? strlen(type_name) : 0;
  ssize_t len;
  while ((len = getline(&line, &cap, fp)) > 0) {
    while (len > 0 && (line[len - 1] == '\n' || line[len - 1] == '\r')) line[--len] = 0;
    char* p = line;
    char* tab = strchr(p, '\t');
    if (!tab) continue;
    if (tn && ((size_t)(tab - p) != tn || strncmp(p, type_name, tn) != 0)) continue;
    p = tab + 1;
    char* e = nullptr;
    uint64_t id = strtoull(p, &e, 10);
    if (e == p) continue;
    t.ids.push_back(id);
    p = e;
    while (*p == '\t') {
      ++p;
      long sl = strtol(p, &e, 10);
      if (e == p || *e != ':') break;
      p = e + 1;
      while (true) {
        uint64_t v = strtoull(p, &e, 10);
        if (e == p) break;
        t.slot.push_back((int32_t)sl);
        t.val.push_back(v);
        p = e;
        if (*p == sep && sep != '\t') ++p; else break;
      }
    }
    t.off.push_back((int64_t)t.val.size());
    ++cnt;
  }
  free(line);
  fclose(fp);
  t.sorted = false;
  if (n_loaded) *n_loaded = cnt;
  return 0;
}

int64_t pbh_graph_num_typed_nodes(void* h, int ntype) { return (int64_t)((GraphStore*)h)->ntype(ntype).ids.size(); }

int pbh_graph_export_typed_nodes(void* h, int ntype, uint64_t* out) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  t.finalize();
  for (size_t i = 0; i < t.order.size(); ++i) out[i] = t.ids[t.order[i]];
  return 0;
}

// get_feature_of_nodes: out [n, n_slots] = the first feasign of slot slot_ids[j] of node i (0 when the node is
// unknown or has no value in that slot) — the fixed slot_num layout the reference's feed consumes
int pbh_graph_node_features(void* h, int ntype, const uint64_t* nodes, int64_t n, const int32_t* slot_ids, int n_slots,
                            uint64_t* out) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  t.finalize();
  graph_pool()->parallel_range((size_t)n, [&](size_t b, size_t e, int) {
    for (size_t i = b; i < e; ++i) {
      uint64_t* o = out + i * (size_t)n_slots;
      std::fill(o, o + n_slots, 0ull);
      const int64_t at = t.find(nodes[i]);
      if (at < 0) continue;
      for (int64_t j = t.off[at]; j < t.off[at + 1]; ++j)
        for (int q = 0; q < n_slots; ++q)
          if (t.slot[j] == slot_ids[q] && o[q] == 0) o[q] = t.val[j];
    }
  });
  return 0;
}

// get_all_feature_ids: sorted unique feasigns of the node type; call with out == NULL to get the count
int pbh_graph_all_feature_ids(void* h, int ntype, uint64_t* out, int64_t* n_out) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  std::vector<uint64_t> v(t.val);
  std::sort(v.begin(), v.end());
  v.erase(std::unique(v.begin(), v.end()), v.end());
  if (n_out) *n_out = (int64_t)v.size();
  if (out) memcpy(out, v.data(), v.size() * 8);
  return 0;
}

// load_node_weight: "<id>\t<weight>" lines; nodes of the type that the file does not name keep weight 1
int pbh_graph_load_node_weight(void* h, int ntype, const char* path, int64_t* n_set) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  t.finalize();
  t.weight.assign(t.ids.size(), 1.f);
  FILE* fp = fopen(path, "r");
  PBH_REQUIRE(fp != nullptr, "cannot open %s", path);
  char* line = nullptr;
  size_t cap = 0;
  int64_t cnt = 0;
  while (getline(&line, &cap, fp) > 0) {
    char* e = nullptr;
    uint64_t id = strtoull(line, &e, 10);
    if (e == line) continue;
    char* p = e;
    float w = strtof(p, &e);
    if (e == p) continue;
    const int64_t at = t.find(id);
    if (at >= 0) { t.weight[at] = w; ++cnt; }
  }
  free(line);
  fclose(fp);
  if (n_set) *n_set = cnt;
  return 0;
}

int pbh_graph_node_weights(void* h, int ntype, const uint64_t* nodes, int64_t n, float* out) {
  NodeType& t = ((GraphStore*)h)->ntype(ntype);
  t.finalize();
  for (int64_t i = 0; i < n; ++i) {
    const int64_t at = t.find(nodes[i]);
    out[i] = (at >= 0 && !t.weight.empty()) ? t.weight[at] : (at >= 0 ? 1.f : 0.f);
  }
  return 0;
}

// ---- partitions: paging a graph that does not fit in HBM -----------------------------------
// make_partitions(idx, byte_size, device_len): consecutive node-index ranges whose CSR footprint (8 bytes per
// edge + 24 per node) stays within byte_size * device_len; returns the number of partitions.
int pbh_graph_make_partitions(void* h, int type, int64_t byte_size, int device_len, int* n_parts) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(t.built, "edge type %d is not built", type);
  PBH_REQUIRE(byte_size > 0 && device_len > 0, "make_partitions: byte_size and device_len must be positive");
  const int64_t budget = byte_size * device_len;
  t.part_begin.assign(1, 0);
  int64_t used = 0;
  const int64_t n = (int64_t)t.node_ids.size();
  for (int64_t i = 0; i < n; ++i) {
    const int64_t cost = 24 + 8 * (t.row_ptr[i + 1] - t.row_ptr[i]);
    if (used > 0 && used + cost > budget) { t.part_begin.push_back(i); used = 0; }
    used += cost;
  }
  t.part_begin.push_back(n);
  t.next_part = 0;
  if (n_parts) *n_parts = (int)t.part_begin.size() - 1;
  return 0;
}

int pbh_graph_partition_num(void* h, int type) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  return t.part_begin.empty() ? 0 : (int)t.part_begin.size() - 1;
}

// load_next_partition: index of the partition to page in next, -1 after the last one (then it wraps)
int pbh_graph_next_partition(void* h, int type) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  const int np = t.part_begin.empty() ? 0 : (int)t.part_begin.size() - 1;
  if (t.next_part >= np) { t.next_part = 0; return -1; }
  return t.next_part++;
}

// geometry of one partition: its node count and edge count
int pbh_graph_partition_shape(void* h, int type, int part, int64_t* n_nodes, int64_t* n_edges) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(part >= 0 && part + 1 < (int)t.part_begin.size(), "partition %d out of range", part);
  const int64_t b = t.part_begin[part], e = t.part_begin[part + 1];
  *n_nodes = e - b;
  *n_edges = t.row_ptr[e] - t.row_ptr[b];
  return 0;
}

// CSR of one partition: node_ids [n], row_ptr [n + 1] (rebased to 0), col_ids [m] = GLOBAL neighbour ids (a
// neighbour may live in another partition), weights [m] when the type is weighted and `weights` is given
int pbh_graph_export_partition(void* h, int type, int part, uint64_t* node_ids, int64_t* row_ptr, uint64_t* col_ids,
                               float* weights) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(part >= 0 && part + 1 < (int)t.part_begin.size(), "partition %d out of range", part);
  const int64_t b = t.part_begin[part], e = t.part_begin[part + 1];
  const int64_t e0 = t.row_ptr[b];
  for (int64_t i = b; i < e; ++i) node_ids[i - b] = t.node_ids[i];
  for (int64_t i = b; i <= e; ++i) row_ptr[i - b] = t.row_ptr[i] - e0;
  for (int64_t j = e0; j < t.row_ptr[e]; ++j) {
    col_ids[j - e0] = t.node_ids[t.col[j]];
    if (weights && !t.wgt.empty()) weights[j - e0] = t.wgt[j];
  }
  return 0;
}

// export_partition_files: one text file per partition, "<path>/partition_<i>", one node id per line
int pbh_graph_export_partition_files(void* h, int type, const char* dir) {
  EdgeType& t = ((GraphStore*)h)->type(type);
  PBH_REQUIRE(!t.part_begin.empty(), "edge type %d has no partitions (call make_partitions)", type);
  for (int p = 0; p + 1 < (int)t.part_begin.size(); ++p) {
    char name[4096];
    snprintf(name, sizeof(name), "%s/partition_%d", dir, p);
    FILE* fp = fopen(name, "w");
    PBH_REQUIRE(fp != nullptr, "cannot write %s", name);
    for (int64_t i = t.part_begin[p]; i < t.part_begin[p + 1]; ++i) fprintf(fp, "%llu\n", (unsigned long long)t.node_ids[i]);
    fclose(fp);
  }
  return 0;
}

}  // extern "C"
