// Host-memory tier of the sparse table + CPU execution of pull/push + checkpoint I/O.
//
// Roles in the reference: the closed libbox_ps.so memory tier (BeginPass stages the
// pass's working set from host memory into HBM, EndPass writes it back; SaveBase /
// SaveDelta / LoadSSD2Mem / ShrinkTable, fleet/box_wrapper.cc:171-210, :1286-1324) and its
// open twin, pscore's MemorySparseTable + CtrDymfAccessor
// (paddle/fluid/distributed/ps/table/memory_sparse_table.cc:338-393,
// ctr_dymf_accessor.cc:68-142, :279-315).  The CPU optimizer math restates
// heter_ps/optimizer.cuh.h:30-472 and must match csrc/cuda/table.cuh bit-for-bit in
// structure (same layout, same update order).
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <mutex>
#include <shared_mutex>
#include <sstream>
#include <unordered_map>

#include "../../include/pbx_types.h"
#include "util.h"

namespace pbh {

static inline float uniform01(uint64_t key, uint32_t idx, uint64_t seed) {
  uint64_t x = mix64(key + 0x632be59bd9b4e019ull * (uint64_t)(idx + 1) + seed);
  return ((float)((x >> 40) + 1)) * (1.0f / 16777216.0f);
}

// key -> row number, open addressing with linear probing (load <= 0.5); an entry stores row + 1 so
// that 0 marks an empty cell and key 0 needs no special case
class FlatIndex {
 public:
  int64_t find(uint64_t key) const {
    if (cells_.empty()) return -1;
    for (size_t i = mix64(key) & mask_;; i = (i + 1) & mask_) {
      const Cell& c = cells_[i];
      if (c.row1 == 0) return -1;
      if (c.key == key) return (int64_t)c.row1 - 1;
    }
  }
  void emplace(uint64_t key, uint32_t row) {
    if ((n_ + 1) * 2 > cells_.size()) grow();
    put(key, row + 1);
    ++n_;
  }
  void prefetch(uint64_t key) const {
    if (!cells_.empty()) __builtin_prefetch(&cells_[mix64(key) & mask_]);
  }
  size_t size() const { return n_; }

 private:
  struct Cell { uint64_t key; uint32_t row1; };
  void put(uint64_t key, uint32_t row1) {
    size_t i = mix64(key) & mask_;
    while (cells_[i].row1 != 0) i = (i + 1) & mask_;
    cells_[i].key = key;
    cells_[i].row1 = row1;
  }
  void grow() {
    std::vector<Cell> old;
    old.swap(cells_);
    cells_.assign(old.empty() ? 1024 : old.size() * 2, Cell{0, 0});
    mask_ = cells_.size() - 1;
    for (const Cell& c : old)
      if (c.row1) put(c.key, c.row1);
  }
  std::vector<Cell> cells_;
  size_t mask_ = 0, n_ = 0;
};

class HostTable {
 public:
  HostTable(int embedx_dim, int optimizer, const PbxSparseOptConfig& cfg, uint64_t seed, int shards)
      : L(pbx_make_layout(embedx_dim, optimizer)), cfg(cfg), seed(seed), shards_(shards), mu_(shards) {}

  PbxRowLayout L;
  PbxSparseOptConfig cfg;
  uint64_t seed;
  bool test_mode = false;
  // > 0: serving-format saves (modes 1, 2) write embedx as int16 steps of this size and serving loads
  // multiply back — the quantised feature layout of the reference's pull records
  // (EmbedxQuantOp: value = int16 * scale, fleet/box_wrapper.cu:36-43)
  float serving_quant_scale = 0.f;

  struct Shard {
    FlatIndex index;
    std::vector<float> rows;
    std::vector<uint64_t> keys;
    std::vector<uint8_t> dirty;
  };

  // ---- SSD tier: cold rows leave host memory for one append-only file per shard ------------------
  // (role of the closed library's SSD tier behind LoadSSD2Mem / CheckNeedLimitMem,
  //  fleet/box_wrapper.cc:1316-1324, box_wrapper.h:638; open twin: pscore's SSDSparseTable,
  //  paddle/fluid/distributed/ps/table/ssd_sparse_table.cc.)  A record is the key followed by the
  //  full value row; the offset of every live record is indexed in memory (16 B per cold key instead
  //  of a whole row).  A lookup that misses host memory promotes the row back.
  struct SsdShard {
    std::unordered_map<uint64_t, uint64_t> index;   // key -> byte offset of its record
    int fd = -1;
    uint64_t end = 0;                               // append position
    uint64_t dead = 0;                              // records superseded or promoted (garbage)
    std::string path;
  };

  size_t shard_of(uint64_t k) const { return mix64(k ^ 0x5bd1e995u) % shards_.size(); }
  size_t rec_bytes() const { return 8 + (size_t)L.stride * sizeof(float); }

  size_t mem_rows() const {
    size_t n = 0;
    for (auto& s : shards_) n += s.keys.size();
    return n;
  }
  size_t ssd_rows() const {
    size_t n = 0;
    for (auto& s : ssd_) n += s.index.size();
    return n;
  }
  size_t size() const { return mem_rows() + ssd_rows(); }

  ~HostTable() { ssd_close(false); }

  void ssd_close(bool unlink_files) {
    for (auto& s : ssd_) {
      if (s.fd >= 0) close(s.fd);
      if (unlink_files && !s.path.empty()) unlink(s.path.c_str());
    }
    ssd_.clear();
  }

  int ssd_open(const std::string& dir) {
    ssd_close(false);
    ssd_.resize(shards_.size());
    for (size_t i = 0; i < ssd_.size(); ++i) {
      char name[512];
      snprintf(name, sizeof(name), "%s/ssd-%05d.bin", dir.c_str(), (int)i);
      ssd_[i].path = name;
      ssd_[i].fd = open(name, O_RDWR | O_CREAT | O_TRUNC, 0644);
      PBH_REQUIRE(ssd_[i].fd >= 0, "cannot open SSD shard file %s", name);
    }
    return 0;
  }

  bool ssd_read(const SsdShard& d, uint64_t off, uint64_t* key, float* row) const {
    std::vector<char> buf(rec_bytes());
    ssize_t got = pread(d.fd, buf.data(), buf.size(), (off_t)off);
    if (got != (ssize_t)buf.size()) return false;
    memcpy(key, buf.data(), 8);
    memcpy(row, buf.data() + 8, (size_t)L.stride * sizeof(float));
    return true;
  }

  // rewrites a shard file without its garbage once more than half of it is dead
  void ssd_compact(SsdShard& d) {
    if (d.dead < 1024 || d.dead < d.index.size()) return;
    std::string tmp = d.path + ".tmp";
    int nfd = open(tmp.c_str(), O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (nfd < 0) return;
    std::vector<char> buf(rec_bytes());
    std::vector<std::pair<uint64_t, uint64_t>> moved;     // key -> offset in the new file
    moved.reserve(d.index.size());
    uint64_t pos = 0;
    bool ok = true;
    for (auto& kv : d.index) {
      ok = pread(d.fd, buf.data(), buf.size(), (off_t)kv.second) == (ssize_t)buf.size() &&
           pwrite(nfd, buf.data(), buf.size(), (off_t)pos) == (ssize_t)buf.size();
      if (!ok) break;
      moved.emplace_back(kv.first, pos);
      pos += buf.size();
    }
    if (!ok) {                         // disk full etc.: the old file and its offsets stay as they are
      close(nfd);
      unlink(tmp.c_str());
      return;
    }
    for (auto& kv : moved) d.index[kv.first] = kv.second;
    close(d.fd);
    rename(tmp.c_str(), d.path.c_str());
    d.fd = nfd;
    d.end = pos;
    d.dead = 0;
  }

  // Moves rows out of host memory.  Rows unseen for >= min_unseen_days go first (negative: age is
  // ignored); if the table still holds more than max_mem_rows (> 0), the longest-unseen, then
  // lowest-score rows follow until it fits.  Rows touched since the last delta save (dirty) and key 0
  // stay.  Returns the number spilled.
  int64_t spill(float min_unseen_days, int64_t max_mem_rows, ThreadPool* pool) {
    if (ssd_.empty()) return 0;
    const size_t S = shards_.size();
    std::vector<int64_t> moved(S, 0);
    const int64_t quota = max_mem_rows > 0 ? std::max<int64_t>(1, max_mem_rows / (int64_t)S) : -1;
    auto work = [&](size_t si) {
      Shard& s = shards_[si];
      SsdShard& d = ssd_[si];
      const size_t n = s.keys.size();
      std::vector<uint8_t> go(n, 0);
      size_t stay = 0;
      for (size_t i = 0; i < n; ++i) {
        const float* row = s.rows.data() + i * L.stride;
        go[i] = s.keys[i] != 0 && !s.dirty[i] && min_unseen_days >= 0.f && row[L.off_tail + 2] >= min_unseen_days;
        stay += !go[i];
      }
      if (quota >= 0 && (int64_t)stay > quota) {
        std::vector<uint32_t> cand;
        for (size_t i = 0; i < n; ++i)
          if (!go[i] && s.keys[i] != 0 && !s.dirty[i]) cand.push_back((uint32_t)i);
        size_t extra = std::min(cand.size(), stay - (size_t)quota);
        auto colder = [&](uint32_t a, uint32_t b) {
          const float* ra = s.rows.data() + (size_t)a * L.stride;
          const float* rb = s.rows.data() + (size_t)b * L.stride;
          if (ra[L.off_tail + 2] != rb[L.off_tail + 2]) return ra[L.off_tail + 2] > rb[L.off_tail + 2];
          return ra[L.off_tail + 0] < rb[L.off_tail + 0];
        };
        if (extra < cand.size()) std::nth_element(cand.begin(), cand.begin() + extra, cand.end(), colder);
        for (size_t j = 0; j < extra; ++j) go[cand[j]] = 1;
      }
      std::vector<char> out;
      Shard ns;
      for (size_t i = 0; i < n; ++i) {
        const float* row = s.rows.data() + i * L.stride;
        if (go[i]) {
          if (d.index.count(s.keys[i])) ++d.dead;
          d.index[s.keys[i]] = d.end + out.size();
          size_t at = out.size();
          out.resize(at + rec_bytes());
          memcpy(out.data() + at, &s.keys[i], 8);
          memcpy(out.data() + at + 8, row, (size_t)L.stride * sizeof(float));
          ++moved[si];
        } else {
          ns.index.emplace(s.keys[i], (uint32_t)ns.keys.size());
          ns.keys.push_back(s.keys[i]);
          ns.dirty.push_back(s.dirty[i]);
          ns.rows.insert(ns.rows.end(), row, row + L.stride);
        }
      }
      if (out.empty()) return;
      if (pwrite(d.fd, out.data(), out.size(), (off_t)d.end) != (ssize_t)out.size()) {
        // the write failed: forget the new offsets, keep every row in memory
        uint64_t key;
        for (size_t at = 0; at < out.size(); at += rec_bytes()) { memcpy(&key, out.data() + at, 8); d.index.erase(key); }
        moved[si] = 0;
        return;
      }
      d.end += out.size();
      s = std::move(ns);
      ssd_compact(d);
    };
    if (pool) pool->parallel_range(S, [&](size_t b, size_t e, int) { for (size_t i = b; i < e; ++i) work(i); });
    else for (size_t i = 0; i < S; ++i) work(i);
    int64_t total = 0;
    for (auto m : moved) total += m;
    return total;
  }

  // row of `key` comes back from its shard file into host memory (nullptr when it is not there)
  float* promote(Shard& s, SsdShard& d, uint64_t key) {
    auto it = d.index.find(key);
    if (it == d.index.end()) return nullptr;
    std::vector<float> row(L.stride);
    uint64_t k2 = 0;
    if (!ssd_read(d, it->second, &k2, row.data()) || k2 != key) return nullptr;
    d.index.erase(it);
    ++d.dead;
    uint32_t idx = (uint32_t)s.keys.size();
    s.index.emplace(key, idx);
    s.keys.push_back(key);
    s.dirty.push_back(0);
    s.rows.insert(s.rows.end(), row.begin(), row.end());
    return s.rows.data() + (size_t)idx * L.stride;
  }

  // LoadSSD2Mem: every cold row returns to host memory (before a day-end shrink / base save)
  int64_t load_ssd2mem(ThreadPool* pool) {
    if (ssd_.empty()) return 0;
    const size_t S = shards_.size();
    std::vector<int64_t> cnt(S, 0);
    auto work = [&](size_t si) {
      // one sequential read of the shard file in large blocks; a record is live when the index still
      // points at its offset (superseded and promoted records are skipped)
      SsdShard& d = ssd_[si];
      Shard& s = shards_[si];
      const size_t rb = rec_bytes();
      const size_t per_block = std::max<size_t>(1, (size_t)(8 << 20) / rb);
      std::vector<char> buf(per_block * rb);
      s.rows.reserve(s.rows.size() + d.index.size() * L.stride);
      for (uint64_t off = 0; off < d.end && !d.index.empty();) {
        size_t want = (size_t)std::min<uint64_t>(buf.size(), d.end - off);
        ssize_t got = pread(d.fd, buf.data(), want, (off_t)off);
        if (got <= 0) break;
        size_t recs = (size_t)got / rb;
        if (recs == 0) break;
        for (size_t r = 0; r < recs; ++r) {
          uint64_t key;
          memcpy(&key, buf.data() + r * rb, 8);
          auto it = d.index.find(key);
          if (it == d.index.end() || it->second != off + r * rb) continue;
          d.index.erase(it);
          uint32_t idx = (uint32_t)s.keys.size();
          s.index.emplace(key, idx);
          s.keys.push_back(key);
          s.dirty.push_back(0);
          const float* row = (const float*)(buf.data() + r * rb + 8);
          s.rows.insert(s.rows.end(), row, row + L.stride);
          ++cnt[si];
        }
        off += recs * rb;
      }
      std::vector<uint64_t> ks;                       // anything a short read left behind
      for (auto& kv : d.index) ks.push_back(kv.first);
      for (uint64_t k : ks) cnt[si] += promote(s, d, k) != nullptr;
      if (d.index.empty() && d.fd >= 0 && ftruncate(d.fd, 0) == 0) { d.end = 0; d.dead = 0; }
    };
    if (pool) pool->parallel_range(S, [&](size_t b, size_t e, int) { for (size_t i = b; i < e; ++i) work(i); });
    else for (size_t i = 0; i < S; ++i) work(i);
    int64_t total = 0;
    for (auto c : cnt) total += c;
    return total;
  }

  void init_row(float* row, uint64_t key) const {
    std::fill(row, row + L.stride, 0.f);
    row[2] = cfg.initial_range != 0.f ? uniform01(key, 0xffffu, seed) * cfg.initial_range : 0.f;
    if (L.optimizer != PBX_OPT_ADAGRAD) {
      row[L.off_es + 2] = cfg.beta1_decay_rate;
      row[L.off_es + 3] = cfg.beta2_decay_rate;
    }
  }

  float* find_or_create(uint64_t key, bool create, bool* fresh = nullptr) {
    Shard& s = shards_[shard_of(key)];
    int64_t at = s.index.find(key);
    if (fresh) *fresh = false;
    if (at >= 0) return s.rows.data() + (size_t)at * L.stride;
    if (!ssd_.empty()) {
      if (float* cold = promote(s, ssd_[shard_of(key)], key)) return cold;
    }
    if (!create) return nullptr;
    uint32_t idx = (uint32_t)s.keys.size();
    s.index.emplace(key, idx);
    s.keys.push_back(key);
    s.dirty.push_back(0);
    s.rows.resize(s.rows.size() + L.stride);
    float* row = s.rows.data() + (size_t)idx * L.stride;
    init_row(row, key);
    if (fresh) *fresh = true;
    return row;
  }

  // full rows for HBM staging; missing keys are created with initial state
  void fetch_rows(const uint64_t* keys, int64_t n, float* out, bool create, ThreadPool* pool) {
    // group by shard to lock each shard once per worker
    run_sharded(keys, n, pool, [&](Shard&, int64_t i) {
      float* row = find_or_create(keys[i], create);
      if (row) memcpy(out + i * L.stride, row, L.stride * sizeof(float));
      else init_row(out + i * L.stride, keys[i]);
    }, create);
  }

  void store_rows(const uint64_t* keys, int64_t n, const float* rows, bool mark_dirty, ThreadPool* pool) {
    run_sharded(keys, n, pool, [&](Shard& s, int64_t i) {
      float* row = find_or_create(keys[i], true);
      memcpy(row, rows + i * L.stride, L.stride * sizeof(float));
      if (mark_dirty) s.dirty[s.index.find(keys[i])] = 1;
    }, true);
  }

  // CPU pull records [n, hidden] (key 0 / unknown keys -> zeros)
  void pull(const uint64_t* keys, int64_t n, float* out, int out_stride, ThreadPool* pool) {
    const int H = 3 + L.embedx_dim;
    run_sharded(keys, n, pool, [&](Shard&, int64_t i) {
      float* o = out + i * out_stride;
      float* row = keys[i] == 0 ? nullptr : find_or_create(keys[i], false);
      if (!row) { std::fill(o, o + out_stride, 0.f); return; }
      memcpy(o, row, H * sizeof(float));
      for (int j = H; j < out_stride; ++j) o[j] = 0.f;
    });
  }

  // CPU push of merged records for unique keys
  void push(const uint64_t* keys, int64_t n, const float* push, int push_stride, ThreadPool* pool) {
    if (test_mode) return;
    run_sharded(keys, n, pool, [&](Shard& s, int64_t i) {
      if (keys[i] == 0) return;
      float* row = find_or_create(keys[i], false);
      if (!row) return;
      apply_push(row, keys[i], push + i * push_stride);
      s.dirty[s.index.find(keys[i])] = 1;
    });
  }

  void adagrad(float* w, float* g2sum, const float* g, int n, float scale, float lr, float g2_0, float lo,
               float hi) const {
    float ratio = lr * std::sqrt(g2_0 / (g2_0 + *g2sum));
    float add = 0.f, inv = 1.f / scale;
    for (int i = 0; i < n; ++i) {
      float sg = g[i] * inv;
      w[i] = std::min(std::max(w[i] + sg * ratio, lo), hi);
      add += sg * sg;
    }
    *g2sum += add / (float)n;
  }

  void adam(float* w, float* m, float* v, float* b1p, float* b2p, const float* g, int n, float scale, float lr,
            float b1, float b2, float eps, float lo, float hi, bool shared) const {
    float b1pow = *b1p, b2pow = *b2p;
    float ratio = lr * std::sqrt(1.f - b2pow) / (1.f - b1pow);
    float inv = 1.f / scale;
    float sm = 0.f, sv = 0.f, m0 = *m, v0 = *v;
    for (int i = 0; i < n; ++i) {
      float sg = g[i] * inv;
      float mi = shared ? m0 : m[i], vi = shared ? v0 : v[i];
      float nm = b1 * mi + (1.f - b1) * sg;
      float nv = b2 * vi + (1.f - b2) * sg * sg;
      w[i] = std::min(std::max(w[i] + ratio * (nm / (std::sqrt(nv) + eps)), lo), hi);
      if (!shared) { m[i] = nm; v[i] = nv; }
      sm += nm; sv += nv;
    }
    if (shared) { *m = sm / (float)n; *v = sv / (float)n; }
    *b1p = b1pow * b1;
    *b2p = b2pow * b2;
  }

  void apply_push(float* row, uint64_t key, const float* push) const {
    const PbxSparseOptConfig& c = cfg;
    float g_show = push[1], g_click = push[2];
    if (g_show <= 0.f) return;
    row[L.off_tail + 1] = push[0];
    float show = row[0] + g_show, click = row[1] + g_click;
    row[0] = show; row[1] = click;
    row[L.off_tail + 0] += c.nonclk_coeff * (g_show - g_click) + c.clk_coeff * g_click;
    row[L.off_tail + 2] = 0.f;
    float* es = row + L.off_es;
    float* xs = row + L.off_xs;
    const int D = L.embedx_dim;
    if (L.optimizer == PBX_OPT_ADAGRAD)
      adagrad(row + 2, es, push + 3, 1, g_show, c.learning_rate, c.initial_g2sum, c.min_bound, c.max_bound);
    else
      adam(row + 2, es, es + 1, es + 2, es + 3, push + 3, 1, g_show, c.learning_rate, c.beta1_decay_rate,
           c.beta2_decay_rate, c.ada_epsilon, c.min_bound, c.max_bound, false);
    if (row[L.off_mf] == 0.f) {
      if (c.mf_create_thresholds <= c.nonclk_coeff * (show - click) + c.clk_coeff * click) {
        row[L.off_mf] = (float)D;
        for (int i = 0; i < D; ++i) row[3 + i] = uniform01(key, (uint32_t)i, seed) * c.mf_initial_range;
        if (L.optimizer == PBX_OPT_ADAM) { xs[2 * D] = c.mf_beta1_decay_rate; xs[2 * D + 1] = c.mf_beta2_decay_rate; }
        else if (L.optimizer == PBX_OPT_SHARED_ADAM) { xs[2] = c.mf_beta1_decay_rate; xs[3] = c.mf_beta2_decay_rate; }
      }
    } else if (L.optimizer == PBX_OPT_ADAGRAD) {
      adagrad(row + 3, xs, push + 4, D, g_show, c.mf_learning_rate, c.mf_initial_g2sum, c.mf_min_bound,
              c.mf_max_bound);
    } else if (L.optimizer == PBX_OPT_ADAM) {
      adam(row + 3, xs, xs + D, xs + 2 * D, xs + 2 * D + 1, push + 4, D, g_show, c.mf_learning_rate,
           c.mf_beta1_decay_rate, c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound, false);
    } else {
      adam(row + 3, xs, xs + 1, xs + 2, xs + 3, push + 4, D, g_show, c.mf_learning_rate, c.mf_beta1_decay_rate,
           c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound, true);
    }
  }

  // decay + evict (CtrDymfAccessor::Shrink); returns survivors
  int64_t shrink(float decay, float delete_threshold, float delete_after_unseen_days, ThreadPool* pool = nullptr) {
    load_ssd2mem(pool);           // decay and ageing apply to every row, cold ones included
    // shards in parallel; inside a shard the survivors are compacted in place (no second copy of the shard)
    // and the index is rebuilt only when something was evicted
    std::atomic<int64_t> kept{0};
    auto shrink_shard = [&](size_t si) {
      Shard& s = shards_[si];
      const size_t n = s.keys.size();
      size_t w = 0;
      for (size_t i = 0; i < n; ++i) {
        float* row = s.rows.data() + i * L.stride;
        row[0] *= decay;
        row[1] *= decay;
        row[L.off_tail + 2] += 1.f;
        float score = cfg.nonclk_coeff * (row[0] - row[1]) + cfg.clk_coeff * row[1];
        bool evict = s.keys[i] != 0 && (score < delete_threshold || row[L.off_tail + 2] > delete_after_unseen_days);
        if (evict) continue;
        if (w != i) {
          memmove(s.rows.data() + w * L.stride, row, L.stride * sizeof(float));
          s.keys[w] = s.keys[i];
          s.dirty[w] = s.dirty[i];
        }
        ++w;
      }
      if (w != n) {
        s.keys.resize(w);
        s.dirty.resize(w);
        s.rows.resize(w * L.stride);
        if (w * 2 < n) { s.rows.shrink_to_fit(); s.keys.shrink_to_fit(); s.dirty.shrink_to_fit(); }
        s.index = FlatIndex();
        for (size_t i = 0; i < w; ++i) s.index.emplace(s.keys[i], (uint32_t)i);
      }
      kept += (int64_t)w;
    };
    const size_t S = shards_.size();
    if (pool) pool->parallel_range(S, [&](size_t b, size_t e, int) { for (size_t i = b; i < e; ++i) shrink_shard(i); });
    else for (size_t i = 0; i < S; ++i) shrink_shard(i);
    return kept.load();
  }

  // ---- text checkpoint, one directory per table, files part-<shard>-<file>
  // line: "<key> <unseen_days> <delta_score> <show> <click> <embed_w> <embed_state...> <slot> <mf_dim>
  //        [<embedx_state...> <embedx_w...>]"   (ctr_dymf_accessor.cc:279-315)
  // mode 0 = all (batch model, full state), 1 = delta (dirty rows, serving format),
  //      2 = base (serving format: no optimizer state), 3 = batch model
  int save(const std::string& dir, int mode, int part_shard_base, float base_threshold, float delta_threshold,
           int64_t* n_saved, int clear_dirty, ThreadPool* pool = nullptr) {
    // every shard writes its own file: the shards are saved in parallel over the pool
    std::atomic<int64_t> total{0};
    std::atomic<int> failed{-1};
    auto save_shard = [&](size_t si) {
      Shard& s = shards_[si];
      char name[512];
      snprintf(name, sizeof(name), "%s/part-%03d-%05d", dir.c_str(), part_shard_base, (int)si);
      FILE* fp = fopen(name, "w");
      if (fp == nullptr) { failed.store((int)si); return; }
      std::vector<char> iobuf(1 << 20);
      setvbuf(fp, iobuf.data(), _IOFBF, iobuf.size());
      int64_t n_rows = 0;
      // rows in host memory first, then the shard's cold rows straight from its SSD file (never dirty)
      std::vector<std::pair<uint64_t, uint64_t>> cold;
      if (!ssd_.empty() && mode != 1) cold.assign(ssd_[si].index.begin(), ssd_[si].index.end());
      std::vector<float> cold_row(L.stride);
      for (size_t i = 0; i < s.keys.size() + cold.size(); ++i) {
        const bool in_mem = i < s.keys.size();
        uint64_t key = in_mem ? s.keys[i] : cold[i - s.keys.size()].first;
        if (!in_mem && !ssd_read(ssd_[si], cold[i - s.keys.size()].second, &key, cold_row.data())) continue;
        const float* row = in_mem ? s.rows.data() + i * L.stride : cold_row.data();
        float score = cfg.nonclk_coeff * (row[0] - row[1]) + cfg.clk_coeff * row[1];
        if (mode == 1 && !(s.dirty[i] && score >= delta_threshold)) continue;
        if (mode == 2 && score < base_threshold) continue;
        const bool serving = (mode == 1 || mode == 2);
        const int D = L.embedx_dim;
        const bool has_mf = row[L.off_mf] != 0.f;
        fprintf(fp, "%llu %g %g %g %g %.9g", (unsigned long long)key, row[L.off_tail + 2], row[L.off_tail + 0],
                row[0], row[1], row[2]);
        if (!serving) for (int j = 0; j < L.embed_state; ++j) fprintf(fp, " %.9g", row[L.off_es + j]);
        fprintf(fp, " %d %d", (int)row[L.off_tail + 1], has_mf ? D : 0);
        if (has_mf) {
          if (!serving) for (int j = 0; j < L.embedx_state; ++j) fprintf(fp, " %.9g", row[L.off_xs + j]);
          if (serving && serving_quant_scale > 0.f) {
            for (int j = 0; j < D; ++j) {
              float q = std::nearbyint(row[3 + j] / serving_quant_scale);
              fprintf(fp, " %d", (int)std::min(32767.f, std::max(-32768.f, q)));
            }
          } else {
            for (int j = 0; j < D; ++j) fprintf(fp, " %.9g", row[3 + j]);
          }
        }
        fputc('\n', fp);
        ++n_rows;
        // clear_dirty: 1 = a delta save clears the flags of the rows it wrote; 2 = so does a full save (mode 0
        // writes every row: the delta baseline is reset as a whole, what a base save wants)
        if (in_mem && ((clear_dirty && mode == 1) || (clear_dirty == 2 && mode == 0))) s.dirty[i] = 0;
      }
      if (fclose(fp) != 0) failed.store((int)si);
      total += n_rows;
    };
    const size_t S = shards_.size();
    if (pool) pool->parallel_range(S, [&](size_t b, size_t e, int) { for (size_t i = b; i < e; ++i) save_shard(i); });
    else for (size_t i = 0; i < S; ++i) save_shard(i);
    PBH_REQUIRE(failed.load() < 0, "cannot write %s/part-%03d-%05d", dir.c_str(), part_shard_base, failed.load());
    if (n_saved) *n_saved = total.load();
    return 0;
  }

  // loads files written by save() with mode 0/3 (full state); serving-format files restore
  // weights with fresh optimizer state.
  // rank / world (world > 1): only the keys this rank owns are kept — a model saved by N ranks can be
  // loaded by M (owner = pbx_owner_of, the function that shards the table)
  int load_file(const std::string& path, bool serving, int64_t* n_loaded, int rank = 0, int world = 1) {
    FILE* fp = fopen(path.c_str(), "r");
    PBH_REQUIRE(fp != nullptr, "cannot open %s", path.c_str());
    char* line = nullptr;
    size_t cap = 0;
    ssize_t len;
    int64_t cnt = 0;
    const int D = L.embedx_dim;
    while ((len = getline(&line, &cap, fp)) > 0) {
      char* p = line;
      uint64_t key = strtoull(p, &p, 10);
      std::vector<float> f;
      for (;;) {
        char* e = nullptr;
        float v = strtof(p, &e);
        if (e == p) break;
        f.push_back(v);
        p = e;
      }
      size_t need = 5 + (serving ? 0 : L.embed_state) + 2;
      if (f.size() < need) continue;
      if (world > 1 && (int)((mix64(key) >> 40) % (uint64_t)world) != rank) continue;
      // (several files may be loaded at once: the row is created and filled under its shard's lock)
      std::lock_guard<std::mutex> shard_guard(mu_[shard_of(key)]);
      float* row = find_or_create(key, true);
      init_row(row, key);
      size_t q = 0;
      row[L.off_tail + 2] = f[q++];
      row[L.off_tail + 0] = f[q++];
      row[0] = f[q++]; row[1] = f[q++]; row[2] = f[q++];
      if (!serving) for (int j = 0; j < L.embed_state; ++j) row[L.off_es + j] = f[q++];
      row[L.off_tail + 1] = f[q++];
      int mf = (int)f[q++];
      if (mf > 0 && f.size() >= q + (serving ? 0 : L.embedx_state) + D) {
        row[L.off_mf] = (float)D;
        if (!serving) for (int j = 0; j < L.embedx_state; ++j) row[L.off_xs + j] = f[q++];
        else if (L.optimizer == PBX_OPT_ADAM) { row[L.off_xs + 2 * D] = cfg.mf_beta1_decay_rate; row[L.off_xs + 2 * D + 1] = cfg.mf_beta2_decay_rate; }
        else if (L.optimizer == PBX_OPT_SHARED_ADAM) { row[L.off_xs + 2] = cfg.mf_beta1_decay_rate; row[L.off_xs + 3] = cfg.mf_beta2_decay_rate; }
        const float dq = (serving && serving_quant_scale > 0.f) ? serving_quant_scale : 1.f;
        for (int j = 0; j < D; ++j) row[3 + j] = f[q++] * dq;
      }
      ++cnt;
    }
    free(line);
    fclose(fp);
    if (n_loaded) *n_loaded = cnt;
    return 0;
  }

  void all_keys(std::vector<uint64_t>* out) const {
    out->clear();
    for (auto& s : shards_) out->insert(out->end(), s.keys.begin(), s.keys.end());
    for (auto& d : ssd_)
      for (auto& kv : d.index) out->push_back(kv.first);
  }

  void clear() {
    for (auto& s : shards_) s = Shard();
    for (auto& d : ssd_) {
      d.index.clear();
      if (d.fd >= 0 && ftruncate(d.fd, 0) == 0) { d.end = 0; d.dead = 0; }
    }
  }

 private:
  template <typename F>
  void run_sharded(const uint64_t* keys, int64_t n, ThreadPool* pool, F fn, bool may_create = false) {
    const size_t S = shards_.size();
    if (!pool || n < 4096) {
      for (int64_t i = 0; i < n; ++i) {
        const size_t q = shard_of(keys[i]);
        std::lock_guard<std::mutex> g(mu_[q]);
        fn(shards_[q], i);
      }
      return;
    }
    // one parallel pass computes every key's shard, a counting sort groups the positions by shard, then
    // each worker walks whole shards: no locks on the maps and every key is hashed once
    std::vector<uint8_t> sh((size_t)n);
    const int W = pool->size();
    std::vector<std::vector<int64_t>> cnt(W, std::vector<int64_t>(S, 0));
    pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
      auto& c = cnt[w];
      for (size_t i = b; i < e; ++i) { sh[i] = (uint8_t)shard_of(keys[i]); ++c[sh[i]]; }
    });
    std::vector<int64_t> start(S + 1, 0);
    for (size_t q = 0; q < S; ++q) {
      int64_t t = 0;
      for (int w = 0; w < W; ++w) t += cnt[w][q];
      start[q + 1] = start[q] + t;
    }
    // position of (worker w, shard q)'s first element: workers scanned contiguous ranges in order
    std::vector<std::vector<int64_t>> at(W, std::vector<int64_t>(S, 0));
    for (size_t q = 0; q < S; ++q) {
      int64_t t = start[q];
      for (int w = 0; w < W; ++w) { at[w][q] = t; t += cnt[w][q]; }
    }
    std::vector<int64_t> order((size_t)n);
    pool->parallel_range((size_t)n, [&](size_t b, size_t e, int w) {
      auto& a = at[w];
      for (size_t i = b; i < e; ++i) order[a[sh[i]]++] = (int64_t)i;
    });
    pool->parallel_range(S, [&](size_t qb, size_t qe, int) {
      for (size_t q = qb; q < qe; ++q) {
        Shard& shard = shards_[q];
        std::lock_guard<std::mutex> g(mu_[q]);
        if (may_create) {            // one growth step instead of repeated doubling while rows are added
          size_t extra = (size_t)(start[q + 1] - start[q]);
          if (shard.rows.capacity() < shard.rows.size() + extra * L.stride)
            shard.rows.reserve(std::max(shard.rows.capacity() * 2, shard.rows.size() + extra * L.stride));
        }
        for (int64_t j = start[q]; j < start[q + 1]; ++j) {
          if (j + 8 < start[q + 1]) shard.index.prefetch(keys[order[j + 8]]);
          fn(shard, order[j]);
        }
      }
    });
  }
  std::vector<Shard> shards_;
  std::vector<SsdShard> ssd_;
  std::vector<std::mutex> mu_;       // one per shard: held while a bulk call walks that shard
 public:
  // Bulk calls (fetch / store / pull / push) hold this shared and lock the shards they walk, so several of
  // them may run at once (the pass builder's fetch threads, a write-back next to the next pass's BuildPull);
  // whole-table operations (save, load, shrink, spill, LoadSSD2Mem, clear, key export) hold it exclusively.
  // Taken in the C entry points below.
  mutable std::shared_mutex table_mu;
};

}  // namespace pbh

using pbh::HostTable;
static pbh::ThreadPool* table_pool() {
  static pbh::ThreadPool pool((int)std::max(2u, std::min(16u, std::thread::hardware_concurrency())));
  return &pool;
}

extern "C" {

int pbh_table_create(void** out, int embedx_dim, int optimizer, const PbxSparseOptConfig* cfg, uint64_t seed) {
  *out = new HostTable(embedx_dim, optimizer, *cfg, seed, 64);
  return 0;
}
int pbh_table_destroy(void* h) { delete (HostTable*)h; return 0; }
int64_t pbh_table_size(void* h) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); return (int64_t)((HostTable*)h)->size(); }
int pbh_table_set_config(void* h, const PbxSparseOptConfig* cfg) { ((HostTable*)h)->cfg = *cfg; return 0; }
int pbh_table_set_test_mode(void* h, int on) { ((HostTable*)h)->test_mode = on != 0; return 0; }
int pbh_table_set_serving_quant(void* h, float scale) { ((HostTable*)h)->serving_quant_scale = scale; return 0; }
int pbh_table_clear(void* h) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); ((HostTable*)h)->clear(); return 0; }
int pbh_table_fetch_rows(void* h, const uint64_t* keys, int64_t n, float* out, int create) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  ((HostTable*)h)->fetch_rows(keys, n, out, create != 0, table_pool());
  return 0;
}
int pbh_table_store_rows(void* h, const uint64_t* keys, int64_t n, const float* rows, int mark_dirty) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  ((HostTable*)h)->store_rows(keys, n, rows, mark_dirty != 0, table_pool());
  return 0;
}
int pbh_table_pull(void* h, const uint64_t* keys, int64_t n, float* out, int out_stride) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  ((HostTable*)h)->pull(keys, n, out, out_stride, table_pool());
  return 0;
}
int pbh_table_push(void* h, const uint64_t* keys, int64_t n, const float* push, int push_stride) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  ((HostTable*)h)->push(keys, n, push, push_stride, table_pool());
  return 0;
}
int64_t pbh_table_shrink(void* h, float decay, float delete_threshold, float delete_after_unseen_days) {
  std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  return ((HostTable*)h)->shrink(decay, delete_threshold, delete_after_unseen_days, table_pool());
}
int pbh_table_save(void* h, const char* dir, int mode, int part_shard_base, float base_threshold,
                   float delta_threshold, int clear_dirty, int64_t* n_saved) {
  std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  return ((HostTable*)h)->save(dir, mode, part_shard_base, base_threshold, delta_threshold, n_saved,
                               clear_dirty, table_pool());
}
int pbh_table_load_file(void* h, const char* path, int serving, int64_t* n_loaded) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);      // (files may load concurrently: rows are filled under shard locks)
  return ((HostTable*)h)->load_file(path, serving != 0, n_loaded);
}
int pbh_table_load_file_owned(void* h, const char* path, int serving, int rank, int world, int64_t* n_loaded) {
  std::shared_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  return ((HostTable*)h)->load_file(path, serving != 0, n_loaded, rank, world);
}
int pbh_table_ssd_open(void* h, const char* dir) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); return ((HostTable*)h)->ssd_open(dir); }
int pbh_table_ssd_close(void* h, int unlink_files) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); ((HostTable*)h)->ssd_close(unlink_files != 0); return 0; }
int64_t pbh_table_spill(void* h, float min_unseen_days, int64_t max_mem_rows) {
  std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  return ((HostTable*)h)->spill(min_unseen_days, max_mem_rows, table_pool());
}
int64_t pbh_table_load_ssd2mem(void* h) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); return ((HostTable*)h)->load_ssd2mem(table_pool()); }
int64_t pbh_table_mem_rows(void* h) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); return (int64_t)((HostTable*)h)->mem_rows(); }
int64_t pbh_table_ssd_rows(void* h) { std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu); return (int64_t)((HostTable*)h)->ssd_rows(); }
int64_t pbh_table_keys(void* h, uint64_t* out, int64_t max_n) {
  std::vector<uint64_t> ks;
  std::unique_lock<std::shared_mutex> lk(((HostTable*)h)->table_mu);
  ((HostTable*)h)->all_keys(&ks);
  int64_t n = std::min<int64_t>(max_n, (int64_t)ks.size());
  if (out && n > 0) memcpy(out, ks.data(), (size_t)n * 8);
  return (int64_t)ks.size();
}

}  // extern "C"
