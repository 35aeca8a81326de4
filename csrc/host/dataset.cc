// See dataset.h for the reference map.
#include "dataset.h"
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <dlfcn.h>
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <sstream>

// ---- InputTable key index: string key -> row offset (0 = miss) ---------------------------------
// (InputTable::GetIndexOffset / AddIndexData, fleet/box_wrapper.h:124-197)
namespace {
struct InputIndex {
  std::mutex mu;
  std::unordered_map<std::string, uint64_t> map;
  std::atomic<uint64_t> miss{0};
};
}  // namespace
extern "C" {
int pbh_input_index_create(void** out) { *out = new InputIndex(); return 0; }
int pbh_input_index_destroy(void* h) { delete (InputIndex*)h; return 0; }
// offset of `key`, adding it at the next offset (1, 2, ...) when new; *fresh tells which
uint64_t pbh_input_index_add(void* h, const char* key, size_t len, int* fresh) {
  InputIndex* ix = (InputIndex*)h;
  std::lock_guard<std::mutex> g(ix->mu);
  auto r = ix->map.emplace(std::string(key, len), (uint64_t)ix->map.size() + 1);
  if (fresh) *fresh = r.second ? 1 : 0;
  return r.first->second;
}
uint64_t pbh_input_index_get(void* h, const char* key, size_t len) {
  InputIndex* ix = (InputIndex*)h;
  std::lock_guard<std::mutex> g(ix->mu);
  auto it = ix->map.find(std::string(key, len));
  if (it == ix->map.end()) { ++ix->miss; return 0; }
  return it->second;
}
uint64_t pbh_input_index_size(void* h) { InputIndex* ix = (InputIndex*)h; std::lock_guard<std::mutex> g(ix->mu); return ix->map.size(); }
uint64_t pbh_input_index_miss(void* h) { return ((InputIndex*)h)->miss.load(); }
}
static uint64_t pbh_input_index_get_cb(void* user, const char* key, size_t len) { return pbh_input_index_get(user, key, len); }

namespace pbh {

static thread_local std::string g_err;
void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
}
const char* last_error() { return g_err.c_str(); }

bool ThreadPool::set_cpu_affinity(std::thread& t, int core) {
  cpu_set_t set;
  CPU_ZERO(&set);
  CPU_SET(core, &set);
  return pthread_setaffinity_np(t.native_handle(), sizeof(set), &set) == 0;
}

// ---------------------------------------------------------------- KeyAgent
// Keys are bucketed into thread-local per-shard buffers; a buffer goes to its shard's set (one
// lock) only when it holds a few thousand keys.  Every thread that added keys must call
// flush_local() before the shards are read (merge_loop does).
static thread_local std::vector<std::vector<uint64_t>> tl_agent_local;

void KeySet::grow() {
  std::vector<uint64_t> old;
  old.swap(tab);
  tab.assign(old.empty() ? 1024 : old.size() * 2, 0);
  const size_t mask = tab.size() - 1;
  for (uint64_t k : old) {
    if (!k) continue;
    size_t p = (size_t)(mix64(k) >> 20) & mask;
    while (tab[p]) p = (p + 1) & mask;
    tab[p] = k;
  }
}

inline void KeySet::insert(uint64_t k) {
  if (k == 0) { has_zero = true; return; }
  if ((cnt + 1) * 2 > tab.size()) grow();
  const size_t mask = tab.size() - 1;
  size_t p = (size_t)(mix64(k) >> 20) & mask;       // (the low bits chose the shard)
  while (tab[p]) {
    if (tab[p] == k) return;
    p = (p + 1) & mask;
  }
  tab[p] = k;
  ++cnt;
}

void KeySet::dump_sorted(std::vector<uint64_t>* out) const {
  out->clear();
  out->reserve(cnt + 1);
  if (has_zero) out->push_back(0);
  for (uint64_t k : tab) if (k) out->push_back(k);
  radix_sort_u64(*out);
}

void KeyAgent::flush_bucket(size_t s, std::vector<uint64_t>& b) {
  std::lock_guard<std::mutex> g(mu_[s]);
  KeySet& set = shards_[s];
  // make room first so the table does not move under the prefetches, then insert with the home
  // slot of the key 8 positions ahead already on its way from memory (random-insert bound)
  while ((set.cnt + b.size() + 1) * 2 > set.tab.size()) set.grow();
  const size_t mask = set.tab.size() - 1, nb = b.size();
  for (size_t i = 0; i < nb; ++i) {
    if (i + 8 < nb) __builtin_prefetch(&set.tab[(size_t)(mix64(b[i + 8]) >> 20) & mask], 1, 1);
    set.insert(b[i]);
  }
  b.clear();
}

// A pass repeats its hot keys thousands of times.  Each adding thread keeps a small direct-mapped table of
// the keys it has sent recently (256 KB, stays in L2): a key found there is already in its shard's buffer or
// set and is dropped at once, so only first sightings (and evicted keys) pay for the bucket + shard insert.
// The table is wiped when the agent's stamp changes: cleared for a new pass, or another agent on this thread.
static constexpr size_t kRecentSlots = 1 << 15;
static thread_local std::vector<uint64_t> tl_recent;
static thread_local uint64_t tl_recent_epoch = 0;

void KeyAgent::add_keys(const uint64_t* keys, size_t n) {
  const size_t S = shards_.size();
  const bool pow2 = (S & (S - 1)) == 0;
  auto& local = tl_agent_local;
  if (local.size() != S) local.assign(S, {});
  auto& recent = tl_recent;
  if (recent.size() != kRecentSlots || tl_recent_epoch != epoch_.load(std::memory_order_relaxed)) {
    recent.assign(kRecentSlots, 0);
    tl_recent_epoch = epoch_.load(std::memory_order_relaxed);
  }
  for (size_t i = 0; i < n; ++i) {
    const uint64_t k = keys[i];
    const uint64_t h = mix64(k);
    if (k != 0) {
      uint64_t& slot = recent[(size_t)(h >> 40) & (kRecentSlots - 1)];
      if (slot == k) continue;
      slot = k;
    }
    const size_t s = pow2 ? (size_t)(h & (S - 1)) : (size_t)(h % S);
    auto& b = local[s];
    b.push_back(k);
    if (b.size() >= 4096) flush_bucket(s, b);
  }
}

void KeyAgent::flush_local() {
  auto& local = tl_agent_local;
  for (size_t s = 0; s < local.size() && s < shards_.size(); ++s)
    if (!local[s].empty()) flush_bucket(s, local[s]);
}

void KeyAgent::finalize(ThreadPool* pool, std::vector<uint64_t>* out) {
  std::vector<std::vector<uint64_t>> parts(shards_.size());
  pool->parallel_range(shards_.size(), [&](size_t b, size_t e, int) {
    for (size_t s = b; s < e; ++s) shards_[s].dump_sorted(&parts[s]);
  });
  size_t total = 0;
  for (auto& v : parts) total += v.size();
  out->clear();
  out->reserve(total);
  for (auto& v : parts) out->insert(out->end(), v.begin(), v.end());
}

// ---------------------------------------------------------------- batch dealing
static void left_batches(int64_t ins_num, int nworkers, std::vector<BatchRange>* out, int64_t start) {
  int64_t cur = start;
  int64_t bs = ins_num / nworkers, left = ins_num % nworkers;
  for (int i = 0; i < nworkers; ++i) {
    int64_t n = bs + (i == 0 ? left : 0);
    out->push_back({cur, (int)n});
    cur += n;
  }
}

void compute_batch_ranges(int64_t ins_num, int batch_size, int nworkers, std::vector<BatchRange>* out) {
  out->clear();
  const int64_t round = (int64_t)batch_size * nworkers;
  if (round > ins_num) {  // less than one full round: split evenly
    left_batches(ins_num, nworkers, out, 0);
    return;
  }
  int64_t cur = 0;
  int64_t full = ins_num / round * nworkers;
  int64_t left = ins_num % round;
  if (left > 0 && left < 2 * nworkers && full > 1) {
    // tiny tail: fold it into the last full round and split that into two even rounds
    full -= nworkers;
    left += round;
    for (int64_t i = 0; i < full; ++i) { out->push_back({cur, batch_size}); cur += batch_size; }
    left_batches(left, nworkers * 2, out, cur);
  } else {
    for (int64_t i = 0; i < full; ++i) { out->push_back({cur, batch_size}); cur += batch_size; }
    if (left > 0) left_batches(left, nworkers, out, cur);
  }
}

// ---------------------------------------------------------------- Dataset
Dataset::Dataset() : chan_(64) {}
Dataset::~Dataset() {
  wait_preload_done();
  release_memory();
  if (plugin_handle_) dlclose(plugin_handle_);
}

int Dataset::finalize_slots() {
  sparse_idx_.assign(slots.size(), -1);
  float_idx_.assign(slots.size(), -1);
  float_dim_.clear();
  float_off_.clear();
  n_used_u64_ = n_used_f_ = 0;
  dense_dim_ = 0;
  for (size_t i = 0; i < slots.size(); ++i) {
    if (!slots[i].is_used) continue;
    if (slots[i].type == 0) {
      sparse_idx_[i] = n_used_u64_++;
    } else {
      float_idx_[i] = n_used_f_++;
      float_dim_.push_back(std::max(1, slots[i].dim));
      float_off_.push_back(dense_dim_);
      dense_dim_ += std::max(1, slots[i].dim);
    }
  }
  // compact sparse_idx_ list of used sparse slots for n_sparse()
  std::vector<int> used;
  for (size_t i = 0; i < slots.size(); ++i) if (sparse_idx_[i] >= 0) used.push_back((int)i);
  sparse_idx_.swap(used);  // now: list of slot indices that are used sparse slots
  // rebuild per-slot lookup
  slot_to_sparse_.assign(slots.size(), -1);
  for (size_t p = 0; p < sparse_idx_.size(); ++p) slot_to_sparse_[sparse_idx_[p]] = (int)p;
  return bind_plugin();
}

// opens the parser plugin (once) and hands it the InputTable lookup when both exist
int Dataset::bind_plugin() {
  if (!parser_plugin.empty() && !plugin_handle_) {
    plugin_handle_ = dlopen(parser_plugin.c_str(), RTLD_NOW | RTLD_LOCAL);
    PBH_REQUIRE(plugin_handle_ != nullptr, "cannot dlopen parser plugin %s: %s", parser_plugin.c_str(), dlerror());
    plugin_fn_ = dlsym(plugin_handle_, "pbx_parse_line");
    plugin_file_fn_ = dlsym(plugin_handle_, "pbx_parse_file");
    PBH_REQUIRE(plugin_fn_ != nullptr || plugin_file_fn_ != nullptr,
                "parser plugin %s exports neither pbx_parse_line nor pbx_parse_file", parser_plugin.c_str());
  }
  if (plugin_handle_ && input_index) {
    // optional: void pbx_set_offset_func(uint64_t (*fn)(void* user, const char* key, size_t len), void* user)
    typedef void (*set_fn_t)(uint64_t (*)(void*, const char*, size_t), void*);
    if (auto set_fn = (set_fn_t)dlsym(plugin_handle_, "pbx_set_offset_func")) set_fn(&pbh_input_index_get_cb, input_index);
  }
  return 0;
}

static inline const char* skip_ws(const char* p, const char* e) {
  while (p < e && (*p == ' ' || *p == '\t')) ++p;
  return p;
}

static inline bool read_token(const char*& p, const char* e, const char** tb, size_t* tl) {
  p = skip_ws(p, e);
  if (p >= e) return false;
  *tb = p;
  while (p < e && *p != ' ' && *p != '\t' && *p != '\n' && *p != '\r') ++p;
  *tl = (size_t)(p - *tb);
  return *tl > 0;
}

static inline bool read_u64(const char*& p, const char* e, uint64_t* v) {
  p = skip_ws(p, e);
  if (p >= e || *p < '0' || *p > '9') return false;
  uint64_t x = 0;
  while (p < e && *p >= '0' && *p <= '9') { x = x * 10 + (uint64_t)(*p - '0'); ++p; }
  *v = x;
  return true;
}

static inline bool read_float(const char*& p, const char* e, float* v) {
  p = skip_ws(p, e);
  if (p >= e) return false;
  // fast path: [-+]digits[.digits] with at most 18 significant digits and no exponent (what data
  // generators emit); anything else (exponents, inf/nan, long mantissas) goes through strtof
  const char* q = p;
  bool neg = false;
  if (*q == '-' || *q == '+') { neg = *q == '-'; ++q; }
  uint64_t mant = 0;
  int digits = 0, frac = 0;
  const char* d0 = q;
  while (q < e && *q >= '0' && *q <= '9') { mant = mant * 10 + (uint64_t)(*q - '0'); ++q; ++digits; }
  if (q < e && *q == '.') {
    ++q;
    while (q < e && *q >= '0' && *q <= '9') { mant = mant * 10 + (uint64_t)(*q - '0'); ++q; ++digits; ++frac; }
  }
  const bool ends = q >= e || *q == ' ' || *q == '\t' || *q == '\n' || *q == '\r';
  if (digits > 0 && digits <= 18 && frac <= 18 && ends && q > d0) {
    static const double kPow10[19] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13,
                                      1e14, 1e15, 1e16, 1e17, 1e18};
    const double x = (double)mant / kPow10[frac];     // exact mantissa, one correctly rounded division
    *v = (float)(neg ? -x : x);
    p = q;
    return true;
  }
  char* end = nullptr;
  *v = strtof(p, &end);
  if (end == p) return false;
  p = end;
  return true;
}

// plugin ABI: the plugin parses one line and reports every slot through `emit`
//   int pbx_parse_line(const char* line, size_t len, void* ctx,
//                      void (*emit_u64)(void*, int slot, const uint64_t*, size_t),
//                      void (*emit_f32)(void*, int slot, const float*, size_t),
//                      void (*emit_meta)(void*, const char* ins_id, uint64_t search_id, uint32_t rank, uint32_t cmatch));
// Role of the reference's ISlotParser .so plugins (data_feed.cc:3604-3670).
struct PluginCtx {
  std::vector<std::vector<uint64_t>> u64;
  std::vector<std::vector<float>> f32;
  std::string ins_id;
  uint64_t search_id = 0;
  uint32_t rank = 0, cmatch = 0;
  void* owner = nullptr;         // whole-file mode: the FileCtx this belongs to
};
typedef int (*plugin_fn_t)(const char*, size_t, void*, void (*)(void*, int, const uint64_t*, size_t),
                           void (*)(void*, int, const float*, size_t),
                           void (*)(void*, const char*, uint64_t, uint32_t, uint32_t));
static void plug_emit_u64(void* c, int slot, const uint64_t* v, size_t n) {
  auto* ctx = (PluginCtx*)c;
  if (slot >= 0 && slot < (int)ctx->u64.size()) ctx->u64[slot].assign(v, v + n);
}
static void plug_emit_f32(void* c, int slot, const float* v, size_t n) {
  auto* ctx = (PluginCtx*)c;
  if (slot >= 0 && slot < (int)ctx->f32.size()) ctx->f32[slot].assign(v, v + n);
}
static void plug_emit_meta(void* c, const char* ins, uint64_t sid, uint32_t rank, uint32_t cmatch) {
  auto* ctx = (PluginCtx*)c;
  if (ins) ctx->ins_id = ins;
  ctx->search_id = sid; ctx->rank = rank; ctx->cmatch = cmatch;
}

// one record from what a parser plugin emitted (zero feasigns / tiny floats of sparse slots dropped like
// the text parser does)
bool Dataset::record_from_plugin(void* vctx, SlotRecord* rec) const {
  PluginCtx& ctx = *(PluginCtx*)vctx;
  rec->reset();
  rec->u64_offs.assign(n_used_u64_ + 1, 0);
  rec->f_offs.assign(n_used_f_ + 1, 0);
  thread_local std::vector<std::vector<uint64_t>> u64s;
  thread_local std::vector<std::vector<float>> f32s;
  u64s.resize(n_used_u64_);
  f32s.resize(n_used_f_);
  for (auto& v : u64s) v.clear();
  for (auto& v : f32s) v.clear();
  rec->ins_id = ctx.ins_id; rec->search_id = ctx.search_id; rec->rank = ctx.rank; rec->cmatch = ctx.cmatch;
  for (size_t s = 0; s < slots.size(); ++s) {
    if (!slots[s].is_used) continue;
    if (slots[s].type == 0) {
      auto& dst = u64s[slot_to_sparse_[s]];
      for (uint64_t v : ctx.u64[s]) if (v != 0 || slots[s].is_dense) dst.push_back(v);
    } else {
      auto& dst = f32s[float_idx_[s]];
      for (float v : ctx.f32[s]) if (slots[s].is_dense || std::fabs(v) >= 1e-6f) dst.push_back(v);
    }
  }
  size_t total_u64 = 0;
  for (int s = 0; s < n_used_u64_; ++s) {
    rec->u64_vals.insert(rec->u64_vals.end(), u64s[s].begin(), u64s[s].end());
    total_u64 += u64s[s].size();
    rec->u64_offs[s + 1] = (uint32_t)total_u64;
  }
  size_t total_f = 0;
  for (int s = 0; s < n_used_f_; ++s) {
    rec->f_vals.insert(rec->f_vals.end(), f32s[s].begin(), f32s[s].end());
    total_f += f32s[s].size();
    rec->f_offs[s + 1] = (uint32_t)total_f;
  }
  return !(n_used_u64_ > 0 && total_u64 == 0);          // no feasign at all (data_feed.cc:4133)
}

bool Dataset::parse_line(const char* line, size_t len, SlotRecord* rec) const {
  const char* p = line;
  const char* e = line + len;
  while (e > p && (e[-1] == '\n' || e[-1] == '\r')) --e;
  if (plugin_fn_) {
    PluginCtx ctx;
    ctx.u64.resize(slots.size());
    ctx.f32.resize(slots.size());
    if (((plugin_fn_t)plugin_fn_)(line, (size_t)(e - line), &ctx, plug_emit_u64, plug_emit_f32, plug_emit_meta) != 0)
      return false;
    return record_from_plugin(&ctx, rec);
  }
  rec->reset();
  rec->u64_offs.assign(n_used_u64_ + 1, 0);
  rec->f_offs.assign(n_used_f_ + 1, 0);
  {
    const char* tb; size_t tl; uint64_t n;
    if (parse_ins_id) {
      if (!read_u64(p, e, &n) || n != 1 || !read_token(p, e, &tb, &tl)) return false;
      rec->ins_id.assign(tb, tl);
    }
    if (parse_logkey) {
      if (!read_u64(p, e, &n) || n != 1 || !read_token(p, e, &tb, &tl)) return false;
      // logkey hex layout: [11,14) cmatch, [14,16) rank, [16,32) search_id (data_feed.cc:2385-2396)
      rec->ins_id.assign(tb, tl);         // the logkey becomes the instance id (data_feed.cc:4071)
      if (tl >= 32) {
        std::string lk(tb, tl);
        rec->search_id = strtoull(lk.substr(16, 16).c_str(), nullptr, 16);
        rec->cmatch = (uint32_t)strtoul(lk.substr(11, 3).c_str(), nullptr, 16);
        rec->rank = (uint32_t)strtoul(lk.substr(14, 2).c_str(), nullptr, 16);
      }
    }
    for (size_t s = 0; s < slots.size(); ++s) {
      if (!read_u64(p, e, &n) || n == 0) return false;  // every slot needs >= 1 value (data_feed.cc:4086-4092)
      const SlotDesc& sd = slots[s];
      if (!sd.is_used) {  // skip token-wise
        for (uint64_t i = 0; i < n; ++i) if (!read_token(p, e, &tb, &tl)) return false;
        continue;
      }
      if (sd.type == 0) {
        // used sparse slots are numbered in file order: append straight into the record
        for (uint64_t i = 0; i < n; ++i) {
          uint64_t v;
          if (!read_u64(p, e, &v)) return false;
          if (v != 0 || sd.is_dense) rec->u64_vals.push_back(v);  // zero feasigns dropped (data_feed.cc:4099)
        }
        rec->u64_offs[slot_to_sparse_[s] + 1] = (uint32_t)rec->u64_vals.size();
      } else {
        for (uint64_t i = 0; i < n; ++i) {
          float v;
          if (!read_float(p, e, &v)) return false;
          if (sd.is_dense || std::fabs(v) >= 1e-6f) rec->f_vals.push_back(v);  // (data_feed.cc:4111)
        }
        rec->f_offs[float_idx_[s] + 1] = (uint32_t)rec->f_vals.size();
      }
    }
    if (n_used_u64_ > 0 && rec->u64_vals.empty()) return false;  // no feasign at all (data_feed.cc:4133)
    return true;
  }
}

// ---- whole-file parser plugins (FLAGS_enable_ins_parser_file: ISlotParser::ParseFileInstance,
// data_feed.cc:3795-3883).  Optional plugin export:
//   int pbx_parse_file(const char* path, void* ctx,
//                      void (*begin_record)(void*), emit_u64, emit_f32, emit_meta,   // as in pbx_parse_line
//                      int (*end_record)(void*));      // returns 0 when the record was kept
// The plugin opens and reads the file itself (any container format) and reports record after record.
struct FileCtx {
  PluginCtx cur;                 // the plugin's ctx pointer is &cur; cur.owner leads back here
  const Dataset* ds = nullptr;
  std::function<void(SlotRecord*)> sink;
  std::function<SlotRecord*()> alloc;
  std::function<void(SlotRecord*)> discard;
  int64_t good = 0, bad = 0;
};
static void file_begin_record(void* c) {
  auto* f = (FileCtx*)((PluginCtx*)c)->owner;
  for (auto& v : f->cur.u64) v.clear();
  for (auto& v : f->cur.f32) v.clear();
  f->cur.ins_id.clear();
  f->cur.search_id = 0; f->cur.rank = 0; f->cur.cmatch = 0;
}
static int file_end_record(void* c) {
  auto* f = (FileCtx*)((PluginCtx*)c)->owner;
  SlotRecord* rec = f->alloc();
  if (!f->ds->record_from_plugin(&f->cur, rec)) { f->discard(rec); ++f->bad; return 1; }
  f->sink(rec);
  ++f->good;
  return 0;
}
typedef int (*plugin_file_fn_t)(const char*, void*, void (*)(void*), void (*)(void*, int, const uint64_t*, size_t),
                                void (*)(void*, int, const float*, size_t),
                                void (*)(void*, const char*, uint64_t, uint32_t, uint32_t), int (*)(void*));

int Dataset::read_file_with_plugin(const std::string& path) {
  FileCtx fc;
  fc.ds = this;
  fc.cur.owner = &fc;
  fc.cur.u64.resize(slots.size());
  fc.cur.f32.resize(slots.size());
  std::vector<SlotRecord*> block;
  block.reserve(4096);
  int64_t feas = 0;
  fc.alloc = [this] { return pool_.get(); };
  fc.discard = [this](SlotRecord* r) { std::vector<SlotRecord*> one{r}; pool_.put(one); };
  fc.sink = [&](SlotRecord* rec) {
    feas += (int64_t)rec->u64_vals.size();
    block.push_back(rec);
    if (block.size() >= 4096) {
      std::vector<SlotRecord*> out;
      out.swap(block);
      chan_.put(std::move(out));
      block.reserve(4096);
    }
  };
  int rc = ((plugin_file_fn_t)plugin_file_fn_)(path.c_str(), &fc.cur, file_begin_record, plug_emit_u64, plug_emit_f32,
                                               plug_emit_meta, file_end_record);
  if (!block.empty()) chan_.put(std::move(block));
  lines_read_ += fc.good + fc.bad;
  lines_bad_ += fc.bad;
  total_feasigns_ += feas;
  return rc == 0 ? 0 : -1;
}

int Dataset::read_file(const std::string& path) {
  if (parser_file_mode && plugin_file_fn_) return read_file_with_plugin(path);
  FILE* fp = nullptr;
  bool piped = false;
  if (pipe_command.empty() || pipe_command == "cat") {
    fp = fopen(path.c_str(), "r");
  } else {
    std::string cmd = "cat '" + path + "' | " + pipe_command;
    fp = popen(cmd.c_str(), "r");
    piped = true;
  }
  if (!fp) return -1;
  // Block reads into one buffer, lines found with memchr and parsed in place (BufferedLineFileReader; no
  // per-line copy).  Records come from a stash refilled from the pool 256 at a time (one lock per refill) and
  // new ones get their value vectors reserved at the size the previous lines needed (a fresh vector would
  // otherwise grow eight times per line).
  std::vector<char> buf(4 << 20);
  size_t have = 0;                    // bytes of an unfinished line carried to the front of the buffer
  std::mt19937_64 rng(mix64(std::hash<std::string>()(path)));
  std::uniform_real_distribution<float> uni(0.f, 1.f);
  std::vector<SlotRecord*> block, stash;
  block.reserve(4096);
  int64_t feas = 0;
  size_t hint_u64 = 64, hint_f = 16;
  auto next_record = [&]() -> SlotRecord* {
    if (stash.empty()) pool_.get_many(256, &stash);
    SlotRecord* r;
    if (!stash.empty()) { r = stash.back(); stash.pop_back(); }
    else r = new SlotRecord();
    if (r->u64_vals.capacity() == 0) { r->u64_vals.reserve(hint_u64); r->f_vals.reserve(hint_f); }
    return r;
  };
  SlotRecord* rec = nullptr;
  auto handle_line = [&](const char* line, size_t len) {
    if (sample_rate < 1.0f && uni(rng) > sample_rate) return;  // Bernoulli sampling
    lines_read_++;
    if (!rec) rec = next_record();
    if (!parse_line(line, len, rec)) { lines_bad_++; return; }   // (the record is reused for the next line)
    feas += (int64_t)rec->u64_vals.size();
    hint_u64 = std::max(hint_u64, rec->u64_vals.size());
    hint_f = std::max(hint_f, rec->f_vals.size());
    block.push_back(rec);
    rec = nullptr;
    if (block.size() >= 4096) {
      std::vector<SlotRecord*> out;
      out.swap(block);
      chan_.put(std::move(out));
      block.reserve(4096);
    }
  };
  for (;;) {
    if (have == buf.size()) buf.resize(buf.size() * 2);          // a single line longer than the buffer
    size_t got = fread(buf.data() + have, 1, buf.size() - have, fp);
    size_t end = have + got;
    const char* p = buf.data();
    const char* e = p + end;
    while (p < e) {
      const char* nl = (const char*)memchr(p, '\n', (size_t)(e - p));
      if (!nl) break;
      handle_line(p, (size_t)(nl - p));
      p = nl + 1;
    }
    have = (size_t)(e - p);
    if (got == 0) {                                              // end of input: the last line may lack its newline
      if (have) handle_line(p, have);
      break;
    }
    if (have && p != buf.data()) memmove(buf.data(), p, have);
  }
  if (rec) stash.push_back(rec);
  if (!stash.empty()) pool_.put(stash);
  if (!block.empty()) chan_.put(std::move(block));
  total_feasigns_ += feas;
  if (piped) pclose(fp); else fclose(fp);
  return 0;
}

void Dataset::reader_loop(std::atomic<size_t>* next_file, std::atomic<int>* errors) {
  for (;;) {
    size_t i = next_file->fetch_add(1);
    // files are dealt round-robin over nodes first (SetFileList, data_set.cc:1963-1975)
    size_t gi = i * (size_t)world + (size_t)rank;
    if (gi >= files.size()) break;
    if (read_file(files[gi]) != 0) (*errors)++;
  }
}

void Dataset::collect_record_keys(const SlotRecord* r) {
  if (omit_slots.empty()) {
    agent_.add_keys(r->u64_vals.data(), r->u64_vals.size());
    return;
  }
  for (size_t p = 0; p + 1 < r->u64_offs.size() && p < sparse_idx_.size(); ++p) {
    if (std::find(omit_slots.begin(), omit_slots.end(), sparse_idx_[p]) != omit_slots.end()) continue;
    agent_.add_keys(r->u64_vals.data() + r->u64_offs[p], r->u64_offs[p + 1] - r->u64_offs[p]);
  }
}

void Dataset::merge_loop() {
  std::vector<SlotRecord*> block;
  std::vector<SlotRecord*> mine;
  while (chan_.get(block)) {
    if (collect_keys)
      for (SlotRecord* r : block) collect_record_keys(r);
    mine.insert(mine.end(), block.begin(), block.end());
    block.clear();
    if (mine.size() >= 65536) {
      std::lock_guard<std::mutex> g(rec_mu_);
      records_.insert(records_.end(), mine.begin(), mine.end());
      mine.clear();
    }
  }
  agent_.flush_local();
  std::lock_guard<std::mutex> g(rec_mu_);
  records_.insert(records_.end(), mine.begin(), mine.end());
}

int Dataset::load_into_memory() {
  drop_arena();
  PBH_REQUIRE(!slots.empty(), "dataset has no slots (call set_use_var first)");
  if (slot_to_sparse_.size() != slots.size()) {
    int rc = finalize_slots();
    if (rc) return rc;
  }
  if (int rc = bind_plugin()) return rc;
  Timer t;
  t.start();
  chan_.open();
  agent_.clear();
  if (collect_keys) agent_.add_key(0);  // key 0 is always registered (box_wrapper.h:1107)
  std::atomic<size_t> next_file{0};
  std::atomic<int> errors{0};
  std::vector<std::thread> readers, mergers;
  int nr = std::max(1, thread_num), nm = std::max(1, merge_thread_num);
  // every thread pins ITSELF: pinning from here through the std::thread handle would race with short-lived
  // threads — a thread that has already exited has tid 0 in glibc, and sched_setaffinity(0, ...) then pins
  // the caller instead
  auto pin_self = [this](size_t k) {
    if (reader_cores.empty()) return;
    cpu_set_t set;
    CPU_ZERO(&set);
    CPU_SET(reader_cores[k % reader_cores.size()], &set);
    pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
  };
  for (int i = 0; i < nm; ++i) mergers.emplace_back([this, pin_self, i, nr] { pin_self((size_t)(nr + i)); merge_loop(); });
  for (int i = 0; i < nr; ++i)
    readers.emplace_back([&, pin_self, i] { pin_self((size_t)i); reader_loop(&next_file, &errors); });
  for (auto& th : readers) th.join();
  chan_.close();
  for (auto& th : mergers) th.join();
  t.pause();
  load_seconds_ = t.seconds();
  Timer t2;
  t2.start();
  if (collect_keys) {
    ThreadPool pool(nm);
    agent_.finalize(&pool, &unique_keys_);
    agent_.clear();
  }
  t2.pause();
  merge_seconds_ = t2.seconds();
  PBH_REQUIRE(errors.load() == 0, "%d input file(s) could not be opened", errors.load());
  return 0;
}

int Dataset::preload_into_memory() {
  wait_preload_done();
  preload_rc_ = 0;
  preload_thread_.reset(new std::thread([this] {
    preload_rc_ = load_into_memory();
    if (preload_rc_) preload_err_ = last_error();
  }));
  return 0;
}

int Dataset::wait_preload_done() {
  if (preload_thread_) {
    preload_thread_->join();
    preload_thread_.reset();
    if (preload_rc_) set_error("%s", preload_err_.c_str());
    return preload_rc_;
  }
  return 0;
}

void Dataset::release_memory() {
  drop_arena();
  pool_.put(records_);
  records_.shrink_to_fit();
  worker_batches_.clear();
  unique_keys_.clear();
  unique_keys_.shrink_to_fit();
  total_feasigns_ = 0;
  lines_read_ = 0;
  lines_bad_ = 0;
}

void Dataset::local_shuffle(uint64_t seed) {
  if (arena_only_) return;          // fixed record order; prepare_train shuffles the batches
  drop_arena();
  std::mt19937_64 rng(seed);
  std::shuffle(records_.begin(), records_.end(), rng);
}

// AUC-runner: permute the feasigns of the chosen slots across records, keeping every other
// slot in place, to measure how much the model relies on those slots
// (reference RecordReplace / slots_shuffle, data_feed.h:1772-1962, box_wrapper.cc:212-368).
int Dataset::slots_shuffle(const std::vector<int>& slot_idx, uint64_t seed, bool inverse) {
  PBH_REQUIRE(!arena_only_, "slots_shuffle needs records: not available for a pass loaded from a pass file");
  const bool had_arena = arena_ok_;
  drop_arena();
  const size_t n = records_.size();
  if (n < 2) return 0;
  std::vector<size_t> perm(n);
  for (size_t i = 0; i < n; ++i) perm[i] = i;
  std::mt19937_64 rng(seed);
  std::shuffle(perm.begin(), perm.end(), rng);
  if (inverse) {  // undo a previous shuffle with the same seed (reference RecordReplaceBack)
    std::vector<size_t> inv(n);
    for (size_t i = 0; i < n; ++i) inv[perm[i]] = i;
    perm.swap(inv);
  }
  std::vector<char> chosen(n_used_u64_, 0);
  for (int s : slot_idx) {
    PBH_REQUIRE(s >= 0 && s < (int)slots.size() && slot_to_sparse_[s] >= 0, "slot %d is not a used sparse slot", s);
    chosen[slot_to_sparse_[s]] = 1;
  }
  // snapshot donors' values of the chosen slots
  std::vector<std::vector<std::vector<uint64_t>>> donor(n);
  for (size_t i = 0; i < n; ++i) {
    donor[i].resize(n_used_u64_);
    const SlotRecord* r = records_[i];
    for (int p = 0; p < n_used_u64_; ++p)
      if (chosen[p]) donor[i][p].assign(r->u64_vals.begin() + r->u64_offs[p], r->u64_vals.begin() + r->u64_offs[p + 1]);
  }
  for (size_t i = 0; i < n; ++i) {
    SlotRecord* r = records_[i];
    std::vector<uint64_t> nv;
    std::vector<uint32_t> no(n_used_u64_ + 1, 0);
    for (int p = 0; p < n_used_u64_; ++p) {
      if (chosen[p]) nv.insert(nv.end(), donor[perm[i]][p].begin(), donor[perm[i]][p].end());
      else nv.insert(nv.end(), r->u64_vals.begin() + r->u64_offs[p], r->u64_vals.begin() + r->u64_offs[p + 1]);
      no[p + 1] = (uint32_t)nv.size();
    }
    r->u64_vals.swap(nv);
    r->u64_offs.swap(no);
  }
  if (had_arena) build_arena();   // same order, same batches: only the values changed
  return 0;
}

int Dataset::slot_first_values(int slot, uint64_t* out, uint8_t* has) const {
  PBH_REQUIRE(!arena_only_, "slot_first_values needs records: not available for a pass loaded from a pass file");
  PBH_REQUIRE(slot >= 0 && slot < (int)slots.size() && slot_to_sparse_[slot] >= 0, "slot %d is not a used sparse slot", slot);
  const int p = slot_to_sparse_[slot];
  for (size_t i = 0; i < records_.size(); ++i) {
    const SlotRecord* r = records_[i];
    const bool ok = p + 1 < (int)r->u64_offs.size() && r->u64_offs[p + 1] > r->u64_offs[p];
    has[i] = ok ? 1 : 0;
    out[i] = ok ? r->u64_vals[r->u64_offs[p]] : 0;
  }
  return 0;
}

int Dataset::expand_records(int slot, int label_slot, const int64_t* counts, const uint64_t* ids, const uint8_t* neg) {
  PBH_REQUIRE(!arena_only_, "expand_records needs records: not available for a pass loaded from a pass file");
  PBH_REQUIRE(slot >= 0 && slot < (int)slots.size() && slot_to_sparse_[slot] >= 0, "slot %d is not a used sparse slot", slot);
  int lp = -1, lf = -1;
  if (label_slot >= 0) {
    PBH_REQUIRE(label_slot < (int)slots.size(), "label slot %d out of range", label_slot);
    lp = slot_to_sparse_[label_slot];
    lf = lp < 0 ? float_idx_[label_slot] : -1;
    PBH_REQUIRE(lp >= 0 || lf >= 0, "label slot %d is not a used slot", label_slot);
  }
  drop_arena();
  const int p = slot_to_sparse_[slot];
  std::vector<SlotRecord*> out, dropped;
  int64_t k = 0, total = 0;
  for (size_t i = 0; i < records_.size(); ++i) {
    PBH_REQUIRE(counts[i] >= 0, "negative copy count for record %zu", i);
    total += counts[i];
  }
  out.reserve((size_t)total);
  for (size_t i = 0; i < records_.size(); ++i) {
    SlotRecord* r = records_[i];
    const bool has = p + 1 < (int)r->u64_offs.size() && r->u64_offs[p + 1] > r->u64_offs[p];
    if (counts[i] == 0 || !has) { dropped.push_back(r); k += counts[i]; continue; }
    for (int64_t c = 0; c < counts[i]; ++c, ++k) {
      SlotRecord* n = (c + 1 == counts[i]) ? r : new SlotRecord(*r);   // the original is the last copy
      n->u64_vals[n->u64_offs[p]] = ids[k];
      if (neg[k]) {
        if (lp >= 0 && lp + 1 < (int)n->u64_offs.size() && n->u64_offs[lp + 1] > n->u64_offs[lp]) n->u64_vals[n->u64_offs[lp]] = 0;
        if (lf >= 0 && lf + 1 < (int)n->f_offs.size() && n->f_offs[lf + 1] > n->f_offs[lf]) n->f_vals[n->f_offs[lf]] = 0.f;
      }
      out.push_back(n);
    }
  }
  records_.swap(out);
  if (!dropped.empty()) pool_.put(dropped);
  int64_t feas = 0;
  for (const SlotRecord* r : records_) feas += (int64_t)r->u64_vals.size();
  total_feasigns_ = feas;
  if (collect_keys) {      // the sampled node ids are embedding keys of this pass too
    unique_keys_.insert(unique_keys_.end(), ids, ids + total);
    radix_sort_u64(unique_keys_);
    unique_keys_.erase(std::unique(unique_keys_.begin(), unique_keys_.end()), unique_keys_.end());
  }
  return 0;
}

// Merge the records that share an instance id into one (reference MultiSlotDataset::MergeByInsId,
// data_set.cc:1387-1581): groups whose size differs from merge_size (> 0) are dropped; a sparse
// slot may carry values in at most one record of a group (else the group is dropped as
// conflicting); of a float slot the first non-empty, non-zero copy wins.  Returns the number of
// dropped instances.
int64_t Dataset::merge_by_ins_id(int merge_size) {
  if (arena_only_) return 0;        // a pass file holds no instance ids
  drop_arena();
  std::stable_sort(records_.begin(), records_.end(),
                   [](const SlotRecord* a, const SlotRecord* b) { return a->ins_id < b->ins_id; });
  std::vector<SlotRecord*> out, dropped;
  int64_t n_drop = 0;
  for (size_t i = 0; i < records_.size();) {
    size_t j = i + 1;
    while (j < records_.size() && records_[j]->ins_id == records_[i]->ins_id) ++j;
    const size_t g = j - i;
    bool keep = !(merge_size > 0 && (int)g != merge_size);
    if (keep && g > 1) {
      SlotRecord* m = records_[i];
      std::vector<std::vector<uint64_t>> u(n_used_u64_);
      std::vector<std::vector<float>> f(n_used_f_);
      for (size_t k = i; k < j && keep; ++k) {
        const SlotRecord* r = records_[k];
        for (int p = 0; p < n_used_u64_ && p + 1 < (int)r->u64_offs.size(); ++p) {
          const uint32_t a = r->u64_offs[p], e = r->u64_offs[p + 1];
          if (a == e) continue;
          if (!u[p].empty()) { keep = false; break; }          // conflicting sparse slot
          u[p].assign(r->u64_vals.begin() + a, r->u64_vals.begin() + e);
        }
        for (int p = 0; p < n_used_f_ && p + 1 < (int)r->f_offs.size(); ++p) {
          const uint32_t a = r->f_offs[p], e = r->f_offs[p + 1];
          if (a == e || !f[p].empty()) continue;
          bool nonzero = false;
          for (uint32_t q = a; q < e; ++q) nonzero |= r->f_vals[q] != 0.f;
          if (nonzero || k + 1 == j) f[p].assign(r->f_vals.begin() + a, r->f_vals.begin() + e);
        }
      }
      if (keep) {
        m->u64_vals.clear();
        m->u64_offs.assign(n_used_u64_ + 1, 0);
        for (int p = 0; p < n_used_u64_; ++p) {
          m->u64_vals.insert(m->u64_vals.end(), u[p].begin(), u[p].end());
          m->u64_offs[p + 1] = (uint32_t)m->u64_vals.size();
        }
        m->f_vals.clear();
        m->f_offs.assign(n_used_f_ + 1, 0);
        for (int p = 0; p < n_used_f_; ++p) {
          m->f_vals.insert(m->f_vals.end(), f[p].begin(), f[p].end());
          m->f_offs[p + 1] = (uint32_t)m->f_vals.size();
        }
        out.push_back(m);
        for (size_t k = i + 1; k < j; ++k) dropped.push_back(records_[k]);
      }
    } else if (keep) {
      out.push_back(records_[i]);
    }
    if (!keep) {
      n_drop += (int64_t)g;
      for (size_t k = i; k < j; ++k) dropped.push_back(records_[k]);
    }
    i = j;
  }
  records_.swap(out);
  pool_.put(dropped);
  return n_drop;
}

void Dataset::drop_arena() {
  arena_ok_ = false;
  if (map_base_) munmap(map_base_, map_size_);
  map_base_ = nullptr;
  map_size_ = 0;
  arena_only_ = false;
  arena_n_ = 0;
  pk_ = nullptr; pko_ = nullptr; pso_ = nullptr; pd_ = nullptr;
  std::vector<uint64_t>().swap(ar_keys_);
  std::vector<int64_t>().swap(ar_koff_);
  std::vector<uint32_t>().swap(ar_soff_);
  std::vector<float>().swap(ar_dense_);
}

void Dataset::build_arena() {
  drop_arena();
  const size_t n = records_.size();
  if (compact_batches == 0 || n == 0) return;
  const int S = n_used_u64_;
  ar_koff_.assign(n + 1, 0);
  for (size_t i = 0; i < n; ++i) ar_koff_[i + 1] = ar_koff_[i] + (int64_t)records_[i]->u64_vals.size();
  const int64_t bytes = ar_koff_[n] * 8 + (int64_t)n * ((S + 1) * 4 + dense_dim_ * 4);
  if (compact_batches == 1 && bytes > compact_limit_mb * (1ll << 20)) { drop_arena(); return; }
  ar_keys_.resize((size_t)ar_koff_[n]);
  ar_soff_.resize(n * (size_t)(S + 1));
  ar_dense_.assign(n * (size_t)dense_dim_, 0.f);
  ThreadPool pool(std::max(1, merge_thread_num));
  pool.parallel_range(n, [&](size_t b, size_t e, int) {
    for (size_t i = b; i < e; ++i) {
      const SlotRecord* rec = records_[i];
      if (!rec->u64_vals.empty())
        memcpy(ar_keys_.data() + ar_koff_[i], rec->u64_vals.data(), rec->u64_vals.size() * sizeof(uint64_t));
      uint32_t* so = ar_soff_.data() + i * (size_t)(S + 1);
      for (int s = 0; s <= S; ++s) so[s] = s < (int)rec->u64_offs.size() ? rec->u64_offs[s] : (uint32_t)rec->u64_vals.size();
      float* d = ar_dense_.data() + i * (size_t)dense_dim_;
      for (int f = 0; f < n_used_f_ && f + 1 < (int)rec->f_offs.size(); ++f) {
        const uint32_t a = rec->f_offs[f], en = rec->f_offs[f + 1];
        const int dim = float_dim_[f];
        for (int j = 0; j < dim; ++j) d[float_off_[f] + j] = (a + j < en) ? rec->f_vals[a + j] : 0.f;
      }
    }
  });
  pk_ = ar_keys_.data(); pko_ = ar_koff_.data(); pso_ = ar_soff_.data(); pd_ = ar_dense_.data();
  arena_n_ = (int64_t)n;
  arena_ok_ = true;
}

// ---- pass file ------------------------------------------------------------------------------------
namespace {
struct ArenaHeader {
  char magic[8];            // "PBXARENA"
  uint32_t version, S, dense_dim, reserved;
  int64_t n, nkeys, n_unique;
  int64_t off_koff, off_soff, off_dense, off_keys, off_unique, file_size;
};
inline int64_t align64(int64_t x) { return (x + 63) / 64 * 64; }
}  // namespace

int Dataset::save_arena(const std::string& path) {
  PBH_REQUIRE(!arena_only_, "this pass was loaded from a pass file already");
  if (!arena_ok_) {
    const int keep = compact_batches;
    compact_batches = 2;
    build_arena();
    compact_batches = keep;
  }
  PBH_REQUIRE(arena_ok_, "no records to save");
  ArenaHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "PBXARENA", 8);
  h.version = 1; h.S = (uint32_t)n_used_u64_; h.dense_dim = (uint32_t)dense_dim_;
  h.n = arena_n_; h.nkeys = pko_[arena_n_]; h.n_unique = (int64_t)unique_keys_.size();
  h.off_koff = align64(sizeof(h));
  h.off_soff = align64(h.off_koff + (h.n + 1) * 8);
  h.off_dense = align64(h.off_soff + h.n * (int64_t)(h.S + 1) * 4);
  h.off_keys = align64(h.off_dense + h.n * (int64_t)h.dense_dim * 4);
  h.off_unique = align64(h.off_keys + h.nkeys * 8);
  h.file_size = h.off_unique + h.n_unique * 8;
  FILE* fp = fopen(path.c_str(), "wb");
  PBH_REQUIRE(fp != nullptr, "cannot open %s for writing", path.c_str());
  bool ok = true;
  auto put = [&](int64_t off, const void* p, size_t bytes) {
    ok = ok && fseek(fp, (long)off, SEEK_SET) == 0 && (bytes == 0 || fwrite(p, 1, bytes, fp) == bytes);
  };
  put(0, &h, sizeof(h));
  put(h.off_koff, pko_, (size_t)(h.n + 1) * 8);
  put(h.off_soff, pso_, (size_t)h.n * (h.S + 1) * 4);
  put(h.off_dense, pd_, (size_t)h.n * h.dense_dim * 4);
  put(h.off_keys, pk_, (size_t)h.nkeys * 8);
  put(h.off_unique, unique_keys_.data(), (size_t)h.n_unique * 8);
  if (ok && h.n_unique == 0) ok = ftruncate(fileno(fp), (off_t)h.file_size) == 0;
  ok = (fclose(fp) == 0) && ok;
  PBH_REQUIRE(ok, "short write to %s", path.c_str());
  return 0;
}

int Dataset::load_arena(const std::string& path) {
  if (slot_to_sparse_.size() != slots.size()) finalize_slots();
  release_memory();
  int fd = open(path.c_str(), O_RDONLY);
  PBH_REQUIRE(fd >= 0, "cannot open %s", path.c_str());
  struct stat st;
  if (fstat(fd, &st) != 0 || (size_t)st.st_size < sizeof(ArenaHeader)) { close(fd); PBH_REQUIRE(false, "%s is not a pass file", path.c_str()); }
  void* base = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  PBH_REQUIRE(base != MAP_FAILED, "cannot map %s", path.c_str());
  const ArenaHeader& h = *(const ArenaHeader*)base;
  const bool good = memcmp(h.magic, "PBXARENA", 8) == 0 && h.version == 1 && h.file_size <= (int64_t)st.st_size &&
                    h.n >= 0 && h.nkeys >= 0 && h.off_unique + h.n_unique * 8 <= (int64_t)st.st_size;
  if (!good || (int)h.S != n_used_u64_ || (int)h.dense_dim != dense_dim_) {
    const uint32_t fs = good ? h.S : 0, fd_ = good ? h.dense_dim : 0;
    munmap(base, (size_t)st.st_size);
    PBH_REQUIRE(good, "%s is not a pass file (bad header)", path.c_str());
    PBH_REQUIRE(false, "pass file %s holds %u sparse slots / dense width %u, the dataset is set up for %d / %d",
                path.c_str(), fs, fd_, n_used_u64_, dense_dim_);
  }
  const char* b = (const char*)base;
  {
    // the header only says where things are: every section must lie inside the file and the offsets must be
    // what build_arena writes (a damaged file is refused here instead of being read out of bounds later)
    const int64_t fsz = (int64_t)st.st_size, S1 = (int64_t)h.S + 1;
    auto inside = [&](int64_t off, int64_t bytes) { return off >= (int64_t)sizeof(ArenaHeader) && bytes >= 0 && off <= fsz && bytes <= fsz - off; };
    bool ok = h.n <= (int64_t)1 << 40 && h.nkeys <= (int64_t)1 << 44 && h.n_unique <= (int64_t)1 << 44 &&
              inside(h.off_koff, (h.n + 1) * 8) && inside(h.off_soff, h.n * S1 * 4) &&
              inside(h.off_dense, h.n * (int64_t)h.dense_dim * 4) && inside(h.off_keys, h.nkeys * 8) &&
              inside(h.off_unique, h.n_unique * 8) && h.off_koff % 8 == 0 && h.off_soff % 4 == 0 &&
              h.off_dense % 4 == 0 && h.off_keys % 8 == 0 && h.off_unique % 8 == 0;
    if (ok) {
      const int64_t* ko = (const int64_t*)(b + h.off_koff);
      const uint32_t* so = (const uint32_t*)(b + h.off_soff);
      ok = ko[0] == 0 && ko[h.n] == h.nkeys;
      for (int64_t i = 0; ok && i < h.n; ++i) {
        const int64_t len = ko[i + 1] - ko[i];
        const uint32_t* si = so + i * S1;
        ok = len >= 0 && len <= 0x7fffffff && si[0] == 0 && (int64_t)si[h.S] == len;
        for (int64_t q = 0; ok && q < (int64_t)h.S; ++q) ok = si[q] <= si[q + 1];
      }
    }
    if (!ok) {
      munmap(base, (size_t)st.st_size);
      PBH_REQUIRE(false, "pass file %s is damaged (sections or offsets are inconsistent)", path.c_str());
    }
  }
  map_base_ = base;
  map_size_ = (size_t)st.st_size;
  pko_ = (const int64_t*)(b + h.off_koff);
  pso_ = (const uint32_t*)(b + h.off_soff);
  pd_ = (const float*)(b + h.off_dense);
  pk_ = (const uint64_t*)(b + h.off_keys);
  arena_n_ = h.n;
  arena_only_ = true;
  arena_ok_ = true;
  const uint64_t* uk = (const uint64_t*)(b + h.off_unique);
  unique_keys_.assign(uk, uk + h.n_unique);
  total_feasigns_ = h.nkeys;
  lines_read_ = h.n;
  madvise(base, map_size_, MADV_WILLNEED);
  return 0;
}

// multi-node equalisation: every rank must step the same number of times; the caller passes
// allreduce(max) of the local count and the largest batches are split until the count is reached
// (data_set.cc:2708-2750)
static void equalise_batches(std::vector<std::vector<BatchRange>>* worker_batches, int64_t min_batches) {
  if (min_batches <= 0) return;
  for (auto& wb : *worker_batches) {
    while ((int64_t)wb.size() < min_batches && !wb.empty()) {
      size_t big = 0;
      for (size_t i = 1; i < wb.size(); ++i) if (wb[i].len > wb[big].len) big = i;
      if (wb[big].len < 2) break;
      BatchRange a = wb[big];
      int h = a.len / 2;
      wb[big] = {a.start, h};
      wb.push_back({a.start + h, a.len - h});
    }
  }
}

int Dataset::prepare_train(int nworkers, bool shuffle, uint64_t seed, int64_t min_batches) {
  PBH_REQUIRE(nworkers >= 1, "nworkers must be >= 1");
  if (arena_only_) {
    // the record order of a pass file is fixed (it was shuffled before it was saved): a shuffle request
    // permutes the batches instead
    worker_batches_.assign(nworkers, {});
    if (arena_n_ < nworkers) return 0;
    std::vector<BatchRange> all;
    compute_batch_ranges(arena_n_, batch_size, nworkers, &all);
    if (shuffle) {
      // whole rounds of nworkers batches move together so that every worker keeps equal-sized batches
      const size_t rounds = all.size() / (size_t)nworkers;
      std::vector<size_t> perm(rounds);
      for (size_t i = 0; i < rounds; ++i) perm[i] = i;
      std::mt19937_64 rng(seed);
      std::shuffle(perm.begin(), perm.end(), rng);
      std::vector<BatchRange> mixed;
      for (size_t r : perm) mixed.insert(mixed.end(), all.begin() + r * nworkers, all.begin() + (r + 1) * nworkers);
      mixed.insert(mixed.end(), all.begin() + rounds * nworkers, all.end());
      all.swap(mixed);
    }
    for (size_t i = 0; i < all.size(); ++i) worker_batches_[i % nworkers].push_back(all[i]);
    equalise_batches(&worker_batches_, min_batches);
    return 0;
  }
  if (enable_pv_merge) {
    // page-view merge: instances of one search_id stay adjacent, ordered by rank (data_set.cc:2648-2688)
    std::stable_sort(records_.begin(), records_.end(), [](const SlotRecord* a, const SlotRecord* b) {
      return a->search_id != b->search_id ? a->search_id < b->search_id : a->rank < b->rank;
    });
  } else if (shuffle) {
    local_shuffle(seed);
  }
  worker_batches_.assign(nworkers, {});
  const int64_t n = (int64_t)records_.size();
  if (n < nworkers) return 0;  // fewer instances than workers -> zero batches (data_set.cc:2696-2701)
  std::vector<BatchRange> all;
  compute_batch_ranges(n, batch_size, nworkers, &all);
  for (size_t i = 0; i < all.size(); ++i) worker_batches_[i % nworkers].push_back(all[i]);
  equalise_batches(&worker_batches_, min_batches);
  build_arena();
  return 0;
}

int Dataset::batch_shape(int worker, int64_t b, int* B, int64_t* n_keys) const {
  PBH_REQUIRE(worker >= 0 && worker < (int)worker_batches_.size() && b >= 0 &&
                  b < (int64_t)worker_batches_[worker].size(), "batch (%d, %lld) out of range", worker, (long long)b);
  const BatchRange& r = worker_batches_[worker][b];
  int64_t nk = 0;
  if (arena_ok_) nk = pko_[r.start + r.len] - pko_[r.start];
  else for (int i = 0; i < r.len; ++i) nk += (int64_t)records_[r.start + i]->u64_vals.size();
  *B = r.len;
  *n_keys = nk;
  return 0;
}

int Dataset::pack_batch(int worker, int64_t b, uint64_t* keys, int32_t* off, float* dense, int slot_major) const {
  PBH_REQUIRE(worker >= 0 && worker < (int)worker_batches_.size() && b >= 0 &&
                  b < (int64_t)worker_batches_[worker].size(), "batch (%d, %lld) out of range", worker, (long long)b);
  const BatchRange& r = worker_batches_[worker][b];
  const int B = r.len, S = n_used_u64_;
  if (arena_ok_ && !slot_major) {
    // compact arena: three contiguous copies
    const int64_t k0 = pko_[r.start];
    memcpy(keys, pk_ + k0, (size_t)(pko_[r.start + B] - k0) * sizeof(uint64_t));
    const uint32_t* so = pso_ + (size_t)r.start * (S + 1);
    for (int i = 0; i < B; ++i) {
      const int32_t base = (int32_t)(pko_[r.start + i] - k0);
      int32_t* o = off + (int64_t)i * (S + 1);
      const uint32_t* si = so + (size_t)i * (S + 1);
      for (int s = 0; s <= S; ++s) o[s] = base + (int32_t)si[s];
    }
    if (dense && dense_dim_ > 0)
      memcpy(dense, pd_ + (size_t)r.start * dense_dim_, (size_t)B * dense_dim_ * sizeof(float));
    return 0;
  }
  if (arena_only_) {
    // slot-major straight from the mapped pass file
    int64_t cur = 0;
    for (int s = 0; s < S; ++s) {
      for (int i = 0; i < B; ++i) {
        const uint32_t* si = pso_ + (size_t)(r.start + i) * (S + 1);
        off[(int64_t)s * B + i] = (int32_t)cur;
        memcpy(keys + cur, pk_ + pko_[r.start + i] + si[s], (size_t)(si[s + 1] - si[s]) * sizeof(uint64_t));
        cur += si[s + 1] - si[s];
      }
    }
    off[(int64_t)S * B] = (int32_t)cur;
    if (dense && dense_dim_ > 0)
      memcpy(dense, pd_ + (size_t)r.start * dense_dim_, (size_t)B * dense_dim_ * sizeof(float));
    return 0;
  }
  if (!slot_major) {
    int64_t cur = 0;
    for (int i = 0; i < B; ++i) {
      const SlotRecord* rec = records_[r.start + i];
      memcpy(keys + cur, rec->u64_vals.data(), rec->u64_vals.size() * sizeof(uint64_t));
      int32_t* o = off + (int64_t)i * (S + 1);
      for (int s = 0; s <= S; ++s) o[s] = (int32_t)(cur + rec->u64_offs[s]);
      cur += (int64_t)rec->u64_vals.size();
    }
  } else {
    int64_t cur = 0;
    for (int s = 0; s < S; ++s) {
      for (int i = 0; i < B; ++i) {
        const SlotRecord* rec = records_[r.start + i];
        off[(int64_t)s * B + i] = (int32_t)cur;
        uint32_t a = rec->u64_offs[s], e = rec->u64_offs[s + 1];
        memcpy(keys + cur, rec->u64_vals.data() + a, (size_t)(e - a) * sizeof(uint64_t));
        cur += e - a;
      }
    }
    off[(int64_t)S * B] = (int32_t)cur;
  }
  if (dense && dense_dim_ > 0) {
    for (int i = 0; i < B; ++i) {
      const SlotRecord* rec = records_[r.start + i];
      float* d = dense + (int64_t)i * dense_dim_;
      for (int f = 0; f < n_used_f_; ++f) {
        uint32_t a = rec->f_offs[f], e = rec->f_offs[f + 1];
        int dim = float_dim_[f];
        for (int j = 0; j < dim; ++j) d[float_off_[f] + j] = (a + j < e) ? rec->f_vals[a + j] : 0.f;
      }
    }
  }
  return 0;
}

int Dataset::batch_ins_ids(int worker, int64_t b, std::string* joined) const {
  PBH_REQUIRE(worker >= 0 && worker < (int)worker_batches_.size() && b >= 0 &&
                  b < (int64_t)worker_batches_[worker].size(), "batch out of range");
  const BatchRange& r = worker_batches_[worker][b];
  joined->clear();
  PBH_REQUIRE(!arena_only_, "a pass file holds no instance ids");
  for (int i = 0; i < r.len; ++i) {
    if (i) joined->push_back('\n');
    joined->append(records_[r.start + i]->ins_id);
  }
  return 0;
}

int Dataset::batch_meta(int worker, int64_t b, uint64_t* search_ids, uint32_t* ranks, uint32_t* cmatch) const {
  PBH_REQUIRE(worker >= 0 && worker < (int)worker_batches_.size() && b >= 0 &&
                  b < (int64_t)worker_batches_[worker].size(), "batch out of range");
  const BatchRange& r = worker_batches_[worker][b];
  PBH_REQUIRE(!arena_only_, "a pass file holds no search_id / rank / cmatch");
  for (int i = 0; i < r.len; ++i) {
    const SlotRecord* rec = records_[r.start + i];
    if (search_ids) search_ids[i] = rec->search_id;
    if (ranks) ranks[i] = rec->rank;
    if (cmatch) cmatch[i] = rec->cmatch;
  }
  return 0;
}

// ---- binary archive: [magic][n_records] then per record
//   search_id u64 | rank u32 | cmatch u32 | ins_len u32 | ins bytes |
//   n_u64 u32 | u64 vals | n_offs u16 | offs u32[] | n_f u32 | f vals | n_foffs u16 | foffs u32[]
static const uint64_t kArchiveMagic = 0x50425841524348ull;  // "PBXARCH"

template <typename T>
static void put(std::string& buf, const T& v) { buf.append((const char*)&v, sizeof(T)); }

static int write_archive(const std::string& path, const std::vector<const SlotRecord*>& recs) {
  FILE* fp = fopen(path.c_str(), "wb");
  PBH_REQUIRE(fp != nullptr, "cannot open %s for writing", path.c_str());
  std::string buf;
  put(buf, kArchiveMagic);
  put(buf, (uint64_t)recs.size());
  for (const SlotRecord* r : recs) {
    put(buf, r->search_id); put(buf, r->rank); put(buf, r->cmatch);
    put(buf, (uint32_t)r->ins_id.size()); buf.append(r->ins_id);
    put(buf, (uint32_t)r->u64_vals.size());
    buf.append((const char*)r->u64_vals.data(), r->u64_vals.size() * 8);
    put(buf, (uint16_t)r->u64_offs.size());
    buf.append((const char*)r->u64_offs.data(), r->u64_offs.size() * 4);
    put(buf, (uint32_t)r->f_vals.size());
    buf.append((const char*)r->f_vals.data(), r->f_vals.size() * 4);
    put(buf, (uint16_t)r->f_offs.size());
    buf.append((const char*)r->f_offs.data(), r->f_offs.size() * 4);
    if (buf.size() > (8u << 20)) { fwrite(buf.data(), 1, buf.size(), fp); buf.clear(); }
  }
  fwrite(buf.data(), 1, buf.size(), fp);
  fclose(fp);
  return 0;
}

int Dataset::save_archive(const std::string& path) const {
  PBH_REQUIRE(!arena_only_, "save_archive needs records: not available for a pass loaded from a pass file");
  std::vector<const SlotRecord*> all(records_.begin(), records_.end());
  return write_archive(path, all);
}

// Global shuffle, sending side: split the records into `world` partition files
// "<prefix>.<this rank>.<dst>" (the reference hash-partitions records to ranks and ships them
// over the PaddleShuffler RPC: by search_id, by XXH64 of the ins_id, or randomly;
// data_set.cc:2422-2547).  mode 0 = random, 1 = search_id, 2 = hash of ins_id.
int Dataset::save_partitioned(const std::string& prefix, int nparts, int mode, uint64_t seed) const {
  PBH_REQUIRE(nparts >= 1, "nparts must be >= 1");
  PBH_REQUIRE(!arena_only_, "global shuffle needs records: not available for a pass loaded from a pass file");
  std::vector<std::vector<const SlotRecord*>> parts(nparts);
  std::mt19937_64 rng(seed + (uint64_t)rank * 0x9E3779B97F4A7C15ull);
  for (const SlotRecord* r : records_) {
    uint64_t h;
    if (mode == 1) h = mix64(r->search_id);
    else if (mode == 2) h = mix64(std::hash<std::string>()(r->ins_id.substr(0, 32)));
    else h = rng();
    parts[h % (uint64_t)nparts].push_back(r);
  }
  for (int d = 0; d < nparts; ++d) {
    int rc = write_archive(prefix + "." + std::to_string(rank) + "." + std::to_string(d), parts[d]);
    if (rc) return rc;
  }
  return 0;
}

int Dataset::load_archive(const std::string& path) {
  drop_arena();
  // The file is mapped, record boundaries are found in one cheap sequential scan of the length
  // fields, and the records are decoded in parallel (one fread per field used to dominate).
  int fd = open(path.c_str(), O_RDONLY);
  PBH_REQUIRE(fd >= 0, "cannot open %s", path.c_str());
  struct stat st;
  if (fstat(fd, &st) != 0 || st.st_size < 16) { close(fd); PBH_REQUIRE(false, "archive %s is truncated", path.c_str()); }
  const size_t fsz = (size_t)st.st_size;
  const uint8_t* base = (const uint8_t*)mmap(nullptr, fsz, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  PBH_REQUIRE(base != MAP_FAILED, "cannot map %s", path.c_str());
  uint64_t magic = 0, n = 0;
  memcpy(&magic, base, 8);
  memcpy(&n, base + 8, 8);
  bool ok = magic == kArchiveMagic;
  std::vector<size_t> start;
  if (ok) {
    start.reserve((size_t)n + 1);
    size_t p = 16;
    auto need = [&](size_t k) { return p + k <= fsz; };
    for (uint64_t i = 0; ok && i < n; ++i) {
      start.push_back(p);
      uint32_t len = 0, cnt = 0;
      uint16_t no = 0;
      ok = need(20);                                   // search_id, rank, cmatch, ins_id length
      if (!ok) break;
      memcpy(&len, base + p + 16, 4); p += 20;
      ok = need((size_t)len + 4); if (!ok) break;
      p += len; memcpy(&cnt, base + p, 4); p += 4;
      ok = need((size_t)cnt * 8 + 2); if (!ok) break;
      p += (size_t)cnt * 8; memcpy(&no, base + p, 2); p += 2;
      ok = need((size_t)no * 4 + 4); if (!ok) break;
      p += (size_t)no * 4; memcpy(&cnt, base + p, 4); p += 4;
      ok = need((size_t)cnt * 4 + 2); if (!ok) break;
      p += (size_t)cnt * 4; memcpy(&no, base + p, 2); p += 2;
      ok = need((size_t)no * 4); if (!ok) break;
      p += (size_t)no * 4;
    }
  }
  if (!ok) { munmap((void*)base, fsz); PBH_REQUIRE(false, "archive %s is truncated or not a paddlebox_b200 archive", path.c_str()); }
  if (slot_to_sparse_.size() != slots.size()) finalize_slots();
  agent_.clear();
  if (collect_keys) agent_.add_key(0);
  const size_t first = records_.size();
  records_.resize(first + (size_t)n, nullptr);
  std::atomic<int64_t> feas{0};
  std::atomic<bool> corrupt{false};
  ThreadPool pool(std::max(1, merge_thread_num));
  pool.parallel_range((size_t)n, [&](size_t b0, size_t e0, int) {
    int64_t local_feas = 0;
    for (size_t i = b0; i < e0; ++i) {
      const uint8_t* q = base + start[i];
      SlotRecord* r = pool_.get();
      uint32_t len = 0, cnt = 0;
      uint16_t no = 0;
      memcpy(&r->search_id, q, 8); memcpy(&r->rank, q + 8, 4); memcpy(&r->cmatch, q + 12, 4);
      memcpy(&len, q + 16, 4); q += 20;
      r->ins_id.assign((const char*)q, len); q += len;
      memcpy(&cnt, q, 4); q += 4;
      r->u64_vals.resize(cnt); if (cnt) memcpy(r->u64_vals.data(), q, (size_t)cnt * 8); q += (size_t)cnt * 8;
      memcpy(&no, q, 2); q += 2;
      r->u64_offs.resize(no); if (no) memcpy(r->u64_offs.data(), q, (size_t)no * 4); q += (size_t)no * 4;
      memcpy(&cnt, q, 4); q += 4;
      r->f_vals.resize(cnt); if (cnt) memcpy(r->f_vals.data(), q, (size_t)cnt * 4); q += (size_t)cnt * 4;
      memcpy(&no, q, 2); q += 2;
      r->f_offs.resize(no); if (no) memcpy(r->f_offs.data(), q, (size_t)no * 4);
      records_[first + i] = r;
      // the length fields were consistent with the file size; the offset tables must also be consistent
      // with the values they index and with this dataset's slots (a flipped bit in an offset would
      // otherwise send build_arena / the packers out of bounds)
      auto offsets_ok = [](const std::vector<uint32_t>& offs, size_t n_vals, int n_slots) {
        if ((int)offs.size() != n_slots + 1 || offs.front() != 0 || offs.back() != n_vals) return false;
        for (size_t k = 1; k < offs.size(); ++k) if (offs[k] < offs[k - 1]) return false;
        return true;
      };
      if (!offsets_ok(r->u64_offs, r->u64_vals.size(), n_used_u64_) ||
          !offsets_ok(r->f_offs, r->f_vals.size(), n_used_f_)) { corrupt.store(true); continue; }
      if (collect_keys) collect_record_keys(r);
      local_feas += (int64_t)r->u64_vals.size();
    }
    if (collect_keys) agent_.flush_local();
    feas += local_feas;
  });
  munmap((void*)base, fsz);
  if (corrupt.load()) {
    // refuse the whole file: give its records back and leave the dataset as it was before the call
    std::vector<SlotRecord*> back(records_.begin() + (long)first, records_.end());
    back.erase(std::remove(back.begin(), back.end(), nullptr), back.end());
    records_.resize(first);
    pool_.put(back);
    agent_.clear();
    PBH_REQUIRE(false, "archive %s is damaged (slot offsets do not match its values or this dataset's slots)", path.c_str());
  }
  total_feasigns_ += feas.load();
  if (collect_keys) {
    // several archives may be loaded into one pass (global shuffle): keep the union of their keys
    agent_.add_keys(unique_keys_.data(), unique_keys_.size());
    agent_.flush_local();
    agent_.finalize(&pool, &unique_keys_);
    agent_.clear();
  }
  return 0;
}

std::string Dataset::last_stats() const {
  char buf[512];
  snprintf(buf, sizeof(buf),
           "records=%zu lines=%lld bad_lines=%lld feasigns=%lld unique_keys=%zu read_s=%.6g key_merge_s=%.6g pool_idle=0",
           records_.size(), (long long)lines_read_.load(), (long long)lines_bad_.load(),
           (long long)total_feasigns_.load(), unique_keys_.size(), (double)load_seconds_, (double)merge_seconds_);
  return std::string(buf);
}

}  // namespace pbh
