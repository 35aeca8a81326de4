// C ABI of the host runtime (see include/pbx_host.h): dataset / data feed, async dense
// table.  Host table and metrics export their own entry points in their .cc files.
#include <algorithm>
#include <cmath>
#include <sstream>

#include "dataset.h"

namespace pbh {
const char* last_error();

// Async dense mode: a host-resident flat parameter vector with Adam moments.  Workers
// pull the whole vector before a batch and push their gradient vector after it; one
// background thread merges up to `merge_limit` queued gradients and applies Adam on
// `threads` CPU threads.  Hard-coded hyper-parameters follow the reference
// (BoxPSAsynDenseTable, framework/boxps_worker.cc:61-370; beta1 0.99, beta2 0.9999,
// eps 1e-8 at :287-294).
class AsyncDenseTable {
 public:
  AsyncDenseTable(int64_t n, float lr, int threads, int merge_limit)
      : n_(n), lr_(lr), pool_(threads), merge_limit_(merge_limit), w_(n, 0.f), m_(n, 0.f), v_(n, 0.f),
        queue_(64) {
    worker_ = std::thread([this] { loop(); });
  }
  ~AsyncDenseTable() { stop(); }
  void stop() {
    if (worker_.joinable()) {
      queue_.close();
      worker_.join();
    }
  }
  void set_params(const float* p) { std::lock_guard<std::mutex> g(mu_); std::copy(p, p + n_, w_.begin()); }
  void pull(float* out) { std::lock_guard<std::mutex> g(mu_); std::copy(w_.begin(), w_.end(), out); }
  void push(const float* grad) {
    std::vector<float> g(grad, grad + n_);
    pending_++;
    queue_.put(std::move(g));
  }
  void wait_idle() {
    while (pending_.load() > 0) std::this_thread::sleep_for(std::chrono::microseconds(200));
  }
  int64_t updates() const { return updates_.load(); }
  float lr_scale = 1.f;
  // per-parameter learning rates (lr_map of InitializeGPUAndLoadModel, boxps_worker.cc:199-207)
  void set_lr_range(int64_t begin, int64_t end, float lr) {
    std::lock_guard<std::mutex> g(mu_);
    if (lr_vec_.empty()) lr_vec_.assign((size_t)n_, lr_);
    for (int64_t i = std::max<int64_t>(0, begin); i < std::min(end, n_); ++i) lr_vec_[i] = lr;
  }

 private:
  void loop() {
    std::vector<std::vector<float>> batch;
    for (;;) {
      batch.clear();
      if (queue_.read(batch, (size_t)merge_limit_) == 0) {
        if (queue_.closed()) return;
        continue;
      }
      // merge the queued gradients (mean), then one Adam step
      const float inv = 1.f / (float)batch.size();
      std::vector<float>& acc = batch[0];
      for (size_t k = 1; k < batch.size(); ++k) {
        const std::vector<float>& g = batch[k];
        pool_.parallel_range((size_t)n_, [&](size_t b, size_t e, int) {
          for (size_t i = b; i < e; ++i) acc[i] += g[i];
        });
      }
      step_ += 1;
      const float b1 = 0.99f, b2 = 0.9999f, eps = 1e-8f;
      b1pow_ *= b1;
      b2pow_ *= b2;
      const float corr = lr_scale * std::sqrt(1.f - b2pow_) / (1.f - b1pow_);
      const float lr = lr_ * corr;
      {
        std::lock_guard<std::mutex> g(mu_);
        const float* lrv = lr_vec_.empty() ? nullptr : lr_vec_.data();
        pool_.parallel_range((size_t)n_, [&](size_t b, size_t e, int) {
          for (size_t i = b; i < e; ++i) {
            float gi = acc[i] * inv;
            m_[i] = b1 * m_[i] + (1.f - b1) * gi;
            v_[i] = b2 * v_[i] + (1.f - b2) * gi * gi;
            w_[i] -= (lrv ? lrv[i] * corr : lr) * m_[i] / (std::sqrt(v_[i]) + eps);
          }
        });
      }
      updates_ += 1;
      pending_ -= (int)batch.size();
    }
  }
  int64_t n_;
  float lr_;
  ThreadPool pool_;
  int merge_limit_;
  std::vector<float> w_, m_, v_, lr_vec_;
  Channel<std::vector<float>> queue_;
  std::thread worker_;
  std::mutex mu_;
  std::atomic<int> pending_{0};
  std::atomic<int64_t> updates_{0};
  int64_t step_ = 0;
  float b1pow_ = 1.f, b2pow_ = 1.f;
};

}  // namespace pbh

using pbh::Dataset;
using pbh::AsyncDenseTable;

extern "C" {

const char* pbh_last_error() { return pbh::last_error(); }

// ---------------- dataset ----------------
int pbh_dataset_create(void** out) { *out = new Dataset(); return 0; }
int pbh_dataset_destroy(void* h) { delete (Dataset*)h; return 0; }

// slots: name/type/is_dense/is_used/dim arrays of length n (type 0 = uint64, 1 = float)
int pbh_dataset_set_slots(void* h, int n, const char* const* names, const int* types, const int* is_dense,
                          const int* is_used, const int* dims) {
  Dataset* d = (Dataset*)h;
  d->slots.clear();
  for (int i = 0; i < n; ++i) {
    pbh::SlotDesc s;
    s.name = names[i]; s.type = types[i]; s.is_dense = is_dense[i]; s.is_used = is_used[i]; s.dim = dims[i];
    d->slots.push_back(s);
  }
  return d->finalize_slots();
}
int pbh_dataset_set_filelist(void* h, int n, const char* const* files) {
  Dataset* d = (Dataset*)h;
  d->files.assign(files, files + n);
  return 0;
}
// keys: thread_num, merge_thread_num, batch_size, parse_ins_id, parse_logkey, enable_pv_merge,
//       collect_keys, rank, world
int pbh_dataset_set_int(void* h, const char* key, int64_t v) {
  Dataset* d = (Dataset*)h;
  std::string k(key);
  if (k == "thread_num") d->thread_num = (int)v;
  else if (k == "merge_thread_num") d->merge_thread_num = (int)v;
  else if (k == "batch_size") d->batch_size = (int)v;
  else if (k == "parse_ins_id") d->parse_ins_id = v != 0;
  else if (k == "parse_logkey") d->parse_logkey = v != 0;
  else if (k == "enable_pv_merge") d->enable_pv_merge = v != 0;
  else if (k == "collect_keys") d->collect_keys = v != 0;
  else if (k == "record_reset_shrink") d->set_record_reset_shrink(v != 0);
  else if (k == "compact_batches") d->compact_batches = (int)v;
  else if (k == "compact_limit_mb") d->compact_limit_mb = (int64_t)v;
  else if (k == "parser_file_mode") d->parser_file_mode = v != 0;
  else if (k == "parser_file_mode") d->parser_file_mode = v != 0;
  else if (k == "rank") d->rank = (int)v;
  else if (k == "world") d->world = (int)v;
  else { pbh::set_error("unknown dataset option %s", key); return -1; }
  return 0;
}
int pbh_dataset_set_float(void* h, const char* key, double v) {
  Dataset* d = (Dataset*)h;
  if (std::string(key) == "sample_rate") { d->sample_rate = (float)v; return 0; }
  pbh::set_error("unknown dataset option %s", key);
  return -1;
}
int pbh_dataset_set_str(void* h, const char* key, const char* v) {
  Dataset* d = (Dataset*)h;
  std::string k(key);
  if (k == "pipe_command") d->pipe_command = v;
  else if (k == "parser_plugin") d->parser_plugin = v;
  else { pbh::set_error("unknown dataset option %s", key); return -1; }
  return 0;
}
int pbh_dataset_save_arena(void* h, const char* path) { return ((Dataset*)h)->save_arena(path); }
int pbh_dataset_load_arena(void* h, const char* path) { return ((Dataset*)h)->load_arena(path); }
int pbh_dataset_set_reader_cores(void* h, const int* cores, int n) {
  ((Dataset*)h)->reader_cores.assign(cores, cores + n);
  return 0;
}
int pbh_dataset_set_omit_slots(void* h, const int* slot_indices, int n) {
  ((Dataset*)h)->omit_slots.assign(slot_indices, slot_indices + n);
  return 0;
}
int pbh_dataset_set_input_index(void* h, void* index) {
  Dataset* d = (Dataset*)h;
  d->input_index = index;           // the plugin's offset callback is bound when loading starts
  return 0;
}
// the feed configuration of `src` (slots, parser, threads, flags, plugin, omitted slots, reader cores) onto `dst`;
// the file list and the rank / world of the file dealing stay dst's own (streaming readers deal files themselves)
int pbh_dataset_copy_config(void* dst_h, void* src_h) {
  Dataset* d = (Dataset*)dst_h;
  const Dataset* s = (const Dataset*)src_h;
  d->slots = s->slots;
  d->pipe_command = s->pipe_command;
  d->parser_plugin = s->parser_plugin;
  d->thread_num = s->thread_num;
  d->merge_thread_num = s->merge_thread_num;
  d->batch_size = s->batch_size;
  d->parse_ins_id = s->parse_ins_id;
  d->parse_logkey = s->parse_logkey;
  d->sample_rate = s->sample_rate;
  d->enable_pv_merge = s->enable_pv_merge;
  d->collect_keys = s->collect_keys;
  d->compact_batches = s->compact_batches;
  d->compact_limit_mb = s->compact_limit_mb;
  d->input_index = s->input_index;
  d->reader_cores = s->reader_cores;
  d->omit_slots = s->omit_slots;
  d->parser_file_mode = s->parser_file_mode;
  return d->slots.empty() ? 0 : d->finalize_slots();
}
int pbh_dataset_load_into_memory(void* h) { return ((Dataset*)h)->load_into_memory(); }
int pbh_dataset_preload_into_memory(void* h) { return ((Dataset*)h)->preload_into_memory(); }
int pbh_dataset_wait_preload_done(void* h) { return ((Dataset*)h)->wait_preload_done(); }
int pbh_dataset_release_memory(void* h) { ((Dataset*)h)->release_memory(); return 0; }
// record pool: records freed by release_memory are kept for the next pass up to a cap
// (FLAGS_padbox_record_pool_max_size); trim returns them to the allocator now
int64_t pbh_dataset_trim_record_pool(void* h) { return (int64_t)((Dataset*)h)->trim_record_pool(); }
int64_t pbh_dataset_idle_records(void* h) { return (int64_t)((Dataset*)h)->idle_records(); }
int pbh_dataset_set_record_pool_cap(void* h, int64_t cap) { ((Dataset*)h)->set_record_pool_cap((size_t)(cap < 0 ? 0 : cap)); return 0; }
int pbh_dataset_local_shuffle(void* h, uint64_t seed) { ((Dataset*)h)->local_shuffle(seed); return 0; }
int64_t pbh_dataset_merge_by_ins_id(void* h, int merge_size) { return ((Dataset*)h)->merge_by_ins_id(merge_size); }
int pbh_dataset_slots_shuffle(void* h, const int* slots, int n, uint64_t seed) {
  return ((Dataset*)h)->slots_shuffle(std::vector<int>(slots, slots + n), seed);
}
int pbh_dataset_slots_shuffle_back(void* h, const int* slots, int n, uint64_t seed) {
  return ((Dataset*)h)->slots_shuffle(std::vector<int>(slots, slots + n), seed, true);
}
int pbh_dataset_slot_first_values(void* h, int slot, uint64_t* out, uint8_t* has) {
  return ((Dataset*)h)->slot_first_values(slot, out, has);
}
int pbh_dataset_expand_records(void* h, int slot, int label_slot, const int64_t* counts, const uint64_t* ids, const uint8_t* neg) {
  return ((Dataset*)h)->expand_records(slot, label_slot, counts, ids, neg);
}
int pbh_dataset_prepare_train(void* h, int nworkers, int shuffle, uint64_t seed, int64_t min_batches) {
  return ((Dataset*)h)->prepare_train(nworkers, shuffle != 0, seed, min_batches);
}
int64_t pbh_dataset_num_records(void* h) { return ((Dataset*)h)->num_records(); }
int64_t pbh_dataset_num_batches(void* h, int worker) { return ((Dataset*)h)->num_batches(worker); }
int pbh_dataset_n_sparse(void* h) { return ((Dataset*)h)->n_sparse(); }
int pbh_dataset_dense_dim(void* h) { return ((Dataset*)h)->dense_dim(); }
int64_t pbh_dataset_total_feasigns(void* h) { return ((Dataset*)h)->total_feasigns(); }
int pbh_dataset_batch_shape(void* h, int worker, int64_t b, int* B, int64_t* n_keys) {
  return ((Dataset*)h)->batch_shape(worker, b, B, n_keys);
}
int pbh_dataset_pack_batch(void* h, int worker, int64_t b, uint64_t* keys, int32_t* off, float* dense,
                           int slot_major) {
  return ((Dataset*)h)->pack_batch(worker, b, keys, off, dense, slot_major);
}
int pbh_dataset_batch_meta(void* h, int worker, int64_t b, uint64_t* search_ids, uint32_t* ranks, uint32_t* cmatch) {
  return ((Dataset*)h)->batch_meta(worker, b, search_ids, ranks, cmatch);
}
// writes '\n'-joined ins ids into buf (up to cap bytes incl. NUL); returns needed size
int64_t pbh_dataset_batch_ins_ids(void* h, int worker, int64_t b, char* buf, int64_t cap) {
  std::string s;
  if (((Dataset*)h)->batch_ins_ids(worker, b, &s) != 0) return -1;
  if (buf && cap > 0) {
    size_t n = std::min<size_t>((size_t)cap - 1, s.size());
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return (int64_t)s.size() + 1;
}
int64_t pbh_dataset_unique_keys(void* h, uint64_t* out, int64_t max_n) {
  const auto& ks = ((Dataset*)h)->unique_keys();
  int64_t n = std::min<int64_t>(max_n, (int64_t)ks.size());
  if (out && n > 0) memcpy(out, ks.data(), (size_t)n * 8);
  return (int64_t)ks.size();
}
int pbh_dataset_save_archive(void* h, const char* path) { return ((Dataset*)h)->save_archive(path); }
int pbh_dataset_load_archive(void* h, const char* path) { return ((Dataset*)h)->load_archive(path); }
int pbh_dataset_save_partitioned(void* h, const char* prefix, int nparts, int mode, uint64_t seed) {
  return ((Dataset*)h)->save_partitioned(prefix, nparts, mode, seed);
}
int64_t pbh_dataset_stats(void* h, char* buf, int64_t cap) {
  std::string s = ((Dataset*)h)->last_stats();
  if (buf && cap > 0) {
    size_t n = std::min<size_t>((size_t)cap - 1, s.size());
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return (int64_t)s.size() + 1;
}
// parse one line with the dataset's slot config (unit tests / debugging)
int pbh_dataset_parse_line(void* h, const char* line, int64_t* n_u64, int64_t* n_f32, uint64_t* u64_out,
                           int64_t u64_cap, float* f32_out, int64_t f32_cap, uint64_t* search_id) {
  pbh::SlotRecord rec;
  if (!((Dataset*)h)->parse_line(line, strlen(line), &rec)) return 1;
  *n_u64 = (int64_t)rec.u64_vals.size();
  *n_f32 = (int64_t)rec.f_vals.size();
  for (int64_t i = 0; i < *n_u64 && i < u64_cap; ++i) u64_out[i] = rec.u64_vals[i];
  for (int64_t i = 0; i < *n_f32 && i < f32_cap; ++i) f32_out[i] = rec.f_vals[i];
  if (search_id) *search_id = rec.search_id;
  return 0;
}
// reference compute_batch_num exposed for tests: writes (start, len) pairs
int64_t pbh_compute_batch_ranges(int64_t ins_num, int batch_size, int nworkers, int64_t* starts, int* lens,
                                 int64_t cap) {
  std::vector<pbh::BatchRange> r;
  pbh::compute_batch_ranges(ins_num, batch_size, nworkers, &r);
  for (size_t i = 0; i < r.size() && (int64_t)i < cap; ++i) { starts[i] = r[i].start; lens[i] = r[i].len; }
  return (int64_t)r.size();
}

// ---------------- async dense table ----------------
int pbh_async_dense_create(void** out, int64_t n, float lr, int threads, int merge_limit) {
  *out = new AsyncDenseTable(n, lr, threads < 1 ? 1 : threads, merge_limit < 1 ? 1 : merge_limit);
  return 0;
}
int pbh_async_dense_destroy(void* h) { delete (AsyncDenseTable*)h; return 0; }
int pbh_async_dense_set_params(void* h, const float* p) { ((AsyncDenseTable*)h)->set_params(p); return 0; }
int pbh_async_dense_pull(void* h, float* out) { ((AsyncDenseTable*)h)->pull(out); return 0; }
int pbh_async_dense_push(void* h, const float* g) { ((AsyncDenseTable*)h)->push(g); return 0; }
int pbh_async_dense_set_lr_range(void* h, int64_t begin, int64_t end, float lr) {
  ((AsyncDenseTable*)h)->set_lr_range(begin, end, lr);
  return 0;
}
int pbh_async_dense_wait(void* h) { ((AsyncDenseTable*)h)->wait_idle(); return 0; }
int64_t pbh_async_dense_updates(void* h) { return ((AsyncDenseTable*)h)->updates(); }

}  // extern "C"
