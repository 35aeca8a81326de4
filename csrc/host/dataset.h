// Pass-scoped in-memory SlotRecord dataset + data feed (host C++).
// Reference roles: PadBoxSlotDataset / SlotRecordDataset (framework/data_set.cc:1905-2828,
// :1853-1889), SlotPaddleBoxDataFeed (framework/data_feed.cc:2993-4153), SlotRecord /
// SlotObjPool (framework/data_feed.h:96-428), BufferedLineFileReader (data_feed.cc:54-158),
// PSAgent key collection (fleet/box_wrapper.cc:1185-1196, data_set.cc:2293-2349),
// BinaryArchive (data_feed.cc:2824-2992), compute_batch_num (data_set.cc:280-331).
#pragma once
#include <memory>
#include <random>
#include <unordered_map>

#include "util.h"

namespace pbh {

struct SlotDesc {
  std::string name;
  int type = 0;      // 0 = uint64 feasign slot, 1 = float slot
  int is_dense = 0;  // dense: fixed `dim` values per instance
  int is_used = 1;
  int dim = 1;
};

// CSR-per-record: values of slot s are vals[offs[s] .. offs[s+1])
struct SlotRecord {
  uint64_t search_id = 0;
  uint32_t rank = 0, cmatch = 0;
  std::string ins_id;
  std::vector<uint64_t> u64_vals;
  std::vector<uint32_t> u64_offs;
  std::vector<float> f_vals;
  std::vector<uint32_t> f_offs;
  void reset() {
    search_id = 0; rank = cmatch = 0; ins_id.clear();
    u64_vals.clear(); u64_offs.clear(); f_vals.clear(); f_offs.clear();
  }
  // reset that also gives the value buffers back (FLAGS_enable_slotrecord_reset_shrink, data_feed.h:211)
  void reset_shrink() {
    reset();
    std::vector<uint64_t>().swap(u64_vals);
    std::vector<float>().swap(f_vals);
    std::string().swap(ins_id);
  }
};

// Lock-protected free list with a capacity cap (FLAGS_padbox_record_pool_max_size).
class RecordPool {
 public:
  explicit RecordPool(size_t cap = 2000000) : cap_(cap) {}
  ~RecordPool() { for (auto* r : free_) delete r; }
  SlotRecord* get() {
    {
      std::lock_guard<std::mutex> g(mu_);
      if (!free_.empty()) { SlotRecord* r = free_.back(); free_.pop_back(); return r; }
    }
    return new SlotRecord();
  }
  // up to `want` pooled records under ONE lock (readers keep a small stash: no lock per line)
  void get_many(size_t want, std::vector<SlotRecord*>* out) {
    std::lock_guard<std::mutex> g(mu_);
    while (want-- && !free_.empty()) { out->push_back(free_.back()); free_.pop_back(); }
  }
  void put(std::vector<SlotRecord*>& rs) {
    std::lock_guard<std::mutex> g(mu_);
    for (auto* r : rs) {
      if (free_.size() < cap_) { if (shrink_on_reset) r->reset_shrink(); else r->reset(); free_.push_back(r); } else delete r;
    }
    rs.clear();
  }
  size_t idle() { std::lock_guard<std::mutex> g(mu_); return free_.size(); }
  // free every idle record (FLAGS_enable_force_mem_recyle / enbale_slotpool_auto_clear) / change the cap
  size_t trim() {
    std::vector<SlotRecord*> drop;
    { std::lock_guard<std::mutex> g(mu_); drop.swap(free_); }
    for (auto* r : drop) delete r;
    return drop.size();
  }
  void set_cap(size_t cap) { std::lock_guard<std::mutex> g(mu_); cap_ = cap; }
  bool shrink_on_reset = false;
 private:
  std::mutex mu_;
  std::vector<SlotRecord*> free_;
  size_t cap_;
};

// Thread-sharded unique-key collector for the feed pass.
// Open-addressing set of 64-bit keys (0 is kept aside as "empty" marker).
struct KeySet {
  std::vector<uint64_t> tab;
  size_t cnt = 0;
  bool has_zero = false;
  void clear() { std::vector<uint64_t>().swap(tab); cnt = 0; has_zero = false; }
  void grow();
  inline void insert(uint64_t k);
  void dump_sorted(std::vector<uint64_t>* out) const;
};

// Thread-sharded collector of the pass's unique feasigns (role of boxps::PSAgentBase::AddKeys).
// Duplicates are dropped on arrival (a pass repeats its hot keys many times), so memory is
// proportional to the unique keys, not to the feasigns read.
class KeyAgent {
 public:
  explicit KeyAgent(int shards = 64) : shards_(shards), mu_(shards), epoch_(next_epoch()) {}
  void add_keys(const uint64_t* keys, size_t n);
  void add_key(uint64_t k) { add_keys(&k, 1); flush_local(); }
  void flush_local();   // publish this thread's buffered keys (call before finalize)
  // every shard's keys sorted (parallel), shards concatenated
  void finalize(ThreadPool* pool, std::vector<uint64_t>* out);
  void clear() { for (auto& s : shards_) s.clear(); epoch_.store(next_epoch()); }
 private:
  void flush_bucket(size_t s, std::vector<uint64_t>& b);
  std::vector<KeySet> shards_;
  std::vector<std::mutex> mu_;
  // process-wide unique stamp, renewed by clear(): adding threads wipe their recent-key tables when it changes
  static uint64_t next_epoch() { static std::atomic<uint64_t> g{0}; return ++g; }
  std::atomic<uint64_t> epoch_;
};

struct BatchRange { int64_t start; int len; };

class Dataset {
 public:
  Dataset();
  ~Dataset();
  // ---- configuration ----
  std::vector<SlotDesc> slots;
  std::vector<std::string> files;
  std::string pipe_command = "cat";
  std::string parser_plugin;       // path of a .so implementing the plugin ABI (see dataset.cc)
  int thread_num = 1;              // reader threads
  int merge_thread_num = 4;
  int batch_size = 1;
  bool parse_ins_id = false, parse_logkey = false;
  float sample_rate = 1.0f;
  bool enable_pv_merge = false;
  bool collect_keys = true;
  int compact_batches = 1;         // 1: build the compact arena when it is below compact_limit_mb; 2: always; 0: never
  int64_t compact_limit_mb = 16384;
  int rank = 0, world = 1;         // file round-robin across nodes (data_set.cc:1963-1975)

  int finalize_slots();            // builds the used-slot index maps
  // ---- pass lifecycle ----
  int load_into_memory();
  int preload_into_memory();       // async
  int wait_preload_done();
  void release_memory();
  void local_shuffle(uint64_t seed);
  int64_t merge_by_ins_id(int merge_size);   // merge records sharing an ins_id; returns dropped instances
  int slots_shuffle(const std::vector<int>& slot_idx, uint64_t seed, bool inverse = false);  // AUC runner: permute slots across records
  int prepare_train(int nworkers, bool shuffle, uint64_t seed, int64_t min_batches);
  // TDM sampling support (reference LayerWiseSampler::sample_from_dataset, index_sampler.cc:75-128): read the
  // first feasign of one sparse slot of every record, then replace record i by counts[i] copies whose feasign
  // in that slot is ids[k]; copies flagged neg[k] get the first value of label_slot (sparse or float) zeroed.
  int slot_first_values(int slot, uint64_t* out, uint8_t* has) const;
  int expand_records(int slot, int label_slot, const int64_t* counts, const uint64_t* ids, const uint8_t* neg);
  int64_t num_records() const { return arena_only_ ? arena_n_ : (int64_t)records_.size(); }
  // Pass file: the compact arena (and the pass's unique keys) written as-is and mapped back with no
  // per-record decode — a pre-processed pass loads at page-cache speed (role: the reference's
  // PreLoadIntoDisk / archive mode, data_set.cc:2092-2217, which still deserialises record by record).
  int save_arena(const std::string& path);
  int load_arena(const std::string& path);
  int64_t num_batches(int worker) const { return worker < (int)worker_batches_.size() ? (int64_t)worker_batches_[worker].size() : 0; }
  // sizes of one batch: instances and total uint64 feasigns of used sparse slots
  int batch_shape(int worker, int64_t b, int* B, int64_t* n_keys) const;
  // record-major CSR (device batch_build input): keys [n_keys], off [B*(S+1)] global offsets,
  // dense [B, dense_dim]; or slot-major: keys in (slot, ins) order, off = lod [S*B+1]
  int pack_batch(int worker, int64_t b, uint64_t* keys, int32_t* off, float* dense, int slot_major) const;
  int batch_ins_ids(int worker, int64_t b, std::string* joined) const;
  int batch_meta(int worker, int64_t b, uint64_t* search_ids, uint32_t* ranks, uint32_t* cmatch) const;
  const std::vector<uint64_t>& unique_keys() const { return unique_keys_; }
  int n_sparse() const { return (int)sparse_idx_.size(); }
  int dense_dim() const { return dense_dim_; }
  int64_t total_feasigns() const { return total_feasigns_.load(); }
  // ---- archive (binary spill / reload; preload_into_disk) ----
  int save_archive(const std::string& path) const;
  int save_partitioned(const std::string& prefix, int nparts, int mode, uint64_t seed) const;
  int load_archive(const std::string& path);
  std::string last_stats() const;
  size_t trim_record_pool() { return pool_.trim(); }
  size_t idle_records() { return pool_.idle(); }
  void set_record_pool_cap(size_t cap) { pool_.set_cap(cap); }
  void set_record_reset_shrink(bool v) { pool_.shrink_on_reset = v; }

  bool parse_line(const char* line, size_t len, SlotRecord* rec) const;

 private:
  void reader_loop(std::atomic<size_t>* next_file, std::atomic<int>* errors);
  void merge_loop();
  int read_file(const std::string& path);
  std::vector<int> sparse_idx_;      // slot indices of the used uint64 sparse slots, in order
  std::vector<int> slot_to_sparse_;  // slot index -> position among used sparse slots (-1 unused)
  std::vector<int> float_idx_;       // slot index -> position among used float slots
  std::string preload_err_;
  std::vector<int> float_dim_;   // per used float slot: dense dim
  std::vector<int> float_off_;   // per used float slot: offset into the dense row
  int dense_dim_ = 0;
  int n_used_u64_ = 0, n_used_f_ = 0;
  std::vector<SlotRecord*> records_;
  // Compact copy of the pass in records_ order, built by prepare_train (compact_batches): packing
  // a batch is then three contiguous copies instead of one heap record chase per instance.
  // Dropped (arena_ok_ = false) by anything that reorders or edits records_.
  bool arena_ok_ = false;
  std::vector<uint64_t> ar_keys_;    // all feasigns, record after record
  std::vector<int64_t> ar_koff_;     // [n + 1] first feasign of every record
  std::vector<uint32_t> ar_soff_;    // [n * (S + 1)] slot offsets inside the record
  std::vector<float> ar_dense_;      // [n * dense_dim] densified float slots
  // what pack_batch reads: the vectors above, or a pass file mapped by load_arena (arena_only_: there
  // are no SlotRecords then; record-level operations are refused)
  const uint64_t* pk_ = nullptr;
  const int64_t* pko_ = nullptr;
  const uint32_t* pso_ = nullptr;
  const float* pd_ = nullptr;
  bool arena_only_ = false;
  int64_t arena_n_ = 0;
  void* map_base_ = nullptr;
  size_t map_size_ = 0;
  void build_arena();
  void drop_arena();
  std::mutex rec_mu_;
  RecordPool pool_;
  Channel<std::vector<SlotRecord*>> chan_;
  KeyAgent agent_;
  std::vector<uint64_t> unique_keys_;
  std::vector<std::vector<BatchRange>> worker_batches_;
  std::unique_ptr<std::thread> preload_thread_;
  int preload_rc_ = 0;
  std::atomic<int64_t> total_feasigns_{0};
  std::atomic<int64_t> lines_read_{0}, lines_bad_{0};
  double load_seconds_ = 0, merge_seconds_ = 0;
  void* plugin_handle_ = nullptr;
  void* plugin_fn_ = nullptr;
  void* plugin_file_fn_ = nullptr;
  int bind_plugin();
  int read_file_with_plugin(const std::string& path);

 public:
  // string key -> row offset index of the InputTable (pbh_input_index_*); when set, a parser plugin
  // that exports pbx_set_offset_func receives the lookup callback (GetOffsetFunc, data_feed.cc:4487)
  void* input_index = nullptr;
  // cores the reader / merge threads are pinned to, round-robin (boxps::get_readins_cores,
  // data_set.cc:2028; FLAGS_enable_binding_train_cpu); empty = not pinned
  std::vector<int> reader_cores;
  // slots (indices into `slots`) whose feasigns are NOT parameter-server keys: left out of the feed
  // pass (slot_omit_in_feedpass, data_feed.cc:3119-3146)
  std::vector<int> omit_slots;
  void collect_record_keys(const SlotRecord* r);
  bool parser_file_mode = false;   // the plugin's pbx_parse_file reads whole files (FLAGS_enable_ins_parser_file)
  bool record_from_plugin(void* plugin_ctx, SlotRecord* rec) const;

 private:
};

// reference compute_batch_num (data_set.cc:280-331): every worker gets the same number of batches
void compute_batch_ranges(int64_t ins_num, int batch_size, int nworkers, std::vector<BatchRange>* out);

}  // namespace pbh
