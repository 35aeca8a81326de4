/* paddlebox_b200 — C ABI of the host runtime (libpbx_host.so, csrc/host/*.cc).
 * Every int function returns 0 on success; otherwise pbh_last_error() holds the reason.
 * Reference citations are relative to the PaddleBox tree. */
#ifndef PBX_HOST_H_
#define PBX_HOST_H_
#include <stdint.h>
#include "pbx_types.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* pbh_last_error(void);

/* ---- dataset / data feed (csrc/host/dataset.cc, api.cc) -----------------------------
 * reference: PadBoxSlotDataset (framework/data_set.cc:1905-2828), SlotPaddleBoxDataFeed
 * (framework/data_feed.cc:2993-4153), compute_batch_num (data_set.cc:280-331) */
int pbh_dataset_create(void** out);
int pbh_dataset_destroy(void* ds);
int pbh_dataset_set_slots(void* ds, int n, const char* const* names, const int* types, const int* is_dense,
                          const int* is_used, const int* dims);
int pbh_dataset_set_filelist(void* ds, int n, const char* const* files);
int pbh_dataset_set_int(void* ds, const char* key, int64_t v);
int pbh_dataset_set_float(void* ds, const char* key, double v);
int pbh_dataset_set_str(void* ds, const char* key, const char* v);
int pbh_dataset_load_into_memory(void* ds);
int pbh_dataset_preload_into_memory(void* ds);
int pbh_dataset_wait_preload_done(void* ds);
int pbh_dataset_release_memory(void* ds);
int pbh_dataset_local_shuffle(void* ds, uint64_t seed);
/* merge the loaded records that share an ins_id (merge_size > 0: groups of another size are dropped;
 * groups with a sparse slot filled twice are dropped); returns the number of dropped instances */
int64_t pbh_dataset_merge_by_ins_id(void* ds, int merge_size);
int pbh_dataset_slots_shuffle(void* ds, const int* slots, int n, uint64_t seed);
/* undoes pbh_dataset_slots_shuffle called with the same slots and seed */
int pbh_dataset_slots_shuffle_back(void* ds, const int* slots, int n, uint64_t seed);
/* TDM sampling: first feasign of a sparse slot per record (has[i] = 0 when the slot is empty), and the rewrite of
 * the record list: record i becomes counts[i] copies with that feasign replaced by ids[k]; copies with neg[k] get
 * the first value of label_slot (-1: none) zeroed; records with an empty slot or count 0 are dropped */
int pbh_dataset_slot_first_values(void* ds, int slot, uint64_t* out, uint8_t* has);
int pbh_dataset_expand_records(void* ds, int slot, int label_slot, const int64_t* counts, const uint64_t* ids, const uint8_t* neg);
int pbh_dataset_prepare_train(void* ds, int nworkers, int shuffle, uint64_t seed, int64_t min_batches);
int64_t pbh_dataset_num_records(void* ds);
int64_t pbh_dataset_num_batches(void* ds, int worker);
int pbh_dataset_n_sparse(void* ds);
int pbh_dataset_dense_dim(void* ds);
int64_t pbh_dataset_total_feasigns(void* ds);
int pbh_dataset_batch_shape(void* ds, int worker, int64_t b, int* B, int64_t* n_keys);
int pbh_dataset_pack_batch(void* ds, int worker, int64_t b, uint64_t* keys, int32_t* off, float* dense,
                           int slot_major);
int pbh_dataset_batch_meta(void* ds, int worker, int64_t b, uint64_t* search_ids, uint32_t* ranks,
                           uint32_t* cmatch);
int64_t pbh_dataset_batch_ins_ids(void* ds, int worker, int64_t b, char* buf, int64_t cap);
int64_t pbh_dataset_unique_keys(void* ds, uint64_t* out, int64_t max_n);
int pbh_dataset_save_archive(void* ds, const char* path);
int pbh_dataset_load_archive(void* ds, const char* path);
/* global shuffle, sending side: partition files "<prefix>.<rank>.<dst>"; mode 0 random, 1 search_id, 2 ins_id hash */
int pbh_dataset_save_partitioned(void* ds, const char* prefix, int nparts, int mode, uint64_t seed);
int64_t pbh_dataset_stats(void* ds, char* buf, int64_t cap);
int pbh_dataset_parse_line(void* ds, const char* line, int64_t* n_u64, int64_t* n_f32, uint64_t* u64_out,
                           int64_t u64_cap, float* f32_out, int64_t f32_cap, uint64_t* search_id);
int64_t pbh_compute_batch_ranges(int64_t ins_num, int batch_size, int nworkers, int64_t* starts, int* lens,
                                 int64_t cap);

/* parser plugin ABI (role of ISlotParser .so plugins, data_feed.cc:3604-3670): a plugin exports
 *   int pbx_parse_line(const char* line, size_t len, void* ctx,
 *                      void (*emit_u64)(void*, int slot, const uint64_t*, size_t),
 *                      void (*emit_f32)(void*, int slot, const float*, size_t),
 *                      void (*emit_meta)(void*, const char* ins_id, uint64_t search_id,
 *                                        uint32_t rank, uint32_t cmatch));
 * and returns 0 for a good line.  Whole-file mode (FLAGS_enable_ins_parser_file; ParseFileInstance,
 * data_feed.cc:3795-3883): the plugin may instead / also export
 *   int pbx_parse_file(const char* path, void* ctx, void (*begin_record)(void*),
 *                      emit_u64, emit_f32, emit_meta,           // as above
 *                      int (*end_record)(void*));               // 0 = record kept
 * and reads the file itself (any container format), bracketing every record with begin / end.
 * Dataset option "parser_file_mode" = 1 selects it. */

/* ---- host-memory table tier + CPU pull/push + checkpoints (csrc/host/host_table.cc) ----
 * reference: libbox_ps memory tier, pscore MemorySparseTable::Save
 * (distributed/ps/table/memory_sparse_table.cc:338-393), CtrDymfAccessor (ctr_dymf_accessor.cc) */
int pbh_table_create(void** out, int embedx_dim, int optimizer, const PbxSparseOptConfig* cfg, uint64_t seed);
int pbh_table_destroy(void* t);
int64_t pbh_table_size(void* t);
int pbh_table_set_config(void* t, const PbxSparseOptConfig* cfg);
int pbh_table_set_test_mode(void* t, int on);
int pbh_table_clear(void* t);
/* scale > 0: serving saves (modes 1, 2) store embedx as int16 multiples of `scale`, serving loads
 * dequantise (quantised feature layout, fleet/box_wrapper.cu:36-43); 0 turns it off */
int pbh_table_set_serving_quant(void* t, float scale);
int pbh_table_fetch_rows(void* t, const uint64_t* keys, int64_t n, float* out, int create);
int pbh_table_store_rows(void* t, const uint64_t* keys, int64_t n, const float* rows, int mark_dirty);
int pbh_table_pull(void* t, const uint64_t* keys, int64_t n, float* out, int out_stride);
int pbh_table_push(void* t, const uint64_t* keys, int64_t n, const float* push, int push_stride);
int64_t pbh_table_shrink(void* t, float decay, float delete_threshold, float delete_after_unseen_days);
int pbh_table_save(void* t, const char* dir, int mode, int part_shard_base, float base_threshold,
                   float delta_threshold, int clear_dirty, int64_t* n_saved);
int pbh_table_load_file(void* t, const char* path, int serving, int64_t* n_loaded);
/* same, keeping only the keys rank `rank` of `world` owns: a model saved by N ranks loads into M */
int pbh_table_load_file_owned(void* t, const char* path, int serving, int rank, int world, int64_t* n_loaded);
int64_t pbh_table_keys(void* t, uint64_t* out, int64_t max_n);

/* parallel sort + unique of keys in place (end of a feed pass: PSGPUWrapper::PreBuildTask,
 * fleet/ps_gpu_wrapper.cc:114-335); returns the unique count */
int64_t pbh_sort_unique_u64(uint64_t* keys, int64_t n);

/* keys grouped by owning rank (owner = pbx_owner_of): out = rank 0's keys, rank 1's, ...; counts[r] */
int pbh_partition_by_owner(const uint64_t* keys, int64_t n, int nranks, uint64_t* out, int64_t* counts);

/* ---- graph store (csrc/host/graph.cc) ---------------------------------------------------
 * reference: GraphGpuWrapper / GpuPsGraphTable (heter_ps/graph_gpu_wrapper.h:34-99,
 * graph_gpu_ps_table_inl.cu), GraphDataGenerator walks (data_feed.cu:200-1206).  Edge type = small int;
 * node ids are arbitrary uint64, mapped to dense indices in sorted-id order inside a type. */
int pbh_graph_create(void** out);
int pbh_graph_destroy(void* g);
int pbh_graph_add_edges(void* g, int type, const uint64_t* src, const uint64_t* dst, int64_t n, int reverse);
int pbh_graph_load_edge_file(void* g, int type, const char* path, int reverse, int64_t* n_loaded);
int pbh_graph_build(void* g, int type);                     /* edge list -> CSR */
int64_t pbh_graph_num_nodes(void* g, int type);
int64_t pbh_graph_num_edges(void* g, int type);
int pbh_graph_num_types(void* g);
int pbh_graph_export_csr(void* g, int type, uint64_t* node_ids, int64_t* row_ptr, int64_t* col);
/* up to k neighbours per node without replacement into out[i*k..], counts into out_cnt[i] */
int pbh_graph_sample_neighbors(void* g, int type, const uint64_t* nodes, int64_t n, int k, uint64_t seed,
                               uint64_t* out, int32_t* out_cnt);
/* walks [n, walk_len]; step s follows edge type metapath[s % n_meta]; out_len = valid prefix length */
int pbh_graph_random_walk(void* g, const int* metapath, int n_meta, const uint64_t* starts, int64_t n,
                          int walk_len, uint64_t seed, uint64_t* out, int32_t* out_len);

/* weighted edges (load_edge_file with a weight column; weighted neighbour sample of graph_gpu_ps_table_inl.cu) */
int pbh_graph_add_edges_weighted(void* g, int type, const uint64_t* src, const uint64_t* dst, const float* w,
                                 int64_t n, int reverse);
int pbh_graph_load_weighted_edge_file(void* g, int type, const char* path, int reverse, int64_t* n_loaded);
int pbh_graph_is_weighted(void* g, int type);
int pbh_graph_export_weights(void* g, int type, float* out);
int pbh_graph_sample_neighbors_weighted(void* g, int type, const uint64_t* nodes, int64_t n, int k, uint64_t seed,
                                        uint64_t* out, int32_t* out_cnt);
/* node types with feature slots (load_node_file / get_feature_of_nodes / get_all_feature_ids / load_node_weight,
 * heter_ps/graph_gpu_wrapper.h:50-99) */
int pbh_graph_add_nodes(void* g, int ntype, const uint64_t* ids, int64_t n, const int64_t* offsets,
                        const int32_t* slots, const uint64_t* values);
int pbh_graph_load_node_file(void* g, int ntype, const char* path, const char* type_name, char sep, int64_t* n_loaded);
int64_t pbh_graph_num_typed_nodes(void* g, int ntype);
int pbh_graph_export_typed_nodes(void* g, int ntype, uint64_t* out);
int pbh_graph_node_features(void* g, int ntype, const uint64_t* nodes, int64_t n, const int32_t* slot_ids, int n_slots,
                            uint64_t* out);
int pbh_graph_all_feature_ids(void* g, int ntype, uint64_t* out, int64_t* n_out);
int pbh_graph_load_node_weight(void* g, int ntype, const char* path, int64_t* n_set);
int pbh_graph_node_weights(void* g, int ntype, const uint64_t* nodes, int64_t n, float* out);
/* partitions for paging a graph larger than HBM (make_partitions / get_partition_num / load_next_partition /
 * get_partition / export_partition_files, graph_gpu_wrapper.h:57-62) */
int pbh_graph_make_partitions(void* g, int type, int64_t byte_size, int device_len, int* n_parts);
int pbh_graph_partition_num(void* g, int type);
int pbh_graph_next_partition(void* g, int type);
int pbh_graph_partition_shape(void* g, int type, int part, int64_t* n_nodes, int64_t* n_edges);
int pbh_graph_export_partition(void* g, int type, int part, uint64_t* node_ids, int64_t* row_ptr, uint64_t* col_ids,
                               float* weights);
int pbh_graph_export_partition_files(void* g, int type, const char* dir);

/* InputTable key index: string key -> row offset, 0 = miss (InputTable::AddIndexData / GetIndexOffset,
 * fleet/box_wrapper.h:124-197).  Attached to a dataset, a parser plugin that also exports
 *   void pbx_set_offset_func(uint64_t (*fn)(void* user, const char* key, size_t len), void* user)
 * gets the lookup as its parse-time callback (GetOffsetFunc, data_feed.cc:4487-4492). */
int pbh_input_index_create(void** out);
int pbh_input_index_destroy(void* ix);
uint64_t pbh_input_index_add(void* ix, const char* key, size_t len, int* fresh);
uint64_t pbh_input_index_get(void* ix, const char* key, size_t len);
uint64_t pbh_input_index_size(void* ix);
uint64_t pbh_input_index_miss(void* ix);
/* pass file: the compact arena + unique keys of a prepared pass written as-is; load maps it back with no
 * per-record decode (role of PreLoadIntoDisk / archive mode, data_set.cc:2092-2217).  The loading
 * dataset must have the same used slots; record-level operations (slots_shuffle, ins ids, PV meta,
 * global shuffle) are refused on such a pass. */
int pbh_dataset_save_arena(void* ds, const char* path);
int pbh_dataset_load_arena(void* ds, const char* path);
int pbh_dataset_set_input_index(void* ds, void* ix);
/* reader / merge threads are pinned round-robin to these cores (get_readins_cores, data_set.cc:2028) */
int pbh_dataset_set_reader_cores(void* ds, const int* cores, int n);
/* slots (indices in set_slots order) whose feasigns are not PS keys: skipped by the key collection
 * (slot_omit_in_feedpass of InitializeGPUAndLoadModel, data_feed.cc:3119-3146) */
int pbh_dataset_set_omit_slots(void* ds, const int* slot_indices, int n);
/* feed configuration of src onto dst (not the file list, not rank / world): streaming readers keep several handles */
int pbh_dataset_copy_config(void* dst, void* src);
/* record pool (FLAGS_padbox_record_pool_max_size, enable_force_mem_recyle): idle records kept for the next pass */
int64_t pbh_dataset_trim_record_pool(void* ds);
int64_t pbh_dataset_idle_records(void* ds);
int pbh_dataset_set_record_pool_cap(void* ds, int64_t cap);
/* SSD tier (closed-library LoadSSD2Mem / CheckNeedLimitMem, fleet/box_wrapper.cc:1316-1324): cold rows
 * move to one append-only file per shard under `dir`; a lookup that misses host memory promotes the
 * row back.  spill: rows unseen for >= min_unseen_days (negative: ignore age), then — if
 * max_mem_rows > 0 — the longest-unseen rows until the memory tier fits; dirty rows stay. */
int pbh_table_ssd_open(void* t, const char* dir);
int pbh_table_ssd_close(void* t, int unlink_files);
int64_t pbh_table_spill(void* t, float min_unseen_days, int64_t max_mem_rows);
int64_t pbh_table_load_ssd2mem(void* t);
int64_t pbh_table_mem_rows(void* t);
int64_t pbh_table_ssd_rows(void* t);

/* ---- AUC calculators (csrc/host/metrics.cc); reference fleet/metrics.cc:38-587 ---- */
int pbh_auc_create(void** out, int table_size);
int pbh_auc_destroy(void* a);
int pbh_auc_reset(void* a);
int pbh_auc_add(void* a, const float* pred, const float* label, const float* mask, int64_t n, int float_label);
int pbh_auc_add_uid(void* a, const float* pred, const float* label, const uint64_t* uid, int64_t n);
int pbh_auc_merge_tables(void* a, const double* tables, const double* sums);
int pbh_auc_export(void* a, double* tables, double* sums);
int pbh_auc_import(void* a, const double* tables, const double* sums);
int pbh_auc_compute(void* a, double* out8);
int pbh_auc_compute_wuauc(void* a, double* out4);
int pbh_auc_nan_inf(void* a, double* out2);

/* ---- async dense table (csrc/host/api.cc); reference boxps_worker.cc:61-370 ---- */
int pbh_async_dense_create(void** out, int64_t n, float lr, int threads, int merge_limit);
int pbh_async_dense_destroy(void* t);
int pbh_async_dense_set_params(void* t, const float* p);
int pbh_async_dense_pull(void* t, float* out);
int pbh_async_dense_push(void* t, const float* g);
int pbh_async_dense_set_lr_range(void* t, int64_t begin, int64_t end, float lr);   /* lr_map, boxps_worker.cc:199-207 */
int pbh_async_dense_wait(void* t);
int64_t pbh_async_dense_updates(void* t);

#ifdef __cplusplus
}
#endif
#endif /* PBX_HOST_H_ */
