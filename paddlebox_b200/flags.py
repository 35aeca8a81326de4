"""Runtime flag registry (the reference's exported gflags).

Every flag can be set from the environment as ``FLAGS_<name>`` before import, or at run
time with ``set_flags({...})`` / ``FLAGS.<name> = v`` — the same two channels Paddle's
``PADDLE_DEFINE_EXPORTED_*`` flags offer (paddle/fluid/platform/flags.cc:929-1011,
framework/boxps_worker.cc:43-58, framework/data_set.cc:42).
"""
from __future__ import annotations

import os
from typing import Any


class _Flags:
    def __init__(self):
        object.__setattr__(self, "_defs", {})
        object.__setattr__(self, "_vals", {})

    def define(self, name: str, default: Any, doc: str = "") -> None:
        self._defs[name] = (type(default), default, doc)
        env = os.environ.get("FLAGS_" + name)
        self._vals[name] = self._parse(name, env) if env is not None else default

    def _parse(self, name: str, text: Any) -> Any:
        typ = self._defs[name][0]
        if isinstance(text, typ) and not (typ is int and isinstance(text, bool)):
            return text
        if typ is bool:
            if isinstance(text, (int, float)):
                return bool(text)
            return str(text).strip().lower() in ("1", "true", "yes", "on")
        return typ(text)

    def __getattr__(self, name: str) -> Any:
        try:
            return self._vals[name]
        except KeyError:
            raise AttributeError(f"unknown flag {name!r}") from None

    def __setattr__(self, name: str, value: Any) -> None:
        if name not in self._defs:
            raise AttributeError(f"unknown flag {name!r}")
        self._vals[name] = self._parse(name, value)

    def names(self):
        return sorted(self._defs)

    def describe(self) -> str:
        return "\n".join(f"{n} = {self._vals[n]!r}  # {self._defs[n][2]}" for n in self.names())


FLAGS = _Flags()
_d = FLAGS.define

# --- pull / push -------------------------------------------------------------
_d("enable_pullpush_dedup_keys", True, "dedup keys before pull/push (flags.cc:946); always on here: the pull kernel dedups in its first phase")
_d("enable_pull_box_padding_zero", True, "empty slot -> one zero row (pull_box_sparse_op.h:131-147)")
_d("gpugraph_hbm_table_load_factor", 0.75, "max load factor of the HBM table")
_d("gpugraph_enable_gpu_direct_access", True, "owner writes requester buffers through peer pointers (always on here)")
_d("gpugraph_enable_segment_merge_grads", False, "two-level gradient merge for hot keys; accepted, not needed: the push merges with vector atomics into the unique row")
_d("gpugraph_merge_grads_segment_size", 128, "segment size of the two-level merge (see gpugraph_enable_segment_merge_grads)")
_d("gpugraph_dedup_pull_push_mode", 1, "dedup mode; one mode here (hash dedup inside the pull kernel, restore index reused by the push)")
# --- dense sync ------------------------------------------------------------------
_d("enable_sync_dense_moment", False, "also average Adam moments in K-step sync (boxps_worker.cc:468-475)")
_d("enable_dense_nccl_barrier", False, "barrier before dense sync")
_d("padbox_enable_sharding_stage", False, "ZeRO-1 style dense sharding (boxps_worker.cc:53-56)")
_d("dense_use_nccl", False, "use the NCCL baseline path instead of the symmetric-memory collectives")
# --- dataset -----------------------------------------------------------------------
_d("enable_binding_train_cpu", True, "pin worker threads to cores (boxps_trainer.cc:165-193)")
_d("enable_shuffle_by_searchid", False, "hash records by search_id in global shuffle")
_d("enable_ins_parser_file", False, "parser plugin consumes whole files")
_d("padbox_record_pool_max_size", 2000000, "SlotRecord object pool cap (data_feed.h:241-428); the native pool frees on release_memory")
_d("padbox_slotrecord_extend_dim", 0, "extra floats appended to each record; accepted, records carry no trailing block here")
_d("padbox_dataset_shuffle_thread_num", 20, "shuffle threads; the network shuffle is one all-to-all, the partition runs on the dataset's threads")
_d("padbox_dataset_merge_thread_num", 20, "merge / key-collection threads")
_d("padbox_dataset_disable_shuffle", False, "skip cross-node shuffle")
_d("padbox_dataset_disable_polling", False, "round-robin file assignment instead of polling")
_d("padbox_disable_ins_shuffle", False, "keep instance order in PrepareTrain (data_set.cc:42)")
_d("padbox_auc_runner_mode", False, "slot-shuffle feature evaluation mode")
_d("use_gpu_replica_cache", False, "enable GpuReplicaCache; the cache object (side_tables.GpuReplicaCache) is created by the user script")
_d("gpu_replica_cache_dim", 8, "row width of the replica cache")
_d("fix_dayid", False, "do not advance the day id between passes")
# --- worker -----------------------------------------------------------------------
_d("padbox_enable_gc", True, "eager free of dead activations; PyTorch autograd frees them, the captured step reuses static buffers")
_d("padbox_enable_print_op_debug", False, "log every op before it runs")
_d("padbox_dump_debug_lineid", "", "dump only this line id")
_d("enable_dump_main_program", False, "write the per-device op list to ./device_<id>_ops.txt")
_d("check_nan_inf", False, "abort + dump on NaN/Inf in a batch (boxps_worker.cc:1325-1336)")
_d("enable_debug_print_metrics_info", False, "verbose AUC computation (metrics.cc:29-31)")
# --- this framework -------------------------------------------------------------------
_d("pbx_use_cuda_graph", True, "capture the training step in a CUDA graph when shapes allow")
_d("pbx_pack_threads", 4, "background threads packing mini-batches into staging buffers (per loader)")
_d("pbx_overlap_wgrad", False, "weight-gradient GEMMs on a side stream, concurrently with the dgrad chain (measured: no gain over the grouped launch)")
_d("pbx_deterministic_merge", False, "push gradients merged by a stable sort + fp64 segmented sum (bit-reproducible, "
   "robust to hot keys) instead of the in-kernel merge; single-rank GPU jobs and the CPU path")
_d("pbx_pass_chunk_rows", 1 << 19, "rows per staging slot of the pass build / write-back pipeline (3 pinned slots)")
_d("pbx_pass_fetch_threads", 4, "CPU threads fetching host-table rows into pinned staging slots during begin_pass")
_d("pbx_dense_overlap_mb", 0.0, "> 0: the dense-gradient allreduce runs in buckets of this many MB, each started from an autograd hook as soon as its gradients are final, on a side stream under the rest of backward (large dense models; takes precedence over the fold)")
_d("pbx_dense_overlap_bf16", False, "overlapped dense allreduce: gradients travel as bf16 (cast before, cast back + scale after)")
_d("pbx_cpu_emulate_bf16", False, "CPU towers round every GEMM operand to bf16 (fp32 accumulation), the numerics of the GPU towers: for bf16-vs-fp32 training comparisons without a GPU")
_d("dump_filed_same_as_aibox", False, "dump fields as 'name:v:v' with the name cut at its first '.', instead of 'name:count:v:v' (boxps_worker.cc:1801-1811)")
_d("lineid_have_extend_info", False, "an instance id 'id<space>extra': the id leads the dump line, the extra part is appended as the last column (boxps_worker.cc:1783-1830)")
_d("enable_print_dump_field_debug", False, "print every dumped field (name, width) while dumping")
_d("enable_print_dump_info_debug", False, "print the number of dumped instances per batch")
_d("enable_force_hbm_recyle", False, "end_pass returns cached device memory to the driver (box_wrapper.cc:198)")
_d("enable_force_mem_recyle", False, "end of a pass releases the dataset's record pool to the OS (box_wrapper.cc:130)")
_d("enbale_slotpool_auto_clear", False, "(sic) the record pool frees its idle records when a dataset releases its memory")
_d("padbox_dataset_enable_unrollinstance", False, "the reference lets the parser plugin unroll loaded instances into several (data_set.cc:2275, ISlotParser::UnrollInstance); here a plugin emits as many records per input line as it wants at parse time (pbx_parse_file), so the flag is accepted and has nothing left to do")
_d("gpugraph_load_node_list_into_hbm", True, "graph engine: node id lists are uploaded with the graph (upload_batch)")
_d("graph_get_neighbor_id", False, "graph datasets also collect the neighbours' ids as pass keys (data_set.cc:532)")
_d("enable_opt_get_features", False, "graph feed: features of sampled nodes are fetched in one batched call (data_feed.cu:371)")
_d("enable_slotrecord_reset_shrink", False, "records returned to the pool also give their value buffers back (data_feed.h:211)")
_d("enable_slotpool_wait_release", False, "the reference's pool may release records asynchronously and wait for it; release is synchronous here (accepted)")
_d("padbox_slotpool_thread_num", 1, "threads of the reference's record pool; records are allocated by the reader threads here (accepted)")
_d("enable_ins_parser_add_file_path", False, "whole-file parser plugins get the file path: always the case here (pbx_parse_file(path, ...)) (accepted)")
_d("pbx_fold_dense_allreduce", True, "data-parallel dense-gradient allreduce rides on the sparse push kernel's barriers")
_d("pbx_sparse_dist", False, "GPU worlds: route pull/push through torch.distributed all-to-all (multi-node) instead of the one-node peer-memory engine; CPU worlds with > 1 rank always route this way")
_d("pbx_sparse_cap_factor", 2.0, "mailbox capacity = factor * max_keys / nranks")


def set_flags(d: dict) -> None:
    for k, v in d.items():
        setattr(FLAGS, k[6:] if k.startswith("FLAGS_") else k, v)


def get_flags(names) -> dict:
    if isinstance(names, str):
        names = [names]
    return {n: getattr(FLAGS, n[6:] if n.startswith("FLAGS_") else n) for n in names}
