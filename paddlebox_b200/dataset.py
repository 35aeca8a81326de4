"""Datasets: pass-scoped, in-memory SlotRecord data with the reference's Python surface.

Reference: ``DatasetFactory`` / ``InMemoryDataset`` / ``QueueDataset`` / ``BoxPSDataset`` /
``PadBoxSlotDataset`` / ``InputTableDataset`` (python/paddle/fluid/dataset.py:66-1497; fleet
flavour python/paddle/distributed/fleet/dataset/dataset.py) over the C++ datasets
(framework/data_set.cc) and feeds (framework/data_feed.cc).  Here the C++ side is
``csrc/host/dataset.cc``; this module keeps the setter names (``set_use_var``,
``set_batch_size``, ``set_thread``, ``set_filelist``, ``set_pipe_command``,
``set_parse_ins_id``, ``set_date``, ``preload_into_memory`` ...) so a user script ports by
changing the import.

Batches come out as ``SlotBatch`` objects (``ops/pull_box_sparse.py``).  On a GPU the host
packs a batch record-major into ONE pinned staging buffer, issues ONE H2D copy and the
``batch_build`` kernel produces the slot-major layout (csrc/cuda/batch_build.cu); the next
batch is packed by a background thread while the current one trains.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
import queue
from dataclasses import dataclass, field
from typing import Iterator, Sequence

import numpy as np
import torch

from . import _native as N
from .box_wrapper import BoxWrapper
from .flags import FLAGS
from .host import NativeDataset
from .ops.pull_box_sparse import SlotBatch
from .utils import monitor


@dataclass
class SlotVar:
    """Stand-in for the reference's program ``Variable`` passed to ``set_use_var``:
    int64 + lod_level 1 = sparse id slot, float32 + lod_level 0 = dense slot of ``shape[-1]``
    values (dataset.py:270-302)."""
    name: str
    dtype: str = "int64"
    lod_level: int = 1
    shape: tuple = (-1, 1)
    is_used: bool = True

    def desc(self) -> dict:
        is_float = self.dtype in ("float32", "float")
        dim = int(self.shape[-1]) if self.lod_level == 0 and len(self.shape) else 1
        return {"name": self.name, "type": 1 if is_float else 0, "is_dense": int(self.lod_level == 0),
                "is_used": int(self.is_used), "dim": max(dim, 1)}


def sparse_slot(name: str) -> SlotVar:
    return SlotVar(name, "int64", 1, (-1, 1))


def dense_slot(name: str, dim: int = 1) -> SlotVar:
    return SlotVar(name, "float32", 0, (-1, dim))


def _is_id_list_file(path: str) -> bool:
    """A text file of item ids (digits and white space) rather than a binary tree-index file."""
    try:
        with open(path, "rb") as f:
            head = f.read(256)
    except OSError:
        return False
    return bool(head) and all(c in b"0123456789 \t\r\n" for c in head)


class DatasetBase:
    def __init__(self):
        self._native = NativeDataset()
        self.use_vars: list = []
        self.all_slots: list = []
        self.batch_size = 1
        self.thread_num = 1
        self.filelist: list = []
        self.pipe_command = "cat"
        self.fs_name = self.fs_ugi = ""
        self.download_cmd = "cat"
        self.trainer_num = 1
        # FLAGS_padbox_dataset_merge_thread_num (default 20 in the reference): capped by the cores there are
        self._native.set("merge_thread_num", max(1, min(FLAGS.padbox_dataset_merge_thread_num, os.cpu_count() or 1, 8)))
        self._prepared_workers = 0
        self._rank, self._world = 0, 1
        self._device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
            else torch.device("cpu")

    # ---- configuration (same names as the reference) ------------------------------------
    def init(self, batch_size=1, thread_num=1, use_var=(), pipe_command="cat", input_type=0, fs_name="",
             fs_ugi="", download_cmd="cat", **kw):
        self.set_batch_size(batch_size)
        self.set_thread(thread_num)
        if use_var:
            self.set_use_var(use_var)
        self.set_pipe_command(pipe_command)
        self.set_hdfs_config(fs_name, fs_ugi)
        self.set_download_cmd(download_cmd)
        for k, v in kw.items():
            setter = getattr(self, "set_" + k, None)
            if setter is None:
                raise TypeError(f"unknown dataset option {k!r}")
            setter(v)

    def set_use_var(self, var_list: Sequence[SlotVar], all_slots: Sequence[SlotVar] | None = None):
        """``var_list`` are the variables the program consumes; ``all_slots`` (optional) is the
        full slot order of the data files — slots not in ``var_list`` are parsed over
        (``is_used = False``)."""
        self.use_vars = list(var_list)
        used = {v.name for v in var_list}
        slots = list(all_slots) if all_slots is not None else list(var_list)
        self.all_slots = [SlotVar(s.name, s.dtype, s.lod_level, s.shape, s.name in used) for s in slots]
        self._native.set_slots([s.desc() for s in self.all_slots])

    def set_batch_size(self, batch_size: int):
        self.batch_size = int(batch_size)
        self._native.set("batch_size", self.batch_size)

    def set_thread(self, thread_num: int):
        self.thread_num = int(thread_num)
        self._native.set("thread_num", self.thread_num)

    def set_filelist(self, filelist: Sequence[str]):
        self.filelist = list(filelist)
        self._native.set_filelist(self.filelist)

    def set_pipe_command(self, pipe_command: str):
        """A shell filter every file is piped through, or the path of a parser plugin ``.so``
        (reference: ``pipe_command`` may name a parser library, data_feed.cc:3604-3670)."""
        self.pipe_command = pipe_command
        if pipe_command.endswith(".so"):
            self._native.set("parser_plugin", pipe_command)
            self._native.set("pipe_command", "cat")
            # FLAGS_enable_ins_parser_file: the plugin's pbx_parse_file reads whole files
            self._native.set("parser_file_mode", int(FLAGS.enable_ins_parser_file))
        else:
            self._native.set("pipe_command", pipe_command)

    # ---- the rest of the reference's DatasetBase surface (python/paddle/fluid/dataset.py:66-395) -------------------
    def set_uid_slot(self, uid_slot: str):
        """Name of the slot whose first feasign is the user id of the instance (``MultiSlotDesc.uid_slot``,
        data_feed.cc:1101, :1222): the per-user metrics read it (``uid_of_batch``) and ``set_shuffle_by_uid``
        partitions by it."""
        self.uid_slot = uid_slot

    def uid_of_batch(self, batch) -> "torch.Tensor":
        """[B] user ids of a ``SlotBatch``: the first key of the uid slot of every instance (0 for an empty bag)."""
        names = [s.name for s in self.all_slots if s.is_used and s.lod_level > 0]
        si = names.index(self.uid_slot)
        B = batch.batch_size
        lod = batch.lod[si * B:(si + 1) * B + 1].long()
        first = batch.keys[lod[:-1].clamp(max=max(batch.keys.numel() - 1, 0))]
        return torch.where(lod[1:] > lod[:-1], first, torch.zeros_like(first))

    def set_input_type(self, input_type: int):
        """0 = one sample per generated record, 1 = one batch (``DataFeedDesc.input_type``)."""
        self.input_type = int(input_type)

    def set_so_parser_name(self, so_parser_name: str):
        """Parser plugin library (``DataFeedDesc.so_parser_name``): same as giving the ``.so`` as pipe command."""
        self.so_parser_name = so_parser_name
        if so_parser_name:
            self.set_pipe_command(so_parser_name)

    def set_feed_type(self, data_feed_type: str):
        """Name of the native feed class.  The SlotRecord feed is the one implementation here; the Record-based
        ``MultiSlot*`` feeds are ``paddlebox_b200.legacy_feed``."""
        self.data_feed_type = data_feed_type

    def _set_use_ps_gpu(self, psgpu):
        """Bind the dataset to a ``PSGPUWrapper`` (``set_date`` then reaches it; dataset.py:347-359)."""
        self.use_ps_gpu, self.psgpu = psgpu is not None, psgpu

    def _set_heter_ps(self, enable_heter_ps: bool = False):
        self.enable_heter_ps = bool(enable_heter_ps)

    def _set_trainer_num(self, trainer_num: int):
        self.trainer_num = int(trainer_num)

    def set_graph_config(self, config: dict):
        """Graph-walk feed configuration (walk_degree, walk_len, window, batch_size, meta_path ...): kept for
        ``paddlebox_b200.graph.GraphDataGenerator(**config)``."""
        self.graph_config = dict(config)

    def _prepare_to_run(self):
        """Before a pass: no more reader threads than files, at least one (dataset.py:336-345, :426-447)."""
        n = max(1, min(int(self.thread_num), max(len(self.filelist), 1)))
        if n != self.thread_num:
            self.set_thread(n)

    def _finish_to_run(self):
        """After a pass the reference destroys its readers; reader threads here end with ``load_into_memory``."""

    def _dynamic_adjust_before_train(self, thread_num: int):
        """The trainer runs with ``thread_num`` workers: deal the loaded pass to that many (dataset.py:450-460)."""
        self._adjusted_from = getattr(self, "_prepared_workers", 0)
        if getattr(self, "_prepared_workers", 0) not in (0, int(thread_num)) and self._native.num_records() > 0:
            self.prepare_train(int(thread_num))

    def _dynamic_adjust_after_train(self):
        self._adjusted_from = 0

    def set_hdfs_config(self, fs_name: str, fs_ugi: str):
        self.fs_name, self.fs_ugi = fs_name, fs_ugi

    def set_download_cmd(self, download_cmd: str):
        self.download_cmd = download_cmd

    def set_trainer_num(self, n: int):
        self.trainer_num = n

    def set_thread_num(self, thread_num: int):
        """pybind name of ``set_thread`` (data_set_py.cc)."""
        self.set_thread(thread_num)

    def _desc(self) -> str:
        return self.desc()

    # getters of the native dataset object (pybind/data_set_py.cc) --------------------------------------------------
    def get_data_feed_desc(self) -> str:
        return self.desc()

    def get_filelist(self) -> list:
        return list(self.filelist)

    def get_thread_num(self) -> int:
        return int(self.thread_num)

    def get_trainer_num(self) -> int:
        return int(self.trainer_num)

    def get_hdfs_config(self) -> tuple:
        return (self.fs_name, self.fs_ugi)

    def get_download_cmd(self) -> str:
        return self.download_cmd

    def get_fleet_send_batch_size(self) -> int:
        return int(getattr(self, "fleet_send_batch_size", 1024))

    def register_client2client_msg_handler(self) -> None:
        """The reference hooks the shuffle receiver into the PS client here; the shuffle transport of this
        framework (parallel/shuffle.py) sets itself up when ``global_shuffle`` runs."""

    def set_data_feed_desc(self, desc: str):
        """Configure from the protobuf TEXT form of a ``DataFeedDesc`` (data_feed.proto:18-57; what ``desc()``
        prints and what the reference hands its native dataset): name, batch size, pipe command, threads, slots
        (+ ``uid_slot``), ``rank_offset``, ``ads_offset``, ``pv_batch_size``, ``input_type``, ``so_parser_name``,
        ``index_parser``, ``sample_rate``, ``graph_config``."""
        from .fleet_desc import parse_text_format
        d = parse_text_format(desc)
        if "name" in d:
            self.set_feed_type(str(d["name"]))
        if "batch_size" in d:
            self.set_batch_size(int(d["batch_size"]))
        if "pipe_command" in d:
            self.set_pipe_command(str(d["pipe_command"]))
        if "thread_num" in d:
            self.set_thread(int(d["thread_num"]))
        msd = d.get("multi_slot_desc") or {}
        raw = msd.get("slots", [])
        raw = raw if isinstance(raw, list) else [raw]
        slots, used = [], []
        for f in raw:
            is_float = str(f.get("type", "uint64")) == "float"
            dense = bool(f.get("is_dense", False))
            shape = f.get("shape", 1)
            dim = int(shape[-1] if isinstance(shape, list) else shape)
            v = SlotVar(str(f["name"]), "float32" if is_float else "int64", 0 if dense else 1, (-1, max(dim, 1)))
            slots.append(v)
            if bool(f.get("is_used", False)):
                used.append(v)
        if slots:
            self.set_use_var(used, slots)
        if msd.get("uid_slot"):
            self.set_uid_slot(str(msd["uid_slot"]))
        for key, setter in (("rank_offset", "set_rank_offset"), ("ads_offset", "set_ads_offset"),
                            ("pv_batch_size", "set_pv_batch_size"), ("input_type", "set_input_type"),
                            ("so_parser_name", "set_so_parser_name"), ("index_parser", "set_index_parser"),
                            ("sample_rate", "set_sample_rate"), ("graph_config", "set_graph_config")):
            if key in d and hasattr(self, setter):
                getattr(self, setter)(d[key])

    def _check_use_var_with_data_generator(self, var_list, data_generator_class, test_file: str) -> None:
        """Run the job's data generator over ``test_file`` and check every sample against ``var_list``: same
        number of slots, no empty slot, float slots hold floats and id slots ints (fleet dataset.py:275-335)."""
        n = len(var_list)
        with open(test_file) as f:
            for line in f:
                for sample in data_generator_class.generate_sample(line)():
                    if len(sample) != n:
                        raise ValueError(f"var length mismatch error: var_list = {n} vs data_generator = {len(sample)}")
                    for var, (name, vals) in zip(var_list, sample):
                        if len(vals) == 0:
                            raise ValueError(f"var length error: var {name}'s length in data_generator is 0")
                        is_float = str(getattr(var, "dtype", "")).lower().replace("torch.", "") in ("float32", "float", "fp32")
                        if is_float and not all(isinstance(e, float) for e in vals):
                            raise TypeError(f"var dtype mismatch error: var name = {name}, var type in var_list = float, "
                                            f"while var in data_generator contains non-float value, which is {vals}")
                        if not is_float and not all(isinstance(e, int) for e in vals):
                            raise TypeError(f"var dtype mismatch error: var name = {name}, var type in var_list = int, "
                                            f"while var in data_generator contains non-int value, which is {vals}")

    def desc(self) -> str:
        """Protobuf text of the feed configuration (``DataFeedDesc``, data_feed.proto:45-57); round-trips through
        ``set_data_feed_desc``."""
        q = lambda v: '"' + str(v).replace("\\", "\\\\").replace('"', '\\"') + '"'      # noqa: E731
        b = lambda v: "true" if v else "false"                                                      # noqa: E731
        lines = [f"name: {q(getattr(self, 'data_feed_type', None) or type(self).__name__)}", f"batch_size: {self.batch_size}",
                 "multi_slot_desc {"]
        for s in self.all_slots:
            d = s.desc()
            lines += ["  slots {", f"    name: {q(d['name'])}", f"    type: {q('float' if d['type'] else 'uint64')}",
                      f"    is_dense: {b(d['is_dense'])}", f"    is_used: {b(d['is_used'])}", f"    shape: {d['dim']}", "  }"]
        if getattr(self, "uid_slot", ""):
            lines.append(f"  uid_slot: {q(self.uid_slot)}")
        lines.append("}")
        lines += [f"pipe_command: {q(self.pipe_command)}", f"thread_num: {self.thread_num}"]
        for key, attr in (("rank_offset", "rank_offset_name"), ("ads_offset", "ads_offset_name"),
                          ("so_parser_name", "so_parser_name"), ("index_parser", "index_parser")):
            if getattr(self, attr, ""):
                lines.append(f"{key}: {q(getattr(self, attr))}")
        if getattr(self, "pv_batch_size", None):
            lines.append(f"pv_batch_size: {int(self.pv_batch_size)}")
        if getattr(self, "input_type", None) is not None:
            lines.append(f"input_type: {int(self.input_type)}")
        if getattr(self, "sample_rate", None) not in (None, 1.0):
            lines.append(f"sample_rate: {float(self.sample_rate)}")
        gc = getattr(self, "graph_config", None)
        if gc:
            lines.append("graph_config {")
            for k, v in gc.items():
                lines.append(f"  {k}: {q(v) if isinstance(v, str) else (b(v) if isinstance(v, bool) else v)}")
            lines.append("}")
        return "\n".join(lines)

    # ---- geometry ---------------------------------------------------------------------------
    @property
    def num_sparse_slots(self) -> int:
        return self._native.n_sparse()

    @property
    def dense_dim(self) -> int:
        return self._native.dense_dim()


def _alias_private_setters(cls):
    """The fleet flavour of the dataset API (python/paddle/distributed/fleet/dataset/dataset.py) spells its setters
    ``_set_x`` and has users go through ``init(**kwargs)`` / ``update_settings(**kwargs)``: both spellings work."""
    for name in list(vars(cls)):
        if name.startswith("set_") and not hasattr(cls, "_" + name):
            setattr(cls, "_" + name, getattr(cls, name))
    return cls


class InMemoryDataset(DatasetBase):
    """Loads a pass into host memory, shuffles, deals equal batch counts to the workers."""

    def __init__(self):
        super().__init__()
        self.parse_ins_id = False
        self.parse_logkey = False
        self._merge_by_lineid = False
        self.fleet_send_batch_size = 1024
        self.queue_num = None
        self.fea_eval = False
        self._shuffle_seed = 0
        self._worker = 0

    # setters (dataset.py:372-1105) -----------------------------------------------------------
    def set_date(self, date: str):
        """Training date (``YYYYMMDD``) for pulls, saves and loads — reaches the bound wrapper
        (``_set_use_ps_gpu``) or the process-wide ``BoxWrapper`` (dataset.py:740-761)."""
        self._date = str(date)
        ps = getattr(self, "psgpu", None)
        if ps is None:
            from .box_wrapper import BoxWrapper
            ps = BoxWrapper._instance if hasattr(BoxWrapper, "_instance") else None
        if ps is not None:
            ps.set_date(self._date)

    def set_parse_ins_id(self, v: bool):
        self.parse_ins_id = bool(v)
        self._native.set("parse_ins_id", int(v))

    def set_parse_logkey(self, v: bool):
        self.parse_logkey = bool(v)
        self._native.set("parse_logkey", int(v))

    def set_parse_content(self, v: bool):
        self.parse_content = bool(v)

    def set_merge_by_sid(self, v: bool):
        self.merge_by_sid = bool(v)

    def set_shuffle_by_uid(self, v: bool):
        """Global shuffle keeps the instances of one user (ins_id prefix) on one rank."""
        self.shuffle_by_uid = bool(v)

    def set_current_phase(self, phase: int):
        """1 = join, 0 = update: PV-merged batches are dealt only in the join phase
        (data_set.cc:2782-2783)."""
        self.current_phase = int(phase)
        if getattr(self, "_pv_merge", False):
            self._native.set("enable_pv_merge", int(phase == 1))

    def enable_pv_merge(self) -> bool:
        return bool(getattr(self, "_pv_merge", False))

    def preprocess_instance(self):
        """Group instances into page views before a join-phase pass (``PreprocessInstance``)."""
        if getattr(self, "_pv_merge", False):
            self._native.set("enable_pv_merge", 1)

    def postprocess_instance(self):
        """Back to one instance per record after the join phase (``PostprocessInstance``)."""
        self._native.set("enable_pv_merge", 0)

    def merge_by_lineid(self) -> int:
        """pybind name: apply the ins_id merge now (``MergeByInsId``)."""
        return self.merge_by_ins_id()

    def set_generate_unique_feasigns(self, generate_uni_feasigns: bool = True, shard_num: int = 1):
        self.gen_uni_feasigns, self.local_shard_num = bool(generate_uni_feasigns), int(shard_num)

    def generate_local_tables_unlock(self, table_id: int = 0, fea_dim: int = 0, read_thread_num: int = 1,
                                     consume_thread_num: int = 1, shard_num: int | None = None) -> list:
        """Unique feasigns of the loaded pass, split into ``shard_num`` local tables by ``key % shard_num``
        (``GenerateLocalTablesUnlock``, data_set.cc:771-877).  Returns the list of key arrays."""
        shard_num = shard_num or getattr(self, "local_shard_num", 1)
        keys = self._native.unique_keys()
        self.local_tables = [keys[keys % np.uint64(shard_num) == np.uint64(s)] for s in range(shard_num)]
        return self.local_tables

    def set_preload_thread_num(self, n: int):
        self.preload_thread_num = int(n)

    def create_channel(self):
        """Channels live inside the native dataset and are opened by every load; kept for API parity."""

    def create_readers(self):
        """Readers = the batch dealing of ``prepare_train`` (``CreateReaders``)."""
        if self._prepared_workers == 0 and self._native.num_records():
            self.prepare_train(max(1, self.trainer_num), 0)

    def destroy_readers(self):
        self._prepared_workers = 0

    def create_preload_readers(self):
        """Preload readers are the native reader threads started by ``preload_into_memory``."""

    def destroy_preload_readers(self):
        self._native.wait_preload_done()

    def dynamic_adjust_channel_num(self, channel_num: int, discard_remaining_ins: bool = False):
        self.queue_num = int(channel_num)

    def set_gpu_graph_mode(self, v: bool = True):
        """Batches come from ``graph.GraphDataGenerator`` walks instead of files."""
        self.gpu_graph_mode = bool(v)

    def _tdm_tensor_sampler(self, tree_name: str, tree_path, tdm_layer_counts, start_sample_layer: int = 1,
                            with_hierarchy: bool = False, seed: int = 0, id_slot: int = 0, branch: int = 2):
        """Bind a TDM item tree to the dataset (``tdm_sample`` with a list of item ids instead of a tree file).  ``tree_path``: a file with one
        item id per line (leaf order) or a sequence of item ids.  Returns ``sampler(item_ids) ->
        (nodes, labels, mask)`` (``ops.tdm.tdm_sample`` with ``tdm_layer_counts`` negatives per layer),
        to be applied to the ids of slot ``id_slot`` in the model's forward."""
        from .ops.tdm import TDMTree, tdm_sample
        if isinstance(tree_path, str):
            with open(tree_path) as f:
                items = [int(x) for x in f.read().split()]
        else:
            items = list(tree_path)
        tree = TDMTree(items, branch)
        gen = torch.Generator().manual_seed(seed)
        self.tdm = {"name": tree_name, "tree": tree, "id_slot": id_slot, "with_hierarchy": with_hierarchy}

        def sampler(item_ids: torch.Tensor):
            g = gen if item_ids.device.type == "cpu" else None
            return tdm_sample(tree.to(item_ids.device), item_ids, tdm_layer_counts, start_sample_layer, g)
        return sampler

    def set_enable_pv_merge(self, v: bool):
        self._pv_merge = bool(v)
        self._native.set("enable_pv_merge", int(v))

    def set_merge_by_lineid(self, merge_size: int = 2):
        """Instances sharing an ins_id are merged into one after loading (``MergeByInsId``,
        data_set.cc:1387-1581); needs ``set_parse_ins_id(True)``."""
        self._merge_by_lineid = True
        self.merge_size = merge_size

    def merge_by_ins_id(self) -> int:
        """Apply the merge now; returns the number of dropped instances."""
        return self._native.merge_by_ins_id(getattr(self, "merge_size", 0))

    def set_sample_rate(self, rate: float):
        self.sample_rate = float(rate)
        self._native.set("sample_rate", float(rate))

    def set_queue_num(self, n: int):
        self.queue_num = n

    def set_fleet_send_batch_size(self, n: int = 1024):
        self.fleet_send_batch_size = n

    def set_fleet_send_sleep_seconds(self, s: int = 0):
        self.fleet_send_sleep_seconds = s

    def set_fea_eval(self, record_candidate_size: int, fea_eval: bool = True):
        self.fea_eval = fea_eval
        self.candidate_size = record_candidate_size

    def set_rank_offset(self, name: str):
        self.rank_offset_name = name

    def set_pv_batch_size(self, n: int):
        self.pv_batch_size = n

    def set_ads_offset(self, name: str):
        self.ads_offset_name = name

    def set_compact_batches(self, mode: int = 1, limit_mb: int | None = None):
        """Compact copy of the pass built by ``prepare_train`` so that packing a batch is three
        contiguous copies (0 never, 1 when it fits ``limit_mb`` [default 16 GB], 2 always)."""
        self._native.set("compact_batches", int(mode))
        if limit_mb is not None:
            self._native.set("compact_limit_mb", int(limit_mb))

    def set_merge_thread_num(self, n: int):
        self._native.set("merge_thread_num", int(n))

    def set_shuffle_seed(self, seed: int):
        self._shuffle_seed = int(seed)

    def set_rank_world(self, rank: int, world: int):
        """Files are dealt round-robin over nodes (SetFileList, data_set.cc:1963-1975)."""
        self._native.set("rank", rank)
        self._native.set("world", world)
        self._rank, self._world = int(rank), int(world)

    # pass data ---------------------------------------------------------------------------------
    def _bind_reader_cores(self, overlapped: bool):
        """FLAGS_enable_binding_train_cpu: while a pass is PRE-loaded next to a training pass, reader / merge
        threads go to the read-ins half of this process's cores (data_set.cc:2028) and leave the other half to
        the training threads.  A synchronous load has the machine to itself and stays unpinned."""
        cores = []
        if overlapped and FLAGS.enable_binding_train_cpu:
            from .parallel.cluster import get_readins_cores
            cores = get_readins_cores()
        self._native.set_reader_cores(cores if len(cores) > 1 else [])

    def load_into_memory(self):
        self._bind_reader_cores(False)
        self._native.load_into_memory()
        if self._merge_by_lineid:
            self.merge_by_ins_id()
        monitor.stat_set("STAT_total_feasign_num_in_mem", self._native.total_feasigns())

    def preload_into_memory(self, thread_num: int | None = None):
        thread_num = thread_num or getattr(self, "preload_thread_num", None)
        if thread_num:
            self.set_thread(thread_num)
        self._bind_reader_cores(True)
        self._native.preload_into_memory()

    def wait_preload_done(self):
        self._native.wait_preload_done()
        if self._merge_by_lineid:
            self.merge_by_ins_id()
        monitor.stat_set("STAT_total_feasign_num_in_mem", self._native.total_feasigns())

    def release_memory(self):
        """Give the pass's records back.  They go to the native record pool for the next pass (up to
        ``FLAGS.padbox_record_pool_max_size``); with ``FLAGS.enable_force_mem_recyle`` or
        ``FLAGS.enbale_slotpool_auto_clear`` the pool is emptied, i.e. the memory returns to the allocator
        (box_wrapper.cc:130-133, :178)."""
        self._native.set_record_pool_cap(int(FLAGS.padbox_record_pool_max_size))
        self._native.set("record_reset_shrink", int(FLAGS.enable_slotrecord_reset_shrink))
        self._native.release_memory()
        if FLAGS.enable_force_mem_recyle or FLAGS.enbale_slotpool_auto_clear:
            self._native.trim_record_pool()
        self._prepared_workers = 0

    def local_shuffle(self):
        self._native.local_shuffle(self._shuffle_seed)

    def set_shuffle_dir(self, path: str):
        """Directory every rank can reach: ``global_shuffle`` then exchanges records through archive
        files there instead of over the network."""
        self._shuffle_dir = path

    def global_shuffle(self, fleet=None, thread_num: int = 12, mode: str | None = None):
        """Re-partition the loaded records over all ranks (``ShuffleData``, data_set.cc:2422-2604),
        then shuffle locally.  One process: a local shuffle.  ``mode``: ``random`` (default),
        ``search_id`` (``FLAGS_enable_shuffle_by_searchid``) or ``ins_id``."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            import tempfile
            from .parallel.shuffle import global_shuffle_records
            if mode is None:
                mode = "search_id" if FLAGS.enable_shuffle_by_searchid else \
                    "ins_id" if getattr(self, "shuffle_by_uid", False) else "random"
            shared = getattr(self, "_shuffle_dir", None)
            work = shared or os.path.join(tempfile.gettempdir(), "pbx_global_shuffle")
            self._shuffle_round = getattr(self, "_shuffle_round", 0) + 1
            global_shuffle_records(self, work, mode=mode, seed=self._shuffle_seed,
                                   tag=f"shuffle{self._shuffle_round}", transport="files" if shared else "network")
        self._native.local_shuffle(self._shuffle_seed)

    def slots_shuffle(self, slots: Sequence[str]):
        """Permute the feasigns of ``slots`` across instances (AUC runner feature evaluation)."""
        names = [s.name for s in self.all_slots]
        self._native.slots_shuffle([names.index(s) for s in slots], self._shuffle_seed + 1)

    def slots_shuffle_back(self, slots: Sequence[str]):
        """Undo the last ``slots_shuffle(slots)`` (reference ``RecordReplaceBack``, data_feed.h:1772-1962)."""
        names = [s.name for s in self.all_slots]
        self._native.slots_shuffle_back([names.index(s) for s in slots], self._shuffle_seed + 1)

    def get_memory_data_size(self, fleet=None) -> int:
        return self._native.num_records()

    def get_shuffle_data_size(self, fleet=None) -> int:
        return self._native.num_records()

    def get_pv_data_size(self) -> int:
        return self._native.num_records()

    def unique_keys(self) -> np.ndarray:
        return self._native.unique_keys()

    def tdm_sample(self, tree_name: str, tree_path: str, tdm_layer_counts, start_sample_layer: int = 1,
                   with_hierachy: bool = False, seed: int = 0, id_slot=0, label_slot=None) -> int:
        """Layer-wise TDM sampling of the loaded records (data_set.cc:694-768, index_sampler.cc:75-128): a record
        whose ``id_slot`` holds item i becomes one positive record per node on i's path in the tree (the slot's
        feasign replaced by the node id) plus that layer's negatives (a random node of the layer, label zeroed).
        ``id_slot`` / ``label_slot``: slot index or name; the label defaults to the slot after ``id_slot``.
        Records without a value in ``id_slot`` are dropped.  Returns the new record count.  (``tree_path`` given as
        a sequence of item ids, or a text file of them, binds the tensor-side sampler of the base class instead.)"""
        from .index_dataset import IndexSampler, IndexWrapper
        if not isinstance(tree_path, str) or _is_id_list_file(tree_path):
            # a plain list of item ids (no tree file): the tensor-side sampler to call in the model's forward
            return self._tdm_tensor_sampler(tree_name, tree_path, tdm_layer_counts, start_sample_layer, with_hierachy,
                                            seed, id_slot if isinstance(id_slot, int) else 0)
        names = [s.name for s in self.all_slots]
        idx = names.index(id_slot) if isinstance(id_slot, str) else int(id_slot)
        if label_slot is None:
            lab = idx + 1 if idx + 1 < len(names) else -1
        else:
            lab = names.index(label_slot) if isinstance(label_slot, str) else int(label_slot)
        IndexWrapper().insert_tree_index(tree_name, tree_path)
        sampler = IndexSampler("by_layerwise", tree_name)
        sampler.init_layerwise_conf(list(tdm_layer_counts), start_sample_layer, seed)
        targets, has = self._native.slot_first_values(idx)
        counts = np.where(has, sampler.layer_counts_sum, 0).astype(np.int64)
        codes, labels = sampler.sample_codes(targets[has])
        ids = sampler.tree.node_ids_of(codes.reshape(-1))
        self._native.expand_records(idx, lab, counts, ids, (labels.reshape(-1) == 0))
        return self._native.num_records()

    def _generate_local_tables_unlock(self, table_id=0, fea_dim=0, read_thread_num=1, consume_thread_num=1,
                                      shard_num=1) -> int:
        """The reference builds per-shard key tables from the loaded records for the heterogeneous PS
        (fleet dataset.py:783-787, data_set.cc ``GenerateLocalTablesUnlock``).  Here the reader threads collect
        the keys while loading; this returns how many distinct keys the loaded data holds."""
        return int(self.unique_keys().size)

    def stats(self) -> str:
        return self._native.stats()

    # preload_into_disk / archive mode (data_set.cc:2092-2217) ---------------------------------
    def set_archivefile(self, path: str):
        self._archive = path

    def preload_into_disk(self, path: str | None = None):
        self._native.load_into_memory()
        self._native.save_archive(path or self._archive)
        self._native.release_memory()

    def load_into_disk(self, path: str | None = None):
        self.preload_into_disk(path)

    def wait_load_disk_done(self):
        self._native.load_archive(self._archive)

    # pass files: a prepared pass written as-is, mapped back with no decode ---------------------------
    def save_pass_file(self, path: str):
        """Write the loaded pass (in its current order: call ``local_shuffle`` / ``prepare_train`` first
        for a shuffled file) and its unique keys as one pass file.  A later ``load_pass_file`` maps it
        back at page-cache speed — no text parse, no per-record decode, no key collection."""
        self._native.save_arena(path)

    def load_pass_file(self, path: str):
        """Make the pass file the loaded pass (the dataset must use the same slots).  The instance order
        is the file's; ``prepare_train(shuffle=True)`` permutes whole batches.  Instance ids, PV fields and
        record-level shuffles are not available on such a pass."""
        self._native.load_arena(path)
        self._prepared_workers = 0

    # training iteration ---------------------------------------------------------------------------
    def prepare_train(self, num_workers: int = 1, worker: int = 0, shuffle: bool | None = None,
                      min_batches: int = 0):
        """Shuffle + deal batches so that every worker (= GPU) gets the same number
        (``PrepareTrain``, data_set.cc:2692-2823)."""
        if shuffle is None:
            shuffle = not FLAGS.padbox_disable_ins_shuffle
        self._native.prepare_train(num_workers, shuffle, self._shuffle_seed, min_batches)
        self._prepared_workers = num_workers
        self._worker = worker

    def dynamic_adjust_readers_num(self, thread_num: int):
        self.prepare_train(thread_num)

    def num_batches(self, worker: int | None = None) -> int:
        return self._native.num_batches(self._worker if worker is None else worker)

    def batches(self, worker: int | None = None, device: torch.device | None = None, prefetch: int = 2,
                with_ins_id: bool = False) -> Iterator[SlotBatch]:
        """Iterate this worker's batches as ``SlotBatch`` (double-buffered H2D on GPU)."""
        if self._prepared_workers == 0:
            self.prepare_train(1)
        w = self._worker if worker is None else worker
        device = device or self._device
        loader = _BatchLoader(self._native, w, device, prefetch, with_ins_id, self._want_meta(),
                              getattr(self, "max_rank", 3))
        return iter(loader)

    def _want_meta(self) -> bool:
        """PV batches carry search_id / rank / cmatch and the rank-offset matrix when a rank-offset
        variable was named (``set_rank_offset``) or logkeys are parsed."""
        return bool(getattr(self, "rank_offset_name", "")) or bool(getattr(self, "parse_logkey", False))

    def set_max_rank(self, max_rank: int):
        self.max_rank = int(max_rank)

    def host_batches(self, worker: int | None = None, prefetch: int = 4, with_ins_id: bool = False):
        """Iterate this worker's batches as pinned, record-major :class:`~paddlebox_b200.feed.HostBatch`
        es (no device work): the input of the pipelined feed ``CTRTrainer.train_host_batches``."""
        if self._prepared_workers == 0:
            self.prepare_train(1)
        w = self._worker if worker is None else worker
        loader = _BatchLoader(self._native, w, torch.device("cuda"), prefetch, with_ins_id, self._want_meta(),
                              getattr(self, "max_rank", 3))
        return loader.host_batches()


class QueueDataset(InMemoryDataset):
    """Streaming dataset (reference ``QueueDataset``, python/paddle/fluid/dataset.py:1106-1230; each DataFeed reads
    its share of the files and hands batches to its worker as it goes — nothing is held for a whole pass, nothing
    is shuffled, no pass keys are collected).

    ``batches()`` without a previous ``load_into_memory()`` streams: this rank's files (``filelist[rank::world]``)
    are read ``thread_num`` at a time; while the batches of one group are being consumed, a background thread
    parses the next group into a second native handle, so host memory holds two groups of files at most and the
    reader never waits for a whole pass.  Batches do not span groups (the last batch of a group may be short).
    ``load_into_memory()`` is still accepted (compatibility): the dataset then behaves like an unshuffled
    in-memory dataset."""

    streaming = True

    def __init__(self):
        super().__init__()
        self._native.set("collect_keys", 0)
        self._stream_handles = None

    def local_shuffle(self):
        raise NotImplementedError("QueueDataset does not support local shuffle")

    def global_shuffle(self, fleet=None, thread_num=12):
        raise NotImplementedError("QueueDataset does not support global shuffle")

    def prepare_train(self, num_workers=1, worker=0, shuffle=None, min_batches=0):
        super().prepare_train(num_workers, worker, False, min_batches)

    def is_loaded(self) -> bool:
        return self._native.num_records() > 0 or self._prepared_workers > 0

    def file_groups(self) -> list:
        mine = self.filelist[self._rank::max(1, self._world)]
        g = max(1, int(self.thread_num))
        return [mine[i:i + g] for i in range(0, len(mine), g)]

    def batches(self, worker: int | None = None, device: torch.device | None = None, prefetch: int = 2,
                with_ins_id: bool = False) -> Iterator[SlotBatch]:
        if self.is_loaded():
            return super().batches(worker, device, prefetch, with_ins_id)
        return self._stream(device or self._device, prefetch, with_ins_id, host=False)

    def host_batches(self, worker: int | None = None, prefetch: int = 4, with_ins_id: bool = False):
        if self.is_loaded():
            return super().host_batches(worker, prefetch, with_ins_id)
        return self._stream(torch.device("cuda"), prefetch, with_ins_id, host=True)

    def _stream(self, device, prefetch: int, with_ins_id: bool, host: bool):
        import threading
        from .host import NativeDataset
        groups = self.file_groups()
        if self._stream_handles is None:
            self._stream_handles = [NativeDataset(), NativeDataset()]
        handles = self._stream_handles
        for h in handles:
            h.copy_config_from(self._native)       # rank / world stay (0, 1): the groups are already this rank's
            h.set("collect_keys", 0)
        self.stream_stats = {"groups": len(groups), "records": 0, "max_resident_records": 0}

        def load(h, files):
            h.release_memory()
            h.set_filelist(files)
            h.load_into_memory()
            h.prepare_train(1, False, 0, 0)

        err: list = []

        def load_bg(h, files):
            try:
                load(h, files)
            except BaseException as e:  # noqa: BLE001  (re-raised in the consumer)
                err.append(e)

        if not groups:
            return
        load(handles[0], groups[0])
        for gi in range(len(groups)):
            cur = handles[gi % 2]
            th = None
            if gi + 1 < len(groups):
                th = threading.Thread(target=load_bg, args=(handles[(gi + 1) % 2], groups[gi + 1]), daemon=True)
                th.start()
            n_cur = cur.num_records()
            self.stream_stats["records"] += n_cur
            loader = _BatchLoader(cur, 0, device, prefetch, with_ins_id, self._want_meta(), getattr(self, "max_rank", 3))
            try:
                yield from (loader.host_batches() if host else iter(loader))
            finally:
                if th is not None:       # also when the consumer stops early: the next stream reuses the handles
                    th.join()
            if th is not None:
                if err:
                    raise err[0]
                self.stream_stats["max_resident_records"] = max(self.stream_stats["max_resident_records"],
                                                                n_cur + handles[(gi + 1) % 2].num_records())
            cur.release_memory()
        self.stream_stats["max_resident_records"] = max(self.stream_stats["max_resident_records"], n_cur)


class FileInstantDataset(QueueDataset):
    """``FileInstantDataset`` (python/paddle/fluid/dataset.py:1193-1229, ``MultiSlotFileInstantDataFeed``): files
    are only ever processed as a stream — the streaming path of :class:`QueueDataset` with the in-memory
    fallback removed."""

    def load_into_memory(self):
        raise NotImplementedError("FileInstantDataset processes files as a stream; use InMemoryDataset to load a pass")

    preload_into_memory = load_into_memory

    def local_shuffle(self):
        raise NotImplementedError("FileInstantDataset does not support local shuffle, please use InMemoryDataset "
                                  "for local_shuffle")

    def global_shuffle(self, fleet=None, thread_num=12):
        raise NotImplementedError("FileInstantDataset does not support global shuffle, please use InMemoryDataset "
                                  "for global_shuffle")


class BoxPSDataset(InMemoryDataset):
    """Dataset bound to the BoxWrapper pass lifecycle (``BoxHelper``, box_wrapper.h:1043-1295):
    ``load_into_memory`` = BeginFeedPass -> read + collect keys -> EndFeedPass."""

    def __init__(self):
        super().__init__()
        self._date = ""

    @property
    def boxps(self) -> BoxWrapper:
        return BoxWrapper.get_instance()

    def set_date(self, date: str):
        self._date = str(date)
        self.boxps.set_date(self._date)

    def begin_pass(self):
        self.boxps.begin_pass()

    def end_pass(self, need_save_delta: bool = False, delta_path: str = ""):
        self.boxps.end_pass(need_save_delta, delta_path)

    def _feed_keys(self):
        self.boxps.begin_feed_pass(self._date or None)
        self.boxps.feed_pass(self._native.unique_keys())
        self.boxps.end_feed_pass()

    def _apply_omitted_slots(self):
        """Slots named in ``slot_omit_in_feedpass`` carry ids that are not embedding keys (e.g. input-table
        offsets): their feasigns stay out of the feed pass (data_feed.cc:3119-3146)."""
        omit = {str(x) for x in getattr(self.boxps, "slot_omit_in_feedpass", ())}
        self._native.set_omit_slots([i for i, s in enumerate(self.all_slots) if s.name in omit])

    def load_into_memory(self):
        self._apply_omitted_slots()
        super().load_into_memory()
        self._feed_keys()

    def preload_into_memory(self, thread_num: int | None = None):
        self._apply_omitted_slots()
        super().preload_into_memory(thread_num)

    def wait_preload_done(self):
        super().wait_preload_done()
        self._feed_keys()

    def wait_feed_pass_done(self):
        self.wait_preload_done()

    def load_pass_file(self, path: str):
        super().load_pass_file(path)
        self._feed_keys()

    def read_ins_into_memory(self):
        """``BoxHelper::ReadData2Memory``: BeginFeedPass -> LoadIntoMemory -> EndFeedPass."""
        self.load_into_memory()


class BoxPS:
    """``core.BoxPS(dataset)`` — the reference's ``BoxHelper`` (pybind/box_helper_py.cc:38-67, box_wrapper.h:1043-1295):
    the pass lifecycle calls for one dataset.  ``BoxPSDataset`` carries the same methods itself; this is the object
    scripts that construct the helper directly expect."""

    def __init__(self, dataset: "BoxPSDataset"):
        if not isinstance(dataset, BoxPSDataset):
            raise TypeError("BoxPS needs a BoxPSDataset / PadBoxSlotDataset")
        self.dataset = dataset

    def set_date(self, year, month=None, day=None):
        """``set_date(2021, 5, 17)`` as the helper takes it, or one ``"20210517"`` string."""
        self.dataset.set_date(f"{int(year):04d}{int(month):02d}{int(day):02d}" if month is not None else str(year))

    def begin_pass(self):
        self.dataset.begin_pass()

    def end_pass(self, need_save_delta: bool = False):
        self.dataset.end_pass(need_save_delta)

    def load_into_memory(self):
        self.dataset.load_into_memory()

    def preload_into_memory(self):
        self.dataset.preload_into_memory()

    def wait_feed_pass_done(self):
        self.dataset.wait_feed_pass_done()

    def read_ins_into_memory(self):
        self.dataset.read_ins_into_memory()

    def slots_shuffle(self, slots):
        self.dataset.slots_shuffle(list(slots))


class PadBoxSlotDataset(BoxPSDataset):
    """SlotRecord dataset of PaddleBox (``PadBoxSlotDataset``, dataset.py:1363-1497)."""

    def disable_polling(self, v: bool = True):
        FLAGS.padbox_dataset_disable_polling = v

    def disable_shuffle(self, v: bool = True):
        FLAGS.padbox_dataset_disable_shuffle = v


class InputTableDataset(PadBoxSlotDataset):
    """Dataset whose instances carry string keys of an ``InputTable`` (replica side table,
    data_set.cc:2830-2858); index files are parsed like data files."""

    def set_index_filelist(self, filelist: Sequence[str]):
        self.index_filelist = list(filelist)

    def set_input_table_dim(self, dim: int):
        self.input_table_dim = int(dim)

    def set_index_parser(self, index_parser: str):
        """Parser plugin for the INDEX files (``DataFeedDesc.index_parser``; the data files keep the dataset's own
        parser).  A library given here must export the line-parser ABI of ``csrc/host/dataset.cc``."""
        self.index_parser = index_parser

    @property
    def input_table(self):
        """The ``InputTable`` of the current BoxWrapper (created on first use)."""
        from .side_tables import InputTable
        box = self.boxps
        if getattr(box, "input_table", None) is None:
            box.input_table = InputTable(getattr(self, "input_table_dim", 1))
        return box.input_table

    def load_index_into_memory(self) -> int:
        """Read the index files (``key<TAB>v1 v2 ...``) into the input table
        (``InputTableDataset::LoadIndexIntoMemory``, data_set.cc:2830-2858; feed
        ``InputIndexDataFeed``, data_feed.cc:4574-4633).  Returns the number of rows added."""
        tab = self.input_table
        return sum(tab.load_index_file(f) for f in getattr(self, "index_filelist", []))

    def index_offsets(self, keys: Sequence[str]) -> torch.Tensor:
        """Row offsets of string keys (0 = miss): what a parser plugin obtains through the
        reference's ``GetOffsetFunc`` callback and stores in an id slot for ``lookup_input``."""
        tab = self.input_table
        return torch.tensor([tab.get_index_offset(k) for k in keys], dtype=torch.int64)

    def load_into_memory(self):
        self.load_index_into_memory()
        # a parser plugin that exports pbx_set_offset_func resolves string keys at parse time
        self._native.set_input_index(self.input_table.handle)
        super().load_into_memory()


def _update_settings(self, **kwargs):
    """``update_settings(**kwargs)`` / ``_init_distributed_settings(**kwargs)`` of the fleet flavour: every key
    is a setter without its ``set_`` prefix (``batch_size``, ``thread_num``, ``use_var``, ``pipe_command``,
    ``input_type``, ``fs_name`` + ``fs_ugi``, ``download_cmd``, ``data_feed_type``, ``queue_num``, ``merge_size``,
    ``parse_ins_id``, ``parse_content``, ``fleet_send_batch_size``, ``fleet_send_sleep_seconds``, ``fea_eval`` +
    ``candidate_size``, ...)."""
    kw = dict(kwargs)
    if "fs_name" in kw or "fs_ugi" in kw:
        self.set_hdfs_config(kw.pop("fs_name", getattr(self, "fs_name", "")), kw.pop("fs_ugi", getattr(self, "fs_ugi", "")))
    if "thread_num" in kw:
        self.set_thread(kw.pop("thread_num"))
    if "merge_size" in kw:
        m = kw.pop("merge_size")
        if m and m > 0:
            self.set_merge_by_lineid(m)
    if "data_feed_type" in kw:
        self.set_feed_type(kw.pop("data_feed_type"))
    if kw.pop("fea_eval", False):
        self.set_fea_eval(kw.pop("candidate_size", 10000), True)
    kw.pop("candidate_size", None)
    for k, v in kw.items():
        setter = getattr(self, "set_" + k, None)
        if setter is None:
            raise TypeError(f"unknown dataset setting {k!r}")
        if v is not None:
            setter(v)


for _cls in (DatasetBase, InMemoryDataset, QueueDataset, FileInstantDataset, BoxPSDataset, PadBoxSlotDataset,
             InputTableDataset):
    _alias_private_setters(_cls)
DatasetBase.update_settings = _update_settings
DatasetBase._init_distributed_settings = _update_settings


class DatasetFactory:
    """``DatasetFactory().create_dataset("PadBoxSlotDataset")`` (dataset.py:30-63)."""
    _KINDS = {"InMemoryDataset": InMemoryDataset, "QueueDataset": QueueDataset, "BoxPSDataset": BoxPSDataset,
              "FileInstantDataset": FileInstantDataset,
              "PadBoxSlotDataset": PadBoxSlotDataset, "InputTableDataset": InputTableDataset,
              "SlotRecordDataset": InMemoryDataset, "MultiSlotDataset": InMemoryDataset}

    def create_dataset(self, datafeed_class: str = "QueueDataset"):
        try:
            return self._KINDS[datafeed_class]()
        except KeyError:
            raise ValueError(f"datafeed class {datafeed_class} does not exist") from None


# --------------------------------------------------------------------------------------------
class _PinnedStage:
    """One pinned staging buffer: [keys u64 | offsets i32 | dense f32]."""

    def __init__(self, key_cap: int, off_cap: int, dense_cap: int, pinned: bool):
        self.key_cap, self.off_cap, self.dense_cap = key_cap, off_cap, dense_cap
        nbytes = key_cap * 8 + off_cap * 4 + dense_cap * 4
        self.buf = torch.empty(nbytes, dtype=torch.uint8, pin_memory=bool(pinned) and torch.cuda.is_available())
        base = self.buf.data_ptr()
        self.keys_ptr = base
        self.off_ptr = base + key_cap * 8
        self.dense_ptr = self.off_ptr + off_cap * 4

    def views(self, n_keys: int, n_off: int, n_dense: int):
        k = self.buf[:n_keys * 8].view(torch.int64)
        o = self.buf[self.key_cap * 8:self.key_cap * 8 + n_off * 4].view(torch.int32)
        d = self.buf[self.key_cap * 8 + self.off_cap * 4:self.key_cap * 8 + self.off_cap * 4 + n_dense * 4] \
            .view(torch.float32)
        return k, o, d


class _BatchLoader:
    """Packs this worker's batches into (pinned) staging buffers on ``pack_threads`` background
    threads — thread t packs batches t, t + T, ... and batches are delivered in order.  One
    packer sustains ~500 batches/s of 4096 x 26-slot instances (it chases one heap record per
    instance); the GPU consumes ~3500/s, hence several (the C call releases the GIL)."""

    def __init__(self, native: NativeDataset, worker: int, device: torch.device, prefetch: int, with_ins_id: bool,
                 with_meta: bool = False, max_rank: int = 3, pack_threads: int | None = None):
        self.native, self.worker, self.device = native, worker, device
        self.with_meta, self.max_rank = with_meta, max_rank
        self.S = native.n_sparse()
        self.Dd = native.dense_dim()
        self.nb = native.num_batches(worker)
        self.with_ins_id = with_ins_id
        self.cuda = device.type == "cuda"
        T = int(pack_threads or FLAGS.pbx_pack_threads)
        self.nw = max(1, min(T, max(self.nb, 1)))
        # stages per packer: the consumer may hold LAG + 2 batches, the rest is prefetch depth
        self.depth = max(2, (max(1, prefetch) + 5 + self.nw - 1) // self.nw)
        self.outq = [queue.Queue() for _ in range(self.nw)]
        self.freeq = [queue.Queue() for _ in range(self.nw)]
        self.made = [0] * self.nw
        self.threads = [threading.Thread(target=self._produce, args=(w,), daemon=True) for w in range(self.nw)]
        self._started = False

    def _fits(self, st: _PinnedStage, B: int, n: int) -> bool:
        return st.key_cap >= n and st.off_cap >= B * (self.S + 1) + 1 and st.dense_cap >= B * self.Dd

    def _stage_for(self, w: int, B: int, n: int) -> _PinnedStage:
        st = None
        if self.made[w] >= self.depth:
            st = self.freeq[w].get()                 # wait for the consumer to hand one back
        else:
            try:
                st = self.freeq[w].get_nowait()
            except queue.Empty:
                st = None
        if st is not None and self._fits(st, B, n):
            return st
        if st is None:
            self.made[w] += 1
        st = _PinnedStage(int(n * 1.25) + 1024, B * (self.S + 1) + self.S + 2, max(B * self.Dd, 1), self.cuda)
        st.owner = w
        return st

    def _release(self, st: _PinnedStage) -> None:
        self.freeq[st.owner].put(st)

    def _produce(self, w: int):
        try:
            for b in range(w, self.nb, self.nw):
                B, n = self.native.batch_shape(self.worker, b)
                st = self._stage_for(w, B, n)
                # GPU: record-major (batch_build transposes on the device); CPU: slot-major directly
                self.native.pack_batch(self.worker, b, st.keys_ptr, st.off_ptr, st.dense_ptr, not self.cuda)
                ids = self.native.batch_ins_ids(self.worker, b) if self.with_ins_id else None
                if self.with_meta:
                    ids = (ids, self._meta(b, B))
                self.outq[w].put((b, B, n, st, ids))
        except BaseException as e:       # surfaces in the consumer
            self.outq[w].put(e)

    def _items(self):
        """(b, B, n, stage, ids) of every batch, in batch order."""
        if not self._started:
            self._started = True
            for t in self.threads:
                t.start()
        for b in range(self.nb):
            item = self.outq[b % self.nw].get()
            if isinstance(item, BaseException):
                raise item
            yield item

    def _meta(self, b: int, B: int) -> dict:
        """search_id / rank / cmatch of the batch's instances and the PV rank-offset matrix
        (``GetRankOffset`` + ``CopyRankOffsetKernel``, data_feed.cc:3531-3591, data_feed.cu:1319-1369)."""
        from .ops.ctr_ops import build_rank_offset
        sid, rk, cm = self.native.batch_meta(self.worker, b, B)
        sid_t = torch.from_numpy(sid.astype(np.int64))
        rk_t = torch.from_numpy(rk.astype(np.int32))
        cm_t = torch.from_numpy(cm.astype(np.int32))
        return {"search_id": sid_t, "rank": rk_t, "cmatch": cm_t,
                "rank_offset": build_rank_offset(sid_t, rk_t, self.max_rank, cmatch=cm_t)}

    def _split(self, ids):
        if self.with_meta:
            ids, meta = ids
            if self.cuda:
                meta = {k: v.to(self.device, non_blocking=True) for k, v in meta.items()}
            return ids, meta
        return ids, None

    def host_batches(self):
        """Packed batches left in pinned memory (the staging buffer returns to the pool when the
        consumer calls ``HostBatch.release``)."""
        from .feed import HostBatch
        S, Dd = self.S, self.Dd
        for item in self._items():
            b, B, n, st, ids = item
            ids, meta = self._split(ids)
            k, o, d = st.views(n, B * (S + 1), B * Dd)
            hb = HostBatch(k, o, d.view(B, Dd) if Dd else None, B, S, n, ins_ids=ids,
                           release=lambda st=st: self._release(st))
            hb.meta = meta
            yield hb

    def __iter__(self):
        S, Dd = self.S, self.Dd
        for item in self._items():
            b, B, n, st, ids = item
            ids, meta = self._split(ids)
            if self.cuda:
                n_off = B * (S + 1)
                k, o, d = st.views(n, n_off, B * Dd)
                # ONE H2D of the staging bytes actually used would need one contiguous range;
                # the three regions are contiguous per kind, so copy them back to back
                dk = k.to(self.device, non_blocking=True)
                do = o.to(self.device, non_blocking=True)
                dd = d.to(self.device, non_blocking=True).view(B, Dd) if Dd else None
                keys = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
                lod = torch.empty(S * B + 1, dtype=torch.int32, device=self.device)
                scratch = torch.empty(1024, dtype=torch.int32, device=self.device)
                N.check(N.cuda().pbx_batch_build(N.ptr(dk), N.ptr(do), B, S, N.ptr(keys), N.ptr(lod),
                                                 N.ptr(scratch), N.stream_ptr()), "pbx_batch_build")
                ev = torch.cuda.Event()
                ev.record()
                batch = SlotBatch(keys[:n], lod, B, S, dd, n_keys=n, ins_ids=ids, meta=meta)
                batch._keepalive = (dk, do, scratch)
                yield batch
                ev.synchronize()   # staging buffer may be reused once its copies are done
                self._release(st)
            else:
                k, o, d = st.views(n, S * B + 1, B * Dd)
                batch = SlotBatch(k.clone(), o.clone(), B, S, d.clone().view(B, Dd) if Dd else None, n_keys=n,
                                  ins_ids=ids, meta=meta)
                self._release(st)
                yield batch
