"""BoxWrapper — the parameter-server façade (reference: ``BoxWrapper`` singleton,
fleet/box_wrapper.h:377-1012, box_wrapper.cc:129-210, :916-1046, :1201-1324; its Python
binding box_helper_py.cc:70-160; and the closed ``libbox_ps.so`` behind it).

One instance per process (= per GPU).  It owns

* the HBM shard of the sparse table (``ops.sparse_table.SparseTable``) and the sharded
  pull/push engine (``parallel.sparse_engine.SparseEngine``) when a GPU is present,
* the host-memory tier (``host.HostTable``): everything outside the current pass's working
  set lives there, and it *is* the table in CPU-only mode,
* the pass lifecycle: ``begin_feed_pass -> (dataset adds keys) -> end_feed_pass ->
  begin_pass`` stages the pass's keys from the host tier into HBM, ``end_pass`` writes them
  back (and can save a delta),
* checkpoints: ``save_base`` (batch model + serving base), ``save_delta``, ``load_model``,
  ``shrink_table``; text layout ``part-<shard>-<file>`` compatible with pscore's
  MemorySparseTable,
* the metric registry (``init_metric`` / ``get_metric_msg``) and the join/update phase.
"""
from __future__ import annotations

import json
import os
import time
from typing import Sequence

import numpy as np
import torch

from . import _native as N
from .flags import FLAGS
from .host import HostTable
from .utils.timer import StageTimers

_MASK64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix64_np(k: np.ndarray) -> np.ndarray:
    k = k.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        k ^= k >> np.uint64(33)
        k *= np.uint64(0xff51afd7ed558ccd)
        k ^= k >> np.uint64(33)
        k *= np.uint64(0xc4ceb9fe1a85ec53)
        k ^= k >> np.uint64(33)
    return k


def owner_of(keys: np.ndarray, nranks: int) -> np.ndarray:
    """Shard of every key; identical to ``pbx_owner_of`` in csrc/cuda/common.cuh."""
    return ((mix64_np(keys) >> np.uint64(40)) % np.uint64(nranks)).astype(np.int64)


def make_day_id(date: str) -> int:
    """'YYYYMMDD' -> days since epoch (reference make_day_id, box_wrapper.cc:57-70)."""
    t = time.strptime(date[:8], "%Y%m%d")
    import calendar
    return int(calendar.timegm(t) // 86400)


class BoxWrapper:
    _instance: "BoxWrapper | None" = None

    # ---- singleton plumbing (SetInstance / GetInstance) ----------------------------
    @classmethod
    def set_instance(cls, embedx_dim: int = 8, expand_embed_dim: int = 0, feature_type: int = 0,
                     pull_embedx_scale: float = 1.0, **kw) -> "BoxWrapper":
        if BoxWrapper._instance is None:      # one PS per process, whichever front end created it
            BoxWrapper._instance = cls(embedx_dim, expand_embed_dim, feature_type, pull_embedx_scale, **kw)
        return BoxWrapper._instance

    @classmethod
    def get_instance(cls) -> "BoxWrapper":
        if BoxWrapper._instance is None:
            raise RuntimeError("BoxWrapper.set_instance(...) has not been called")
        return BoxWrapper._instance

    @classmethod
    def reset_instance(cls) -> None:
        if BoxWrapper._instance is not None:
            BoxWrapper._instance.finalize()
        BoxWrapper._instance = None

    def __init__(self, embedx_dim: int = 8, expand_embed_dim: int = 0, feature_type: int = 0,
                 pull_embedx_scale: float = 1.0, *, optimizer: str = "adagrad", device: torch.device | None = None,
                 rank: int | None = None, world_size: int | None = None, group=None):
        """``feature_type``: 0 = fp32 embedx (HeterPS layout), 1 = adam, 2 = shared adam, 3 = quant:
        the serving models (``save_base`` xbox part, ``save_delta``) hold embedx as int16 multiples of
        ``pull_embedx_scale`` and training pools on that grid (``seqpool_quant_ratio``), while the
        table keeps fp32 master weights.  nncross = ``expand_embed_dim`` trailing embedx columns;
        per-slot embedding widths = one table per width (``MultiDimBox``)  (SURVEY §9.10)."""
        self.embedx_dim = embedx_dim
        self.expand_embed_dim = expand_embed_dim
        self.feature_type = feature_type
        self.pull_embedx_scale = pull_embedx_scale
        self.optimizer = {1: "adam", 2: "shared_adam"}.get(feature_type, optimizer)
        self.cvm_offset = 3                      # show, clk, embed_w (box_wrapper.h:880)
        self.hidden = self.cvm_offset + embedx_dim
        import torch.distributed as dist
        if rank is None:
            rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
        if world_size is None:
            world_size = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank, self.world_size, self.group = rank, world_size, group
        self.use_cuda = torch.cuda.is_available() if device is None else device.type == "cuda"
        self.device = device or (torch.device("cuda", torch.cuda.current_device()) if self.use_cuda
                                 else torch.device("cpu"))
        self.config = N.SparseOptConfig()
        self.host_table: HostTable | None = None
        self.table = None          # SparseTable (HBM)
        self.engine = None         # SparseEngine (one node, peer memory)
        self.router = None         # DistSparseRouter (several nodes / CPU worlds)
        self._phase = 1            # 1 = join, 0 = update (FlipPhase, box_wrapper.h:770-773)
        self._metrics: dict = {}
        self._pass_keys: np.ndarray | None = None
        self._feed_keys: list = []
        self._day_id = 0
        self._test_mode = False
        self._in_pass = False
        self.timers = StageTimers()
        self.slot_vector: list = []
        self.slot_omit_in_feedpass: set = set()
        self.lr_map: dict = {}
        self.host_mem_rows_limit = 0
        self.ssd_unseen_days = -1.0
        self._ssd = False
        self._initialized = False

    # ---- initialisation ----------------------------------------------------------------
    def initialize_gpu_and_load_model(self, conf_file: str | dict | None = None, slot_vector: Sequence[int] = (),
                                      slot_omit_in_feedpass: Sequence[int] = (), model_path: str = "",
                                      lr_map: dict | None = None, *, table_capacity: int | None = None,
                                      max_keys_per_batch: int = 1 << 20) -> None:
        """``conf_file``: dict or path of a JSON file with the sparse optimizer options
        (names of ``SparseOptConfig``) plus optional ``table_capacity`` /
        ``max_keys_per_batch``.  (The closed library reads an opaque conf file;
        box_wrapper.cc:1201-1242.)"""
        conf = {}
        if isinstance(conf_file, dict):
            conf = dict(conf_file)
        elif conf_file:
            with open(conf_file) as f:
                conf = json.load(f)
        ssd_dir = conf.pop("ssd_dir", "")
        self.host_mem_rows_limit = int(conf.pop("host_mem_rows_limit", 0))
        self.ssd_unseen_days = float(conf.pop("ssd_unseen_days", -1.0))
        table_capacity = conf.pop("table_capacity", table_capacity)
        max_keys_per_batch = conf.pop("max_keys_per_batch", max_keys_per_batch)
        self.optimizer = conf.pop("optimizer", self.optimizer)
        self.config = N.SparseOptConfig(**conf)
        self.slot_vector = list(slot_vector)
        self.slot_omit_in_feedpass = set(slot_omit_in_feedpass)
        self.lr_map = dict(lr_map or {})
        self.max_keys_per_batch = int(max_keys_per_batch)
        self.host_table = HostTable(self.embedx_dim, self.optimizer, self.config, seed=self.rank)
        if self.feature_type == 3:
            self.host_table.set_serving_quant(self.pull_embedx_scale)
        if ssd_dir:
            self.host_table.ssd_open(os.path.join(ssd_dir, f"rank{self.rank:03d}"))
            self._ssd = True
        dist_route = self.world_size > 1 and (not self.use_cuda or FLAGS.pbx_sparse_dist)
        if self.use_cuda:
            from .ops.sparse_table import SparseTable
            from .parallel.sparse_engine import SparseEngine
            cap = int(table_capacity or (1 << 22))
            self._pass_builder = None
            self.table = SparseTable(cap, self.embedx_dim, self.optimizer, self.config, seed=self.rank,
                                     device=self.device.index)
            cap_pair = 0
            if self.world_size > 1 and not dist_route:
                cap_pair = int(FLAGS.pbx_sparse_cap_factor * self.max_keys_per_batch / self.world_size) + 1024
            # (with the all-to-all router the engine serves this rank's shard only: status / dump)
            self.engine = SparseEngine(self.table, self.max_keys_per_batch, 0 if dist_route else self.rank,
                                       1 if dist_route else self.world_size, cap=cap_pair, group=self.group)
        if dist_route:
            from .parallel.sparse_dist import DistSparseRouter
            self.router = DistSparseRouter(self.rank, self.world_size, self.hidden, self.group,
                                           host_table=None if self.use_cuda else self.host_table,
                                           device_table=self.table if self.use_cuda else None)
        if model_path:
            self.load_model(model_path)
        self._initialized = True

    @property
    def seqpool_quant_ratio(self) -> int:
        """``quant_ratio`` for ``fused_seqpool_cvm`` that puts pooled embedx on the serving grid
        (0 unless ``feature_type`` 3)."""
        return int(round(1.0 / self.pull_embedx_scale)) if self.feature_type == 3 and self.pull_embedx_scale > 0 else 0

    def get_lr_map(self) -> dict:
        """Per-parameter learning rates for the async dense table: an ``lr_map`` entry whose name contains
        ``param`` covers ``<name>.w_0`` and ``<name>.b_0`` (box_wrapper.cc:1233-1240); other entries are
        taken as full parameter names."""
        out = {}
        for name, lr in self.lr_map.items():
            if "param" in name:
                out[name + ".w_0"] = out[name + ".b_0"] = float(lr)
            else:
                out[name] = float(lr)
        return out

    # ---- pass lifecycle ------------------------------------------------------------------
    def set_date(self, date: str) -> None:
        self._day_id = make_day_id(date) if not FLAGS.fix_dayid or self._day_id == 0 else self._day_id

    def begin_feed_pass(self, date: str | None = None) -> None:
        if date:
            self.set_date(date)
        self._feed_keys = [np.zeros(1, dtype=np.uint64)]  # key 0 always registered (box_wrapper.h:1107)

    def feed_pass(self, keys) -> None:
        """Add keys of the coming pass (``PSAgent::AddKeys``)."""
        k = np.ascontiguousarray(keys.cpu().numpy() if isinstance(keys, torch.Tensor) else keys)
        self._feed_keys.append(k.view(np.uint64) if k.dtype != np.uint64 else k)

    def end_feed_pass(self) -> None:
        """Dedup the pass's keys and route every key to the rank that owns it."""
        t0 = time.time()
        from .host import sort_unique_u64
        keys = sort_unique_u64(np.concatenate(self._feed_keys)) if self._feed_keys else np.zeros(0, np.uint64)
        self._feed_keys = []
        if self.world_size > 1:
            from .host import partition_by_owner
            from .parallel.cluster import Cluster
            # every key goes to the rank that owns it: one native partition, one all-to-all of raw keys
            mine = Cluster(self.group).all_to_all_arrays(partition_by_owner(keys, self.world_size))
            keys = sort_unique_u64(np.concatenate(mine)) if mine else np.zeros(0, np.uint64)
        self._pass_keys = keys
        self.timers.add("feed_pass", time.time() - t0)

    def begin_pass(self) -> None:
        """Stage the pass working set host tier -> HBM shard."""
        assert self._initialized, "call initialize_gpu_and_load_model first"
        t0 = time.time()
        keys = self._pass_keys if self._pass_keys is not None else np.zeros(0, np.uint64)
        if self.use_cuda:
            self.table.clear()
            # streamed build (pass_build.py): host fetch of chunk k+1 into a pinned slot overlaps the H2D copy
            # and the insert kernel of chunk k; staging memory is a fixed pool reused by every pass
            # (ps_gpu_wrapper.cc:684-955, heter_comm_inl.h:671-731, mem_pool.h:28-102)
            self.last_pass_build = self.pass_builder(int(keys.size)).build(keys) if keys.size else {}
        else:
            if keys.size:
                self.host_table.fetch_rows(keys, create=True)  # create missing rows
        self._in_pass = True
        self.timers.add("begin_pass", time.time() - t0)

    def pass_builder(self, n_rows: int = 0):
        """The pass build / write-back pipeline of this shard.  Its pinned staging pool is kept across passes and
        only re-made when a pass needs larger slots than any before (up to ``FLAGS.pbx_pass_chunk_rows`` rows per
        slot), so small jobs do not pin hundreds of megabytes."""
        want = int(FLAGS.pbx_pass_chunk_rows)
        if n_rows > 0:
            c = 1024
            while c < n_rows and c < want:
                c <<= 1
            want = min(want, c)
        pb = getattr(self, "_pass_builder", None)
        if pb is None or (n_rows > 0 and pb.pool.chunk < want):
            from .pass_build import PassBuilder
            pb = self._pass_builder = PassBuilder(self.host_table, self.table.insert, self.host_table.layout.stride,
                                                  self.device, chunk=want, depth=3,
                                                  fetch_threads=int(FLAGS.pbx_pass_fetch_threads))
        return pb

    def end_pass(self, need_save_delta: bool = False, delta_path: str = "") -> None:
        """Write the HBM working set back to the host tier (``EndPass``)."""
        t0 = time.time()
        if self.use_cuda and self._in_pass:
            torch.cuda.synchronize(self.device)
            self.engine.check_status()
            dk, dr = self.table.dump(only_dirty=True, clear_dirty=True)
            if dk.numel():
                # bounded pinned chunks: D2H of chunk k+1 overlaps the host-table store of chunk k
                self.last_pass_write_back = self.pass_builder(int(dk.numel())).write_back(dk, dr, mark_dirty=True)
        self._in_pass = False
        if self.use_cuda and FLAGS.enable_force_hbm_recyle:          # cached device blocks back to the driver
            torch.cuda.empty_cache()
        if need_save_delta and delta_path:
            self.save_delta(delta_path)
        if self._ssd and self.check_need_limit_mem():
            # host tier over its budget: the coldest rows go to the SSD tier
            self.host_table.spill(self.ssd_unseen_days, self.host_mem_rows_limit)
        self.timers.add("end_pass", time.time() - t0)

    # ---- pull / push (single call; autograd wrappers live in ops/pull_box_sparse.py) ----------
    def pull_sparse(self, keys: torch.Tensor, n_dev: torch.Tensor | None = None) -> tuple:
        """Returns (values [U or N, >= hidden], restore int32 [N]); occurrence i's record is
        ``values[restore[i], :hidden]``."""
        if self.router is not None:
            n = keys.numel() if n_dev is None else int(n_dev.reshape(-1)[0])
            return self.router.pull(keys[:n])
        if self.use_cuda:
            restore = self.engine.pull(keys, n_dev=n_dev)
            return self.engine.values, restore
        uk, inv = torch.unique(keys, return_inverse=True)
        vals = self.host_table.pull(uk.numpy())
        self._cpu_unique = uk
        return vals, inv.to(torch.int32)

    def push_grad_scale(self, batch_size: int) -> float:
        """Factor applied to the embedding gradients of a push: ``-batch_size`` (the loss is a batch mean, the
        sparse optimizer wants the per-instance sum, sign flipped for "descent"; PushCopy, box_wrapper.cu:344-416)
        or ``-1`` when the trainer desc says ``scale_sparse_gradient_with_batch_size: false``."""
        return -float(batch_size) if getattr(self, "scale_sparse_gradient_with_batch_size", True) else -1.0

    def push_sparse_grad(self, keys: torch.Tensor, restore: torch.Tensor, grads: torch.Tensor,
                         key2slot: torch.Tensor, batch_size: int) -> None:
        """Per-occurrence gradients [N, hidden] (show, clk counts in the first two columns)."""
        if self._test_mode:
            return
        from .ops import reference as R
        if self.router is not None:
            n_unique = self.router._last[0].numel()
            self.router.push(R.merge_push(restore, n_unique, grads, key2slot, 2, self.push_grad_scale(batch_size)))
            return
        if self.use_cuda:
            if FLAGS.pbx_deterministic_merge and self.world_size == 1:
                # sorted, fp64 segmented merge (reference PushMergeCopy, box_wrapper.cu:417-475), then the plain
                # by-key table update: bit-reproducible whatever the key skew
                n_rows = self.engine.n_unique_rows
                merged = R.merge_push_sorted(restore, n_rows, grads, key2slot, 2, self.push_grad_scale(batch_size))
                uk = torch.zeros(n_rows, dtype=torch.int64, device=keys.device)
                uk[restore.long()] = keys[:restore.numel()]
                used = merged[:, 1] > 0                       # rows that received a show count
                self.table.update(uk[used], merged[used].contiguous())
                return
            self.engine.push_keys(restore, grads.contiguous(), key2slot, 2, self.push_grad_scale(batch_size))
            return
        uk = self._cpu_unique
        merge = R.merge_push_sorted if FLAGS.pbx_deterministic_merge else R.merge_push
        merged = merge(restore, uk.numel(), grads, key2slot, 2, self.push_grad_scale(batch_size))
        self.host_table.push(uk.numpy(), merged)

    # ---- model save / load -------------------------------------------------------------------
    def _sync_to_host(self) -> None:
        if self.use_cuda and self._in_pass:
            torch.cuda.synchronize(self.device)
            dk, dr = self.table.dump(only_dirty=True, clear_dirty=True)
            if dk.numel():
                # bounded pinned chunks: D2H of chunk k+1 overlaps the host-table store of chunk k
                self.last_pass_write_back = self.pass_builder(int(dk.numel())).write_back(dk, dr, mark_dirty=True)

    def save_base(self, batch_model_path: str, xbox_model_path: str = "", date: str = "") -> str:
        """Full training checkpoint ("batch model", every optimizer state) + serving base
        ("xbox model")  (box_wrapper.cc:1286-1305)."""
        self._sync_to_host()
        msg_x = ""
        if xbox_model_path:
            n2 = self.host_table.save(xbox_model_path, mode=2, shard_base=self.rank, clear_dirty=False)
            msg_x = f" xbox_base={n2}"
        # the full save goes last and resets the delta baseline (every row is written, every dirty flag cleared);
        # no scratch files: several ranks save into one directory at the same time
        n1 = self.host_table.save(batch_model_path, mode=0, shard_base=self.rank, clear_dirty=2)
        return f"batch_model={n1}" + msg_x

    def save_delta(self, xbox_model_path: str) -> str:
        """Serving-format delta of the keys touched since the last save (box_wrapper.cc:1306-1315)."""
        self._sync_to_host()
        n = self.host_table.save(xbox_model_path, mode=1, shard_base=self.rank, clear_dirty=True)
        return f"xbox_delta={n}"

    def load_model(self, path: str, serving: bool = False) -> int:
        """Load a saved model into the host tier.  With several ranks every rank reads all part files and
        keeps the keys it owns, so the number of ranks may differ from the run that saved it."""
        return self.host_table.load(path, serving=serving, rank=self.rank, world=self.world_size)

    def load_ssd2mem(self, date: str = "") -> int:
        """Bring every row of the SSD tier back into host memory (``LoadSSD2Mem(day)``,
        box_wrapper.cc:1316-1324) — called before the day-end ``shrink_table`` / ``save_base``.
        Returns the number of rows loaded (0 without an SSD tier: conf ``ssd_dir``)."""
        if date:
            self.set_date(date)
        return self.host_table.load_ssd2mem() if self._ssd else 0

    def shrink_table(self, decay: float = 0.98, delete_threshold: float = 0.0,
                     delete_after_unseen_days: float = 30.0) -> int:
        self._sync_to_host()
        return self.host_table.shrink(decay, delete_threshold, delete_after_unseen_days)

    def shrink_resource(self) -> None:
        if self.use_cuda:
            torch.cuda.empty_cache()

    def merge_model(self, path: str) -> int:
        """Merge another checkpoint into the host tier (``MergeModel``): loaded rows replace
        existing ones."""
        return self.host_table.load(path, rank=self.rank, world=self.world_size)

    def merge_multi_models(self, paths, update_type: int = 0, idx: int = 0) -> int:
        """``MergeMultiModels(path, update_type, idx)``: merge several checkpoints into the host
        tier.  ``update_type`` 0: later models replace earlier rows; 1: show/click of rows present
        in several models are summed (weights from the last); ``idx`` >= 0 with a single string
        path selects ``<path>/<idx>``."""
        import os
        if isinstance(paths, str):
            paths = [os.path.join(paths, str(idx))] if os.path.isdir(os.path.join(paths, str(idx))) else [paths]
        n = 0
        acc = {}
        for p in paths:
            if update_type == 1:
                keys = self.host_table.keys()
                if len(keys):
                    rows = self.host_table.fetch_rows(keys, create=False)
                    acc = {int(k): (float(r[0]), float(r[1])) for k, r in zip(keys, rows)}
            n += self.host_table.load(p, rank=self.rank, world=self.world_size)
            if update_type == 1 and acc:
                keys = self.host_table.keys()
                rows = self.host_table.fetch_rows(keys, create=False).copy()
                for i, k in enumerate(keys.tolist()):
                    if k in acc:
                        rows[i, 0] += acc[k][0]
                        rows[i, 1] += acc[k][1]
                self.host_table.store_rows(keys, rows)
        return n

    def set_test_mode(self, on: bool) -> None:
        self._test_mode = bool(on)
        self.host_table.set_test_mode(on)
        if self.table is not None:
            self.table.set_test_mode(on)

    def check_need_limit_mem(self) -> bool:
        """True when a memory tier is over its budget: the host tier holds more rows than conf
        ``host_mem_rows_limit`` (then ``end_pass`` spills to the SSD tier), or the GPU has less
        than 10 % of its memory free."""
        if self.host_mem_rows_limit and self.host_table is not None \
                and self.host_table.mem_rows() > self.host_mem_rows_limit:
            return True
        if not self.use_cuda:
            return False
        free, total = torch.cuda.mem_get_info(self.device)
        return free < 0.1 * total

    # ---- phases & metrics (box_wrapper.cc:916-1046) --------------------------------------------
    def phase(self) -> int:
        return self._phase

    def flip_phase(self) -> None:
        self._phase = (self._phase + 1) % 2

    def set_phase(self, p: int) -> None:
        self._phase = p

    def init_metric(self, method: str, name: str, label_varname: str = "label", pred_varname: str = "pred",
                    cmatch_rank_varname: str = "", mask_varname: str = "", metric_phase: int = 1,
                    cmatch_rank_group: str = "", ignore_rank: bool = False, bucket_size: int = 1000000,
                    **kw) -> None:
        from .metrics import create_metric
        self._metrics[name] = create_metric(method, name=name, label_varname=label_varname,
                                            pred_varname=pred_varname, cmatch_rank_varname=cmatch_rank_varname,
                                            mask_varname=mask_varname, metric_phase=metric_phase,
                                            cmatch_rank_group=cmatch_rank_group, ignore_rank=ignore_rank,
                                            bucket_size=bucket_size, device=self.device, **kw)

    def get_metric(self, name: str):
        return self._metrics[name]

    def get_metric_name_list(self, metric_phase: int = -1) -> list:
        return [n for n, m in self._metrics.items() if metric_phase < 0 or m.metric_phase == metric_phase]

    def add_metric_data(self, tensors: dict) -> None:
        """Feed every metric bound to the current phase (AddAucMonitor, boxps_worker.cc:1267-1277)."""
        for m in self._metrics.values():
            if m.metric_phase == self._phase:
                m.add_data(tensors)

    def _metric_group(self):
        """The process group metrics are reduced over: this wrapper's group, the default (world) group when none
        was given — ``None`` means "one process, nothing to reduce" to the metric classes, so a multi-rank
        wrapper must never hand them ``None``."""
        if self.world_size <= 1:
            return None
        if self.group is not None:
            return self.group
        import torch.distributed as dist
        return dist.group.WORLD if dist.is_available() and dist.is_initialized() else None

    def get_metric_msg(self, name: str) -> list:
        """[auc, bucket_error, mae, rmse, actual_ctr, predicted_ctr, copc, size]; reduces over
        the data-parallel group and resets the metric (box_wrapper.cc:1027-1046)."""
        return self._metrics[name].get_metric_msg(self._metric_group())

    def get_wuauc_metric_msg(self, name: str) -> list:
        """``[user_cnt, size, sum of user AUCs, sum of (user AUC * instances), uauc, wuauc]`` over all ranks
        (``GetWuAucMetricMsg``, metrics.h:774-805); resets the metric."""
        uauc, wuauc, users, size = self._metrics[name].get_metric_msg(self._metric_group())
        return [users, size, uauc * users, wuauc * size, uauc, wuauc]

    def get_continue_metric_msg(self, name: str) -> list:
        """[mae, rmse, actual value, predicted value, size] of a continuous-value metric
        (``GetContinueMetricMsg``, box_wrapper.cc:1048-1065); resets it."""
        m = self._metrics[name]
        grp = self._metric_group()
        if hasattr(m, "get_continue_msg"):
            return m.get_continue_msg(grp)
        msg = m.get_metric_msg(grp)
        if len(msg) == 5:                       # ContinueMaskCalculator: [mae, rmse, predicted, actual, size]
            return [msg[0], msg[1], msg[3], msg[2], msg[4]]
        return [msg[2], msg[3], msg[4], msg[5], msg[7]]

    def get_nan_inf_metric_msg(self, name: str) -> list:
        return self._metrics[name].get_nan_inf_msg()

    # ---- AUC runner / dataset glue (box_helper_py.cc) --------------------------------------------
    def initialize_auc_runner(self, slot_eval, thread_num: int = 1, pool_size: int = 0, slot_list=()):
        """``InitializeAucRunner`` (box_wrapper.cc:212-368): slot groups whose feasigns are shuffled
        across instances to measure each group's contribution.  Returns the runner
        (``auc_runner.AucRunner``), also kept as ``self.auc_runner``."""
        from .auc_runner import AucRunner
        FLAGS.padbox_auc_runner_mode = True
        self.auc_runner = AucRunner(self, slot_eval, thread_num=thread_num, pool_size=pool_size, slot_list=slot_list)
        return self.auc_runner

    def set_dataset_name(self, name: str) -> None:
        self.dataset_name = name

    def set_input_table_dim(self, dim: int) -> None:
        from .side_tables import InputTable
        self.input_table = InputTable(int(dim))

    def init_afs_api(self, fs_name: str, fs_ugi: str, conf_path: str = "") -> bool:
        """``InitAfsAPI``: remote files are then reached through ``self.file_mgr`` (``BoxFileMgr``)."""
        from .side_tables import BoxFileMgr
        self.file_mgr = BoxFileMgr()
        return self.file_mgr.init(fs_name, fs_ugi, conf_path)

    def release_pool(self) -> None:
        """``ReleasePool``: give cached host / device blocks back (record pools are freed by
        ``dataset.release_memory``)."""
        import gc
        gc.collect()
        if self.use_cuda:
            torch.cuda.empty_cache()

    # ---- the accessors ops and trainers of the reference read off the wrapper (box_wrapper.h:600-900) --------
    def get_gpu_num(self) -> int:
        """GPUs of this node that take part in the job (``GetGpuNum``)."""
        import os
        return max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1"))) if self.world_size > 1 else \
            (1 if self.use_cuda else 0)

    get_device_count = get_gpu_num

    def get_mpi_rank(self) -> int:
        """Rank of this NODE (``GetMpiRank``): the reference has one process per node."""
        return self.rank // max(1, self.get_gpu_num()) if self.world_size > 1 else 0

    def get_mpi_size(self) -> int:
        return max(1, self.world_size // max(1, self.get_gpu_num()))

    def get_nccl_rank_id(self, device_id: int | None = None) -> int:
        """``mpi_rank * gpu_num + device_id`` (box_wrapper.h:448-450) = this process's rank."""
        return self.rank if device_id is None else self.get_mpi_rank() * max(1, self.get_gpu_num()) + int(device_id)

    def get_device_id(self) -> int:
        return self.device.index if self.use_cuda and self.device.index is not None else 0

    get_place_device_id = get_device_id

    def get_round_id(self) -> int:
        """Number of passes begun so far (``GetRoundId``)."""
        cnt = getattr(self.timers, "count", None)
        return int(cnt.get("begin_pass", 0)) if isinstance(cnt, dict) else int(getattr(self, "_round_id", 0))

    def get_dataset_id(self) -> int:
        return getattr(self, "_dataset_id", 0)

    def get_slot_vector(self) -> list:
        return list(getattr(self, "slot_vector", []))

    def get_omited_slot(self) -> set:            # (sic)
        return set(getattr(self, "slot_omit_in_feedpass", ()))

    def get_expand_embed_dim(self) -> int:
        return int(self.expand_embed_dim)

    def get_feedpass_thread_num(self) -> int:
        return int(FLAGS.pbx_pass_fetch_threads)

    def gpu_mem_used(self) -> int:
        """Bytes of HBM in use on this process's device (``GpuMemUsed``)."""
        if not self.use_cuda:
            return 0
        free, total = torch.cuda.mem_get_info(self.device)
        return int(total - free)

    def use_afs_api(self) -> bool:
        return getattr(self, "file_mgr", None) is not None

    def get_file_mgr(self):
        if getattr(self, "file_mgr", None) is None:
            from .side_tables import BoxFileMgr
            self.file_mgr = BoxFileMgr()
        return self.file_mgr

    def phase_num(self) -> int:
        """2 when metrics are registered for both phases (join / update), else 1 (``PhaseNum``)."""
        return 2 if len({m.metric_phase for m in self._metrics.values()}) > 1 else 1

    def mode(self) -> int:
        """1 in test mode, else 0 (``Mode``)."""
        return int(bool(self._test_mode))

    def get_q_tensor(self):
        """The q-values stored by ``store_q_value`` in this step (``GetQTensor``)."""
        from .ops.functional import stored_q_values
        return stored_q_values()

    def sync_dense(self, tensor, scale: float | None = None):
        """``SyncDense``: the inter-node leg of the dense allreduce on ``tensor`` (the whole allreduce on CPU
        worlds); sum over ranks, scaled by ``scale`` (default 1 / world)."""
        if self.world_size == 1:
            return tensor
        import torch.distributed as dist
        dist.all_reduce(tensor, group=self.group)
        tensor.mul_(1.0 / self.world_size if scale is None else scale)
        return tensor

    @staticmethod
    def execute_func(n: int, fn) -> None:
        """``ExecuteFunc``: ``fn(i)`` for i < n over the host pool."""
        from .parallel.cluster import execute_func
        execute_func(n, fn)

    @staticmethod
    def exec_range_func(n: int, fn, parts: int | None = None) -> None:
        from .parallel.cluster import exec_range_func
        exec_range_func(n, fn, parts)

    def print_device_info(self) -> str:
        """``PrintDeviceInfo``: one line per visible device with its memory use."""
        lines = []
        if self.use_cuda:
            for d in range(torch.cuda.device_count()):
                free, total = torch.cuda.mem_get_info(d)
                lines.append(f"device {d}: {torch.cuda.get_device_name(d)} used {(total - free) >> 20} MiB of "
                             f"{total >> 20} MiB")
        else:
            lines.append("device cpu")
        if self.host_table is not None:
            lines.append(f"host table rows {self.host_table.mem_rows()} (+ {self.host_table.ssd_rows()} on SSD)")
        msg = "\n".join(lines)
        print(msg)
        return msg

    # ---- misc ---------------------------------------------------------------------------------
    def print_sync_timer(self) -> str:
        """Per-pass stage timers + memory stats (PrintSyncTimer, box_wrapper.cc:1085-1138)."""
        msg = self.timers.report()
        if self.use_cuda:
            msg += f" gpu_mem_used={torch.cuda.memory_allocated(self.device) >> 20}MiB"
        self.timers.reset()
        return msg

    def finalize(self) -> None:
        self.engine = None
        self.router = None
        self.table = None
        self._pass_builder = None
        self.host_table = None
        self._metrics.clear()
        self._initialized = False
