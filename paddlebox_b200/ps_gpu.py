"""The open twin of the PS front end: ``PSGPUWrapper`` with its pass-build pipeline.

The reference ships two front ends for the GPU parameter server: the closed-library
``BoxWrapper`` and ``PSGPUWrapper`` / ``pull_gpups_sparse`` / ``PSGPUTrainer``
(fleet/ps_gpu_wrapper.h:1-728, operators/pull_gpups_sparse_op.h, ps_gpu_trainer.cc:37-420).
Both drive the same sharded HBM table and sparse engine here; what ``PSGPUWrapper`` adds is its own pass
lifecycle — a background build thread fed through three channels
(``data_ready -> buildcpu_ready -> gpu_free``, fleet/ps_gpu_wrapper.cc:885-955):

    load_into_memory(dataset)      dataset loaded, a ``HeterContext`` is queued            (:885-900)
      pre-build thread             PreBuildTask: the pass's unique keys, sharded by owner   (:114-335)
                                   BuildPull: the rows of those keys fetched / created in
                                   the CPU table, into the task's host pool                 (:337-682)
    begin_pass()                   waits for the built task AND for the GPU to be free
                                   (previous pass ended), then BuildGPUTask: host pool ->
                                   HBM, streamed                                             (:684-883, :957-981)
    end_pass()                     HBM -> CPU table, GPU handed back                        (:983-1079)

so the CPU half of pass N+1 (key collection, CPU-table fetch) overlaps the training of pass N.
"""
from __future__ import annotations

import queue
import threading
import time

import numpy as np

from .box_wrapper import BoxWrapper
from .ops.pull_box_sparse import pull_box_sparse, sparse_embedding_pool
from .trainer import CTRTrainer, Executor, TrainerConfig


class HeterContext:
    """One pass's build task (reference fleet/heter_context.h:41-250): the dataset it belongs to, the unique
    keys this rank owns, the rows staged for them on the host, and the time every stage took."""

    def __init__(self, dataset=None, keys=None, date: str | None = None):
        self.dataset, self.date = dataset, date
        self.raw_keys = keys                  # explicit key list (tests / custom feeds) instead of a dataset
        self.keys: np.ndarray | None = None   # PreBuildTask result: unique, owned by this rank, sorted
        self.rows: np.ndarray | None = None   # BuildPull result: [n, stride] host pool of the pass
        self.timing: dict = {}
        self.error: BaseException | None = None


class PSGPUWrapper(BoxWrapper):
    """``PSGPUWrapper.get_instance()`` — the object model of :class:`BoxWrapper` plus the HeterPS pass
    pipeline described in the module docstring."""

    # ---- channels and the build thread (start_build_thread, ps_gpu_wrapper.cc:885-955) -----------------
    def start_build_thread(self) -> None:
        if getattr(self, "_build_thread", None) is not None:
            return
        self._data_ready: "queue.Queue" = queue.Queue(maxsize=3)
        self._buildcpu_ready: "queue.Queue" = queue.Queue(maxsize=3)
        self._gpu_free: "queue.Queue" = queue.Queue(maxsize=1)
        self._gpu_free.put(None)              # the GPU starts out free
        self.current_task: HeterContext | None = None
        self._running = True
        self._tasks_queued = 0               # tasks handed to the pre-build thread and not yet taken by begin_pass
        # The pre-build thread exchanges keys between ranks while the main thread trains the previous pass and
        # issues collectives of its own: two threads on one process group would interleave differently on
        # different ranks.  The build thread gets a group of its own (host arrays: gloo), created here, where
        # every rank arrives in program order.
        self._build_group = self.group
        if self.world_size > 1:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                ranks = dist.get_process_group_ranks(self.group) if self.group is not None else None
                self._build_group = dist.new_group(ranks=ranks, backend="gloo")
        self._build_thread = threading.Thread(target=self._pre_build_loop, name="psgpu-pre-build", daemon=True)
        self._build_thread.start()

    def _pre_build_loop(self) -> None:
        while True:
            task = self._data_ready.get()
            if task is None:
                return
            try:
                self.pre_build_task(task)
                self.build_pull(task)
            except BaseException as e:  # noqa: BLE001  (handed to begin_pass, which raises it)
                task.error = e
            self._buildcpu_ready.put(task)

    def stop_build_thread(self) -> None:
        if getattr(self, "_build_thread", None) is not None:
            self._data_ready.put(None)
            self._build_thread.join(timeout=30)
            self._build_thread = None

    # ---- the three build stages ---------------------------------------------------------------------------
    def pre_build_task(self, task: HeterContext) -> None:
        """PreBuildTask: the unique keys of the pass, each routed to the rank that owns its shard."""
        t0 = time.time()
        from .host import partition_by_owner, sort_unique_u64
        raw = task.raw_keys if task.raw_keys is not None else task.dataset.unique_keys()
        raw = np.ascontiguousarray(raw.cpu().numpy() if hasattr(raw, "cpu") else raw)
        if raw.dtype.kind not in "iu" or raw.dtype.itemsize != 8:
            raise TypeError(f"pass keys must be 64-bit integers, got {raw.dtype}")
        raw = (raw.view(np.uint64) if raw.dtype != np.uint64 else raw).reshape(-1)
        keys = sort_unique_u64(np.concatenate([np.zeros(1, np.uint64), raw]))       # key 0 is always registered
        if self.world_size > 1:
            from .parallel.cluster import Cluster
            mine = Cluster(getattr(self, "_build_group", self.group)).all_to_all_arrays(
                partition_by_owner(keys, self.world_size))
            keys = sort_unique_u64(np.concatenate(mine)) if mine else np.zeros(0, np.uint64)
        task.keys = keys
        task.timing["pre_build_s"] = time.time() - t0

    def build_pull(self, task: HeterContext) -> None:
        """BuildPull: fetch (or create) the rows of the pass's keys in the CPU table into the task's host pool."""
        t0 = time.time()
        task.rows = self.host_table.fetch_rows(task.keys, create=True) if task.keys.size else \
            np.zeros((0, self.host_table.layout.stride), np.float32)
        task.timing["build_pull_s"] = time.time() - t0

    def build_gpu_task(self, task: HeterContext) -> None:
        """BuildGPUTask: the task's host pool -> this rank's HBM shard (streamed through the staging pool)."""
        t0 = time.time()
        if self.use_cuda:
            self.table.clear()
            if task.keys.size:
                pool = _PrefetchedRows(task.keys, task.rows, self.host_table)
                from .pass_build import PassBuilder
                pb = PassBuilder(pool, self.table.insert, self.host_table.layout.stride, self.device,
                                 chunk=min(int(1 << 19), max(int(task.keys.size), 1)), depth=2, fetch_threads=1)
                task.timing["gpu_build"] = pb.build(task.keys)
        self._pass_keys = task.keys
        self._in_pass = True
        task.timing["build_gpu_s"] = time.time() - t0

    # ---- the reference's entry points ------------------------------------------------------------------------
    def load_into_memory(self, dataset, date: str | None = None) -> HeterContext:
        """``LoadIntoMemory``: load the dataset, then queue its build task for the pre-build thread."""
        self.start_build_thread()
        if date:
            self.set_date(date)
        dataset.load_into_memory()
        return self.submit(HeterContext(dataset=dataset, date=date))

    def submit(self, task: HeterContext) -> HeterContext:
        """Hand a build task to the pre-build thread (``data_ready`` channel); ``begin_pass`` takes the built
        tasks in this order."""
        self.start_build_thread()
        self._tasks_queued += 1
        self._data_ready.put(task)
        return task

    def build_task(self, keys=None) -> HeterContext:
        """Queue a build task for an explicit key list (no dataset) and start the pass."""
        self.start_build_thread()
        task = self.submit(HeterContext(keys=np.zeros(0, np.uint64) if keys is None else keys))
        self.begin_pass()
        return task

    def begin_pass(self) -> None:
        """``BeginPass``: take the next built task, wait until the GPU is free, build the HBM table."""
        if getattr(self, "_build_thread", None) is None:         # used as a plain BoxWrapper
            return super().begin_pass()
        assert self._initialized, "call init_gpu_ps / initialize_gpu_and_load_model first"
        t0 = time.time()
        if self._tasks_queued <= 0:
            raise RuntimeError("begin_pass: no pass is being built (call load_into_memory(dataset) or build_task(keys) first)")
        self._tasks_queued -= 1
        task = self._buildcpu_ready.get()
        if task.error is not None:
            raise task.error
        self._gpu_free.get()                                     # previous pass must have ended
        try:
            self.build_gpu_task(task)
        except BaseException:
            self._gpu_free.put(None)                             # a failed build must not leave the GPU "taken"
            raise
        self.current_task = task
        self.timers.add("begin_pass", time.time() - t0)

    def end_pass(self, need_save_delta: bool = False, delta_path: str = "") -> None:
        """``EndPass``: HBM -> CPU table, then the GPU is free for the next built task."""
        had_task = getattr(self, "current_task", None) is not None
        super().end_pass(need_save_delta, delta_path)
        if had_task:
            self.current_task = None
            self._gpu_free.put(None)

    def finalize(self) -> None:
        self.stop_build_thread()
        super().finalize()

    def set_slot_vector(self, slots) -> None:
        self.slot_vector = list(slots)

    def set_slot_dim_vector(self, dims) -> None:
        """Per-slot embedx widths (dynamic-mf mode).  With more than one distinct width, build a
        ``multi_dim.MultiDimBox(self.slot_dim_vector)`` — one table per width."""
        self.slot_dim_vector = list(dims)

    def set_slot_offset_vector(self, offsets) -> None:
        self.slot_offset_vector = list(offsets)

    def set_dataset(self, dataset) -> None:
        self.dataset = dataset

    def init(self, *a, **kw) -> None:
        """``PSGPUWrapper::init``: nothing to pre-create here (tables are built by ``init_gpu_ps``)."""

    def init_gpu_ps(self, gpus=None, conf: dict | None = None, **kw) -> None:
        """``InitializeGPU(dev_ids)``: this process drives the current device only (one process per GPU)."""
        if not self._initialized:
            self.initialize_gpu_and_load_model(conf or {}, **kw)

    # ---- AFS / HDFS client names of ps_gpu_wrapper_py.cc (delegating to BoxFileMgr) -----------------
    def _fm(self):
        fm = getattr(self, "file_mgr", None)
        if fm is None:
            from .side_tables import BoxFileMgr
            fm = self.file_mgr = BoxFileMgr()
        return fm

    def list(self, path: str) -> list:
        return self._fm().list_dir(path)

    def exist(self, path: str) -> bool:
        return self._fm().exists(path)

    def mkdir(self, path: str) -> bool:
        return self._fm().makedir(path)

    def remove(self, path: str) -> bool:
        return self._fm().remove(path)

    def mv(self, src: str, dst: str) -> bool:
        return self._fm().rename(src, dst)

    def touchz(self, path: str) -> bool:
        return self._fm().touch(path)

    def upload(self, local: str, remote: str) -> bool:
        return self._fm().upload(local, remote)

    def download(self, local: str, remote: str) -> bool:
        return self._fm().download(remote, local)

    def cat(self, path: str) -> str:
        fm = self._fm()
        if fm._remote(path):
            return fm._run("-cat", path).stdout
        with open(path) as f:
            return f.read()

    # SetSparseSGD / SetEmbedxSGD (fleet/ps_gpu_wrapper.cu:139-187): the two halves of the sparse
    # optimizer configuration, applied to the host tier and the HBM shard
    def set_sparse_sgd(self, nonclk_coeff=0.1, clk_coeff=1.0, min_bound=-10.0, max_bound=10.0, learning_rate=0.05,
                       initial_g2sum=3.0, initial_range=0.0, beta1_decay_rate=0.9, beta2_decay_rate=0.999,
                       ada_epsilon=1e-8) -> None:
        PSGPUWrapper._update_sparse_config(self, nonclk_coeff=nonclk_coeff, clk_coeff=clk_coeff, min_bound=min_bound,
                                   max_bound=max_bound, learning_rate=learning_rate, initial_g2sum=initial_g2sum,
                                   initial_range=initial_range, beta1_decay_rate=beta1_decay_rate,
                                   beta2_decay_rate=beta2_decay_rate, ada_epsilon=ada_epsilon)

    def set_embedx_sgd(self, mf_create_thresholds=10.0, mf_learning_rate=0.05, mf_initial_g2sum=3.0,
                       mf_initial_range=1e-4, mf_min_bound=-10.0, mf_max_bound=10.0, mf_beta1_decay_rate=0.9,
                       mf_beta2_decay_rate=0.999, mf_ada_epsilon=1e-8) -> None:
        PSGPUWrapper._update_sparse_config(self, mf_create_thresholds=mf_create_thresholds, mf_learning_rate=mf_learning_rate,
                                   mf_initial_g2sum=mf_initial_g2sum, mf_initial_range=mf_initial_range,
                                   mf_min_bound=mf_min_bound, mf_max_bound=mf_max_bound,
                                   mf_beta1_decay_rate=mf_beta1_decay_rate, mf_beta2_decay_rate=mf_beta2_decay_rate,
                                   mf_ada_epsilon=mf_ada_epsilon)

    def _update_sparse_config(self, **kw) -> None:
        from . import _native as N
        cur = self.host_table.config.as_dict() if getattr(self, "host_table", None) is not None else {}
        cur.update(kw)
        cfg = N.SparseOptConfig(**cur)
        if getattr(self, "host_table", None) is not None:
            self.host_table.set_config(cfg)
        if getattr(self, "table", None) is not None:
            self.table.set_config(cfg)

    def init_gpu_server(self, conf: "dict | str | None" = None, **kw) -> None:
        """``InitializeGPUServer``: sparse-SGD config from the fleet desc (ps_gpu_trainer.cc:137-254).  ``conf``:
        a dict of ``SparseOptConfig`` names, or the job's ``fleet_desc`` protobuf text (``fleet_desc.py`` reads the
        first sparse table's accessor out of it, as the reference does)."""
        if isinstance(conf, str) and "downpour_table_param" in conf:
            from .fleet_desc import sparse_config_from_fleet_desc
            c, optimizer, _ = sparse_config_from_fleet_desc(conf)
            conf = {**c, "optimizer": optimizer}
        self.initialize_gpu_and_load_model(conf or {}, **kw)


class _PrefetchedRows:
    """Host-table face of a task's already fetched pool (BuildPull ran in the background): chunk lookups
    are slices of the pool, writes go to the real host table."""

    def __init__(self, keys: np.ndarray, rows: np.ndarray, host_table):
        self.keys, self.rows, self.host_table = keys, rows, host_table
        self.layout = host_table.layout

    def fetch_rows_into(self, keys: np.ndarray, out: np.ndarray, create: bool = True) -> None:
        i0 = int(np.searchsorted(self.keys, keys[0])) if keys.size else 0
        out[:] = self.rows[i0:i0 + keys.size]                   # chunks are contiguous runs of the sorted key list

    def store_rows(self, keys, rows, mark_dirty=True):
        self.host_table.store_rows(keys, rows, mark_dirty=mark_dirty)


def pull_gpups_sparse(batch, size=None, offset: int = 0, box=None):
    """``_pull_gpups_sparse`` (python/paddle/fluid/layers/nn.py): HeterPS twin of ``pull_box_sparse``."""
    return pull_box_sparse(batch, size=size, offset=offset, box=box)


class PSGPUWorker(CTRTrainer):
    """Per-GPU worker of the HeterPS flavour (ps_gpu_worker.cc:36-310): ``TrainFiles`` = one pass over the
    batches the dataset dealt to this worker, on the sparse table the wrapper built for the pass."""

    def train_files(self, dataset, **kw) -> dict:
        return self.train_from_dataset(dataset, **kw)


class PSGPUTrainer:
    """``PSGPUTrainer`` (ps_gpu_trainer.cc:37-420): ``Initialize`` configures the GPU server from the fleet
    desc (``InitializeGPUServer`` :137-254), ``Run`` trains one pass through the wrapper's pipeline
    (``BeginPass -> TrainFiles -> EndPass``), ``Finalize`` merges the dense parameters back
    (``MergeDenseParam`` :359-369 — with one process per GPU every rank already holds the synchronised
    dense parameters, so this returns them)."""

    def __init__(self, model, wrapper: PSGPUWrapper | None = None, config: TrainerConfig | None = None,
                 sparse_conf: dict | None = None, **init_kw):
        self.wrapper = wrapper or PSGPUWrapper.get_instance()
        if not self.wrapper._initialized:
            self.wrapper.init_gpu_server(sparse_conf or {}, **init_kw)
        self.worker = PSGPUWorker(model, self.wrapper, config or TrainerConfig())

    def run(self, dataset, date: str | None = None, need_save_delta: bool = False, delta_path: str = "") -> dict:
        self.wrapper.load_into_memory(dataset, date)
        self.wrapper.begin_pass()
        try:
            stats = self.worker.train_files(dataset)
        finally:
            self.wrapper.end_pass(need_save_delta, delta_path)
        stats["build"] = dict(self.wrapper.timers.summary()) if hasattr(self.wrapper.timers, "summary") else {}
        return stats

    def finalize(self) -> dict:
        self.worker.finalize()
        named = self.worker.model.named_dense_params() if hasattr(self.worker.model, "named_dense_params") \
            else self.worker.model.named_parameters()
        return {n: p.detach().cpu() for n, p in named}


__all__ = ["PSGPUWrapper", "pull_gpups_sparse", "PSGPUTrainer", "PSGPUWorker", "sparse_embedding_pool", "Executor",
           "TrainerConfig"]
