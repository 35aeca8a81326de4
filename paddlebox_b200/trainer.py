"""Training runtime: one process per GPU, one ``CTRTrainer`` per process.

Reference roles: ``Executor.train_from_dataset`` (python/paddle/fluid/executor.py:1787-1933),
``BoxPSTrainer`` (framework/boxps_trainer.cc: one worker thread per GPU) and ``BoxPSWorker``
(framework/boxps_worker.cc:1278-1482: per-batch op loop, dense sync, NaN/Inf check, AUC hook,
dump, per-op profiler).  The op-list interpreter, scopes and eager GC of the reference are
replaced by PyTorch autograd + the CUDA caching allocator; what remains is the per-batch
loop and everything wrapped around it.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field

import torch

from .box_wrapper import BoxWrapper
from . import metrics as _metrics
from .flags import FLAGS
from .parallel.dense import DenseOptimizer, DenseParamStore, DenseSync
from .utils.nan_inf import check_batch_nan_inf
from .utils.timer import CudaStageTimers


@dataclass
class TrainerConfig:
    """Mirror of ``BoxPSWorkerParameter`` / ``TrainerDesc`` fields the reference exposes
    (trainer_desc.proto:78, :121-129; device_worker.py:623-655)."""
    dense_lr: float = 1e-3
    dense_optimizer: str = "adam"
    sync_dense_mode: str = "allreduce"     # allreduce | kstep | sharding | async | none
    sync_weight_step: int = 1
    async_mode: bool = False
    debug: bool = False                    # TrainFilesWithProfiler: per-stage device timing
    check_nan_inf: bool | None = None
    dump_fields: list = field(default_factory=list)
    dump_fields_path: str = ""
    dump_param: list = field(default_factory=list)
    dump_mode: int = 0
    dump_interval: int = 10
    dump_thread_num: int = 1
    is_test: bool = False
    user_define_dump_filename: str = ""    # dump files are named part-<this>-<n> instead of part-<rank>-<n>
    check_nan_var_names: list = field(default_factory=list)   # NaN / Inf check restricted to these outputs
    loss_names: list = field(default_factory=list)
    scale_sparse_gradient_with_batch_size: bool = True        # push scale -batch_size (True) or -1

    @classmethod
    def from_desc(cls, text: str) -> "TrainerConfig":
        """From the protobuf text of a ``TrainerDesc`` (trainer_desc.proto:20-75, ``boxps_param`` :121-129): the
        fields that mean something here — debug, dump_* (``enable_random_dump`` / ``random_with_lineid`` /
        ``dump_interval`` choose the dump mode), ``check_nan_var_names``, ``loss_names``,
        ``scale_sparse_gradient_with_batch_size``, and the BoxPS worker's ``async_mode`` / ``sync_dense_mode`` /
        ``sync_weight_step`` / ``dump_thread_num``."""
        from .fleet_desc import parse_text_format
        d = parse_text_format(text)

        def lst(v):
            return [] if v is None else (list(v) if isinstance(v, list) else [v])
        cfg = cls()
        cfg.debug = bool(d.get("debug", False))
        cfg.dump_fields = [str(x) for x in lst(d.get("dump_fields"))]
        cfg.dump_param = [str(x) for x in lst(d.get("dump_param"))]
        cfg.dump_fields_path = str(d.get("dump_fields_path", ""))
        cfg.user_define_dump_filename = str(d.get("user_define_dump_filename", ""))
        cfg.check_nan_var_names = [str(x) for x in lst(d.get("check_nan_var_names"))]
        cfg.loss_names = [str(x) for x in lst(d.get("loss_names"))]
        cfg.scale_sparse_gradient_with_batch_size = bool(d.get("scale_sparse_gradient_with_batch_size", True))
        if bool(d.get("enable_random_dump", False)):
            cfg.dump_mode = 1 if bool(d.get("random_with_lineid", False)) else 2
            cfg.dump_interval = int(d.get("dump_interval", 10000))
        bp = d.get("boxps_param") or {}
        cfg.async_mode = bool(bp.get("async_mode", False))
        cfg.sync_weight_step = int(bp.get("sync_weight_step", 1))
        cfg.dump_thread_num = int(bp.get("dump_thread_num", 1))
        mode = bp.get("sync_dense_mode", 0)
        cfg.sync_dense_mode = "async" if cfg.async_mode else \
            {0: "allreduce", 1: "kstep", 2: "allreduce"}.get(int(mode), "allreduce") if not isinstance(mode, str) else mode
        if cfg.sync_weight_step > 1 and cfg.sync_dense_mode == "allreduce":
            cfg.sync_dense_mode = "kstep"
        return cfg


class CTRTrainer:
    def __init__(self, model: torch.nn.Module, box: BoxWrapper | None = None, config: TrainerConfig | None = None,
                 rank: int | None = None, world_size: int | None = None, group=None, topo=None):
        """``topo``: a ``parallel.topology.NodeTopology`` when the job spans several nodes — the dense
        sync then runs reduce-scatter inside the node, the inter-node leg on each rank's shard and
        all-gather inside the node (boxps_worker.cc:1220-1233)."""
        self.model = model
        self.box = box or BoxWrapper.get_instance()
        self.cfg = config or TrainerConfig()
        self.rank = self.box.rank if rank is None else rank
        self.world = self.box.world_size if world_size is None else world_size
        mode = "async" if self.cfg.async_mode else self.cfg.sync_dense_mode
        if FLAGS.padbox_enable_sharding_stage:
            mode = "sharding"
        self.store = DenseParamStore(model, self.rank, self.world, group, topo=topo)
        self.opt = DenseOptimizer(self.store, lr=self.cfg.dense_lr, kind=self.cfg.dense_optimizer)
        self.sync = DenseSync(self.store, self.opt, mode, self.cfg.sync_weight_step)
        if mode == "async" and self.box.lr_map:
            # per-parameter learning rates of the async dense table (boxps_worker.cc:199-207)
            lrs = self.box.get_lr_map()
            named = list(model.named_dense_params()) if hasattr(model, "named_dense_params") else \
                list(model.named_parameters())
            offs = {id(p): o for p, o in zip(self.store.params, self.store.offsets)}
            for n, p in named:
                if n in lrs and id(p) in offs:
                    self.sync.async_table.set_lr_range(offs[id(p)], offs[id(p)] + p.numel(), lrs[n])
        # The folded allreduce runs inside the sparse push kernel, i.e. when the pull op's backward
        # runs: every dense gradient must be final by then.  That holds when the pull is the first
        # op of the forward and nothing dense is computed beside it — a property of the MODEL, so the
        # model has to declare it (CTRModelBase does); any other module keeps the separate allreduce.
        if self.world > 1 and not self.cfg.is_test and getattr(model, "dense_grads_final_before_push", False):
            self.sync.fold_into(getattr(self.box, "engine", None))
        if self.world > 1:
            self.store.broadcast_params(0)
            self.store.mark_dirty()
        self.pinned_cores = self._bind_train_cpu()
        if FLAGS.enable_dump_main_program:
            self.dump_main_program()
        self.timers = CudaStageTimers(self.cfg.debug)
        self.dumper = None
        if self.cfg.dump_fields_path and (self.cfg.dump_fields or self.cfg.dump_param):
            from .utils.dump import FieldDumper
            self.dumper = FieldDumper(self.cfg.dump_fields_path, self.cfg.dump_fields, self.cfg.dump_mode,
                                      self.cfg.dump_interval, self.cfg.dump_thread_num, rank=self.rank,
                                      user_filename=self.cfg.user_define_dump_filename)
        self.box.scale_sparse_gradient_with_batch_size = bool(self.cfg.scale_sparse_gradient_with_batch_size)
        self.steps = 0
        self.last_outputs = None
        # data_norm(sync_stats): on GPUs of one node the three statistics vectors are reduced over the
        # ranks and the summaries updated by ONE kernel over symmetric memory (pbx_symm_stat_update);
        # elsewhere (CPU worlds, several nodes, FLAGS.dense_use_nccl) a torch.distributed allreduce
        synced = [m for m in model.modules() if getattr(m, "sync_stats", False)]
        self._stat_symm = None
        if synced and self.world > 1:
            fused = self.store.use_symm and self.store.topo is None
            if fused:
                from .parallel.symm import SymmetricBuffer
                width = max(int(m.batch_size.numel()) for m in synced)
                self._stat_symm = SymmetricBuffer(3 * width + 8, self.rank, self.world, group=self.store.group)
            for m in synced:
                if fused:
                    m.stat_sync = self._sync_stats_fused
                else:
                    m.reduce_fn = self._reduce_stats

    def dump_main_program(self, path: str | None = None) -> str:
        """``FLAGS_enable_dump_main_program``: the reference writes every device's op list to
        ``./device_<id>_ops.txt`` (boxps_worker.cc:860-880).  There is no op list here — the step is autograd plus
        one CUDA graph — so the file holds what plays that role: the module tree, every dense parameter with its
        checkpoint name, shape and offset in the flat buffer, the sparse table and the sync configuration."""
        path = path or f"./device_{self.rank}_ops.txt"
        named = dict((id(p), n) for n, p in (self.model.named_dense_params() if hasattr(self.model, "named_dense_params")
                                             else self.model.named_parameters()))
        lines = [f"# device {self.rank} of {self.world}: dense sync {self.sync.mode}, optimizer {self.opt.kind} lr {self.opt.lr}",
                 f"# sparse table: embedx_dim {self.box.embedx_dim}, optimizer {self.box.optimizer}, "
                 f"feature_type {self.box.feature_type}, hidden {self.box.hidden}", "", "[modules]"]
        for name, mod in self.model.named_modules():
            lines.append(f"{name or '<root>'}: {type(mod).__name__}")
        lines += ["", "[dense parameters]  name shape flat_offset"]
        for p, off in zip(self.store.params, self.store.offsets):
            lines.append(f"{named.get(id(p), '?')} {tuple(p.shape)} {off}")
        lines.append(f"# flat buffer: {self.store.numel} floats, symmetric memory {bool(self.store.use_symm)}")
        with open(path, "w") as f:
            f.write("\n".join(lines) + "\n")
        return path

    def _bind_train_cpu(self) -> list:
        """FLAGS_enable_binding_train_cpu (boxps_trainer.cc:165-193): the training thread of this process is
        pinned to its share of the train cores — the processes of one node split them by local rank.  Only
        with several processes per node; a lone process keeps the scheduler's choice."""
        import os
        from .parallel.cluster import get_train_cores, set_cpu_affinity
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
        if not FLAGS.enable_binding_train_cpu or local_world <= 1:
            return []
        cores = get_train_cores()
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        mine = cores[local_rank::local_world] or cores
        return list(mine) if set_cpu_affinity(mine) else []

    def _sync_stats_fused(self, m, d_size, d_sum, d_sq) -> None:
        """summaries = summaries * decay + mean over ranks of the batch statistics, one launch
        (reference: 3 ncclAllReduce + KernelUpdateParam, data_norm_op.cu:203-231)."""
        Cn = int(m.batch_size.numel())
        buf = self._stat_symm.tensor[:3 * Cn].view(3, Cn)
        torch.stack([d_size.reshape(-1), d_sum.reshape(-1), d_sq.reshape(-1)], out=buf)
        self._stat_symm.stat_update(0, Cn, 1.0 / self.world, float(m.decay), m.batch_size, m.batch_sum,
                                    m.batch_square_sum)

    def _reduce_stats(self, stats: torch.Tensor) -> torch.Tensor:
        import torch.distributed as dist
        dist.all_reduce(stats, group=self.store.group)
        return stats / self.world

    # ---- CUDA-graph replay of the whole step -----------------------------------------------
    def enable_cuda_graph(self, example_batch, max_keys: int | None = None, warmup: int = 3) -> bool:
        """Capture forward + backward/push + dense sync + optimizer of one step into a CUDA graph
        and replay it for every later batch of the same (B, S, dense width).

        Variable-length batches work because every buffer has a fixed capacity and the kernels
        read the real key count from the device (``lod[S*B]``).  The ``warmup`` eager steps and
        the capture run are real training steps on ``example_batch``.  Returns False (and stays
        eager) when capture is not possible on this setup."""
        if not example_batch.keys.is_cuda or self.cfg.is_test or self.sync.mode == "async" \
                or not FLAGS.pbx_use_cuda_graph:
            return False
        from .ops.pull_box_sparse import SlotBatch
        B, S = example_batch.batch_size, example_batch.num_slots
        dev = example_batch.keys.device
        cap = int(max_keys or self.box.max_keys_per_batch)
        keys = torch.zeros(cap, dtype=torch.int64, device=dev)
        lod = torch.zeros(S * B + 1, dtype=torch.int32, device=dev)
        dense = torch.zeros_like(example_batch.dense) if example_batch.dense is not None else None
        static = SlotBatch(keys, lod, B, S, dense, n_keys=cap, n_dev=lod[S * B:])
        self._static = static
        self._copy_into_static(example_batch)
        try:
            # Autograd runs a parameter's gradient accumulation on the stream that was current when
            # the parameter was FIRST used; if that was the default stream, a capture on any other
            # stream fails ("dependency created on uncaptured work in another stream").  A caller that
            # trains on a stream of its own from the start (torch.cuda.set_stream) gets warm-up and
            # capture on that same stream.
            cur = torch.cuda.current_stream(dev)
            side = cur if cur != torch.cuda.default_stream(dev) else torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(max(1, warmup)):
                    self._eager_step(static)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self.store.mark_dirty()
            g = torch.cuda.CUDAGraph()
            # thread_local: a loader thread that pins host memory while we capture must not
            # invalidate the capture
            # (captured on the stream the warm-up ran on: tensors cached during warm-up — the root
            #  gradient, workspaces — then carry no dependency on a stream outside the capture)
            with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                loss = self._eager_step(static)
            self._graph, self._graph_loss = g, loss
            torch.cuda.synchronize(dev)
            return True
        except Exception as e:  # capture unsupported (e.g. cooperative launch in capture): stay eager
            import traceback
            self._graph = None
            self._graph_error = repr(e) + "\n" + traceback.format_exc()
            torch.cuda.synchronize(dev)
            return False

    def _copy_into_static(self, batch) -> None:
        st = self._static
        n = batch.N
        if n > st.keys.numel():
            raise ValueError(f"batch has {n} keys; the captured graph was sized for {st.keys.numel()}")
        k, d = batch.keys, batch.dense
        fused = (k.is_cuda and k.is_contiguous() and batch.lod.is_contiguous() and k.data_ptr() % 16 == 0
                 and (d is None or (d.is_contiguous() and st.dense.is_contiguous())))
        if fused:
            # one launch instead of three device-to-device copies
            from . import _native as N
            import ctypes as C
            N.check(N.cuda().pbx_copy_batch(N.ptr(k), N.ptr(st.keys), C.c_int64(n), N.ptr(batch.lod), N.ptr(st.lod),
                                            C.c_int64(st.lod.numel()), N.ptr(d), N.ptr(st.dense) if d is not None else None,
                                            C.c_int64(d.numel() if d is not None else 0), N.stream_ptr()),
                    "pbx_copy_batch")
        else:
            st.keys[:n].copy_(batch.keys[:n], non_blocking=True)
            st.lod.copy_(batch.lod, non_blocking=True)
            if st.dense is not None:
                st.dense.copy_(batch.dense, non_blocking=True)
        st.ins_ids = batch.ins_ids

    # ---- one mini-batch ------------------------------------------------------------------
    def train_step(self, batch) -> torch.Tensor:
        """forward + backward (= sparse push) + dense sync/update.  Returns the loss tensor
        (device; reading it is the caller's sync point)."""
        eng = getattr(self.box, "engine", None)
        if eng is not None:
            eng.peek_status()        # a mailbox overflow / barrier timeout of an earlier step is fatal here
        g = getattr(self, "_graph", None)
        if g is not None and not self.cfg.is_test and batch.batch_size == self._static.batch_size \
                and batch.num_slots == self._static.num_slots:
            self._copy_into_static(batch)
            g.replay()
            self.steps += 1
            return self._graph_loss
        return self._eager_step(batch)

    def _eager_step(self, batch) -> torch.Tensor:
        t = self.timers
        if FLAGS.padbox_enable_print_op_debug:              # one line per stage (boxps_worker.cc:50-52)
            print(f"[rank {self.rank} step {self.steps}] forward B={batch.batch_size} keys={batch.N}", flush=True)
        self.sync.before_forward()
        with t.span("forward"):
            loss, outs = self.model(batch)
        if self.cfg.is_test:
            self._after_batch(batch, loss, outs)
            return loss
        with t.span("backward+push"):
            # a cached root gradient: autograd does not launch a fill kernel for it every step
            one = getattr(self, "_root_grad", None)
            if one is None or one.device != loss.device or one.dtype != loss.dtype or one.shape != loss.shape:
                one = self._root_grad = torch.ones_like(loss)
            torch.autograd.backward(loss, grad_tensors=one)
        with t.span("dense_sync+opt"):
            self.sync.after_backward()
        self._after_batch(batch, loss, outs)
        return loss

    def infer_step(self, batch):
        with torch.no_grad():
            loss, outs = self.model(batch)
        self.box.add_metric_data(outs)
        if _metrics.Metric._instance is not None:
            _metrics.Metric._instance.add_metric_data(outs)
        return loss, outs

    def _after_batch(self, batch, loss, outs):
        self.steps += 1
        self.last_outputs = outs
        check = FLAGS.check_nan_inf if self.cfg.check_nan_inf is None else self.cfg.check_nan_inf
        if check:
            names = set(self.cfg.check_nan_var_names)
            check_batch_nan_inf({"loss": loss, **{k: v for k, v in outs.items()
                                                  if torch.is_tensor(v) and (not names or k in names)}},
                                self.cfg.dump_fields_path or None, tag=f"rank{self.rank}_step{self.steps}")
        self.box.add_metric_data(outs)           # AddAucMonitor
        if _metrics.Metric._instance is not None:
            _metrics.Metric._instance.add_metric_data(outs)
        if self.dumper is not None and self.cfg.dump_fields and batch.ins_ids is not None:
            self.dumper.dump_batch(batch.ins_ids, outs)

    # ---- one pass ---------------------------------------------------------------------------
    def _deal_batches(self, dataset) -> int:
        """Deals the loaded records into batches and returns the worker index to iterate.

        One process per GPU: whatever this process loaded is ITS share of the pass — the files were split
        between the processes before loading, by ``set_rank_world`` (round-robin, data_set.cc:1963-1975) or by
        the job script (``fleet.util.get_file_shard`` / ``split_files``, the usual idiom).  So ALL local records
        are dealt to a single local worker, and the batch count is equalised across ranks with an
        allreduce(max) (data_set.cc:2708-2750) so that every rank enters the in-kernel barriers of pull / push
        the same number of times.  Only a dataset the caller has explicitly dealt to several workers
        (``prepare_train(n > 1, ...)``: several readers of one in-memory dataset) keeps that deal."""
        if dataset._prepared_workers > 1:
            return self.rank if dataset._prepared_workers == self.world else 0
        if self.world == 1:
            if dataset._prepared_workers == 0:
                dataset.prepare_train(1, 0)
            return 0
        from .parallel.cluster import OP_MAX, Cluster
        dataset.prepare_train(1, 0)
        local = dataset.num_batches(0)
        most = int(Cluster.ins().allreduce(int(local), OP_MAX))
        if most > local:
            if local == 0:
                raise RuntimeError(f"rank {self.rank} loaded no trainable batch while another rank has {most}: "
                                   "every rank needs at least batch_size records (give it more files)")
            dataset.prepare_train(1, 0, min_batches=most)
            if dataset.num_batches(0) != most:
                raise RuntimeError(f"rank {self.rank}: cannot split {local} batches into {most} (too few records)")
        return 0

    def _stream_batches(self, dataset, want_ids: bool):
        """Batches of a streaming dataset.  One rank: as they come.  Several ranks: nobody knows its batch count
        in advance, yet every rank must take the same number of steps (pull / push meet at in-kernel barriers,
        the dense sync is a collective), so the ranks agree before every step that all of them still have a
        batch; the pass ends for everyone when the first rank runs dry (the reference leaves equal shares to
        the user: a QueueDataset job with unequal file shares hangs in its first collective)."""
        it = iter(dataset.batches(with_ins_id=want_ids))
        if self.world == 1:
            yield from it
            return
        from .parallel.cluster import OP_MIN, Cluster
        while True:
            b = next(it, None)
            if not Cluster.ins().allreduce(int(b is not None), OP_MIN):
                return
            yield b

    def train_from_dataset(self, dataset, print_period: int = 0, fetch_handler=None) -> dict:
        """Train every batch the dataset dealt to this worker; returns pass statistics."""
        t0 = time.time()
        n_ins, losses = 0, []
        want_ids = self.dumper is not None and bool(self.cfg.dump_fields)
        if getattr(dataset, "streaming", False) and not dataset.is_loaded():
            batches = self._stream_batches(dataset, want_ids)      # QueueDataset: nothing is loaded up front
        else:
            worker = self._deal_batches(dataset)
            batches = dataset.batches(worker=worker, with_ins_id=want_ids)
        for i, batch in enumerate(batches):
            loss = self.train_step(batch)
            n_ins += batch.batch_size
            if print_period and (i + 1) % print_period == 0:
                losses.append(float(loss))
                if fetch_handler:
                    fetch_handler(i + 1, losses[-1])
        self.sync.finalize()
        if self.box.use_cuda:
            torch.cuda.synchronize()
            if self.box.engine is not None:
                self.box.engine.check_status()
        if self.dumper is not None and self.cfg.dump_param:
            named = dict(self.model.named_dense_params()) if hasattr(self.model, "named_dense_params") \
                else dict(self.model.named_parameters())
            self.dumper.dump_params([(n, named[n]) for n in self.cfg.dump_param if n in named])
        dt = time.time() - t0
        stats = {"instances": n_ins, "seconds": dt, "ins_per_sec": n_ins / max(dt, 1e-9), "steps": self.steps,
                 "losses": losses}
        if self.cfg.debug:
            stats["stage_ms"] = self.timers.summary()   # read_time / cal_time style breakdown
        return stats

    def train_host_batches(self, host_batches, on_loss=None, max_keys: int | None = None,
                           after_enqueue=None) -> dict:
        """Train on pinned host batches through the pipelined feed (``paddlebox_b200.feed``):
        H2D of batch i+1 overlaps step i, the batch is built straight into the buffers of the
        captured step graph (when ``enable_cuda_graph`` was called), and every step's loss is
        read back through a pinned ring one step behind the device."""
        from .feed import PipelinedFeed
        feed = getattr(self, "_feed", None)
        if feed is None:
            feed = self._feed = PipelinedFeed(self, max_keys)
        t0 = time.time()
        steps, n_ins, losses = feed.run(host_batches, on_loss, after_enqueue=after_enqueue)
        self.sync.finalize()
        dt = time.time() - t0
        return {"instances": n_ins, "seconds": dt, "ins_per_sec": n_ins / max(dt, 1e-9), "steps": steps,
                "losses": losses, "h2d_bytes": feed.h2d_bytes, "d2h_bytes": feed.d2h_bytes,
                "host_enqueue_s": feed.host_enqueue_s, "host_wait_s": feed.host_wait_s}

    def finalize(self):
        if self.dumper is not None:
            self.dumper.close()
            self.dumper = None


class FetchHandler:
    """``fluid.executor.FetchHandler`` (executor.py:502-527): while ``train_from_dataset`` runs, ``handler`` is
    called every ``period_secs`` with the current value of every variable of ``var_dict``.  A "variable" here is
    the name of an entry of the model's output dict (``"pred"``, ``"loss"`` ...), a callable returning a tensor /
    number, or a tensor; values arrive as numpy arrays (missing ones as ``None``)."""

    def __init__(self, var_dict=None, period_secs: float = 60):
        assert var_dict is not None
        self.var_dict, self.period_secs = var_dict, period_secs

    def handler(self, res_dict: dict) -> None:
        import sys
        for key, v in res_dict.items():
            if hasattr(v, "reshape") and getattr(v, "size", 0):
                sys.stdout.write(f"{key}[0]: {v.reshape(-1)[0]} ")
        sys.stdout.write("\n")

    @staticmethod
    def help() -> None:
        print("class MyHandler(FetchHandler):\n    def handler(self, res_dict):\n        print(res_dict['pred'][:4])\n"
              "exe.train_from_dataset(model, dataset, fetch_handler=MyHandler(var_dict={'pred': 'pred'}, period_secs=30))")


class _FetchHandlerMonitor:
    """The reference's ``FetchHandlerMonitor`` (trainer_factory.py:88-170): a thread that wakes every
    ``period_secs`` and hands the handler a snapshot."""

    def __init__(self, trainer: "CTRTrainer", handler: FetchHandler):
        import threading
        self.trainer, self.handler = trainer, handler
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._loop, name="fetch-handler", daemon=True)

    def snapshot(self) -> dict:
        import numpy as np
        outs = self.trainer.last_outputs or {}
        res = {}
        for name, var in self.handler.var_dict.items():
            v = var() if callable(var) else outs.get(var) if isinstance(var, str) else var
            if torch.is_tensor(v):
                v = v.detach().float().cpu().numpy()
            elif v is not None and not isinstance(v, np.ndarray):
                v = np.asarray(v)
            res[name] = v
        return res

    def _loop(self) -> None:
        while not self._stop.wait(self.handler.period_secs):
            try:
                self.handler.handler(self.snapshot())
            except Exception as e:  # noqa: BLE001  (a failing user handler must not end the training pass)
                print(f"fetch handler failed: {e}")

    def start(self) -> None:
        self._thread.start()

    def stop(self) -> None:
        self._stop.set()
        self._thread.join(timeout=10)


class Executor:
    """``exe.train_from_dataset(program, dataset)`` facade: ``program`` is a model (or a
    prepared ``CTRTrainer``)."""

    def __init__(self, place=None):
        self.place = place
        self._trainers: dict = {}

    def _trainer_for(self, program, **kw) -> CTRTrainer:
        if isinstance(program, CTRTrainer):
            return program
        t = self._trainers.get(id(program))
        if t is None:
            cfg = kw.pop("config", None) or getattr(program, "_pipeline_opt", None)
            fleet_opt = getattr(program, "_fleet_opt", None)
            if isinstance(cfg, dict) or (cfg is None and isinstance(fleet_opt, dict)):
                # the reference's path: program._fleet_opt / _pipeline_opt dict -> TrainerFactory -> TrainerDesc
                # (executor.py _prepare_trainer); the description is turned into this trainer's config
                from .trainer_desc import TrainerFactory
                info = dict(fleet_opt or {})
                info.setdefault("trainer", "BoxPSTrainer" if isinstance(cfg, dict) else "MultiTrainer")
                info.setdefault("device_worker", "BoxPSWorker" if isinstance(cfg, dict) else "Hogwild")
                desc = TrainerFactory()._create_trainer(info)
                desc._set_program(program)
                desc._gen_trainer_desc()
                cfg = desc.to_config()
            t = self._trainers[id(program)] = CTRTrainer(program, config=cfg)
        return t

    def train_from_dataset(self, program, dataset, scope=None, thread=0, debug=False, fetch_list=None,
                           fetch_info=None, print_period=100, fetch_handler=None, **kw):
        tr = self._trainer_for(program, **kw)
        tr.cfg.debug = tr.cfg.debug or debug
        tr.timers.enabled = tr.cfg.debug and torch.cuda.is_available()
        tr.cfg.is_test = False
        if isinstance(fetch_handler, FetchHandler):
            mon = _FetchHandlerMonitor(tr, fetch_handler)
            mon.start()
            try:
                return tr.train_from_dataset(dataset)
            finally:
                mon.stop()
        return tr.train_from_dataset(dataset, print_period if fetch_handler else 0, fetch_handler)

    def infer_from_dataset(self, program, dataset, **kw):
        tr = self._trainer_for(program, **kw)
        tr.cfg.is_test = True
        self_box = tr.box
        self_box.set_test_mode(True)
        try:
            return tr.train_from_dataset(dataset)
        finally:
            self_box.set_test_mode(False)
            tr.cfg.is_test = False

    def run(self, program=None, feed=None, fetch_list=None, return_numpy: bool = True, is_test: bool = False, **kw):
        """``exe.run(program, feed=batch, fetch_list=[...])``: ONE step on one ``SlotBatch`` (training unless
        ``is_test``); returns the fetched outputs in ``fetch_list`` order (``"loss"`` and the names of the model's
        output dict), as numpy arrays by default.  A startup program (``program`` without a forward, or None) is
        a no-op: parameters are initialised at construction."""
        if program is None or feed is None or not (isinstance(program, CTRTrainer) or callable(program)):
            return []
        tr = self._trainer_for(program, **kw)
        if is_test:
            loss, outs = tr.infer_step(feed)
        else:
            loss = tr.train_step(feed)
            outs = tr.last_outputs or {}
        vals = []
        for name in (fetch_list or []):
            v = loss if name == "loss" else outs.get(name)
            if return_numpy and torch.is_tensor(v):
                v = v.detach().float().cpu().numpy()
            vals.append(v)
        return vals

    def close(self) -> None:
        """Release the trainers this executor created (the reference notifies its pservers here)."""
        self._trainers.clear()
