"""``fleet`` — the role / world helper user scripts call around the PS
(reference: ``paddle.distributed.fleet``: ``fleet.init / worker_index / worker_num / is_first_worker /
barrier_worker / stop_worker`` in python/paddle/distributed/fleet/base/fleet_base.py and
``fleet.util`` in .../base/util_factory.py: ``all_reduce / all_gather / barrier / get_file_shard``).
One process per GPU; the world is the ``torch.distributed`` group (``Cluster``)."""
from __future__ import annotations

import os
from typing import Sequence

import numpy as np
import torch

from .parallel.cluster import OP_MAX, OP_MIN, OP_SUM, Cluster

_cluster: Cluster | None = None


_role_maker = None
_init_strategy = None


def init(role_maker=None, is_collective: bool = True, strategy=None, backend: str | None = None) -> None:
    """Join the job's world.  Without a role maker: the launcher's environment as torchrun sets it (RANK /
    WORLD_SIZE / MASTER_ADDR / MASTER_PORT); a single process needs no environment.  With one
    (``role_maker.py``: ``GeneralRoleMaker``, ``PaddleCloudRoleMaker`` ...): rank, world size and the rendezvous
    address come from the Paddle launcher's variables it reads (``PADDLE_TRAINER_ID``, ``PADDLE_TRAINER_ENDPOINTS``,
    ``TRAINING_ROLE``); a process launched as a parameter server joins nothing (there are no server processes)."""
    global _cluster, _role_maker, _init_strategy
    import datetime
    import torch.distributed as dist
    _role_maker = role_maker
    _init_strategy = strategy                  # fleet 2.0: the strategy may be given here instead of to the optimizer
    backend = backend or os.environ.get("PADDLE_DISTRI_BACKEND") or None
    if backend == "auto":
        backend = None
    rank, world, addr, port, timeout = None, None, None, None, None
    if role_maker is not None:
        role_maker.generate_role()
        if role_maker.is_server():
            _cluster = Cluster.ins()
            return
        rank, world = role_maker.worker_index(), role_maker.worker_num()
        addr, port = role_maker.rendezvous()
        timeout = getattr(role_maker, "_init_timeout_seconds", None)
    else:
        world = int(os.environ.get("WORLD_SIZE", "1"))
    if not dist.is_initialized() and world > 1:
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", str(rank or 0)))) %
                                  max(1, torch.cuda.device_count()))
        kw = {}
        if role_maker is not None:
            if addr is None:
                raise ValueError("the role maker knows no trainer endpoint to rendezvous at")
            kw = {"init_method": f"tcp://{addr}:{port}", "rank": rank, "world_size": world}
            if timeout:
                kw["timeout"] = datetime.timedelta(seconds=min(int(timeout), 7 * 24 * 3600))
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"), **kw)
    _cluster = Cluster.ins()


def _c() -> Cluster:
    return _cluster or Cluster.ins()


def worker_index() -> int:
    return _c().rank()


def worker_num() -> int:
    return _c().size()


def is_first_worker() -> bool:
    return worker_index() == 0


def is_worker() -> bool:
    return True if _role_maker is None else _role_maker.is_worker()


def is_server() -> bool:
    """False for every process that trains: the parameter server lives inside the workers' GPUs.  True only for
    a process the launcher started with ``TRAINING_ROLE=PSERVER`` (it has nothing to do: ``run_server`` returns)."""
    return False if _role_maker is None else _role_maker.is_server()


def barrier_worker() -> None:
    _c().barrier()


def init_worker() -> None:
    """Nothing to connect to: there are no separate server processes."""


def init_server(*a, **kw) -> None:
    """See ``is_server``."""


def stop_worker() -> None:
    import torch.distributed as dist
    from .box_wrapper import BoxWrapper
    BoxWrapper.reset_instance()
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


class _Util:
    def all_reduce(self, value, mode: str = "sum", comm_world: str = "worker"):
        op = {"sum": OP_SUM, "max": OP_MAX, "min": OP_MIN}[mode]
        if isinstance(value, (int, float, np.integer, np.floating)):
            return _c().allreduce(value, op)
        a = np.asarray(value, dtype=np.float64)
        if mode == "sum":
            return _c().allreduce_sum(a.reshape(-1)).reshape(a.shape)
        return np.array([_c().allreduce(float(v), op) for v in a.reshape(-1)]).reshape(a.shape)

    def all_gather(self, value, comm_world: str = "worker") -> list:
        return _c().allgather_object(value)

    def barrier(self, comm_world: str = "worker") -> None:
        _c().barrier()

    def get_file_shard(self, files: Sequence[str]) -> list:
        """This worker's contiguous share of ``files``: the first ``len % n`` workers get one more
        (util_factory.py ``get_file_shard``)."""
        n, r = worker_num(), worker_index()
        per, rem = divmod(len(files), n)
        begin = r * per + min(r, rem)
        return list(files[begin:begin + per + (1 if r < rem else 0)])

    def print_on_rank(self, message: str, rank_id: int = 0) -> None:
        if worker_index() == rank_id:
            print(message)


util = _Util()

from .distributed_strategy import DistributedStrategy  # noqa: E402,F401  ``fleet.DistributedStrategy()``

from . import fleet_metrics as metrics  # noqa: E402,F401  ``fleet.metrics.auc(stat_pos, stat_neg)`` ...


# ---- the parameter-server calls of the reference's pslib fleet ------------------------------------------------
# (python/paddle/fluid/incubate/fleet/parameter_server/pslib/__init__.py:214-800).  Job scripts written for the
# CPU / GPU parameter server call these on ``fleet``; here they act on the process-wide wrapper (there is one
# sparse table, so ``table_id`` arguments are accepted and not needed).
def _box():
    from .box_wrapper import BoxWrapper
    return BoxWrapper.get_instance()


def save_persistables(executor=None, dirname: str = "", main_program=None, mode: int = 0, **kwargs) -> int:
    """``mode`` 0 = full state, 1 = delta (serving format, moves the delta baseline), 2 = serving base,
    3 = batch model (pslib :348-370).  Returns the rows this worker wrote."""
    box = _box()
    box._sync_to_host()
    n = box.host_table.save(dirname, mode=0 if mode == 3 else mode, shard_base=box.rank, clear_dirty=(mode == 1),
                            base_threshold=float(kwargs.get("base_threshold", table_policy.get("base_threshold", 0.0)))
                            if mode == 2 else 0.0,
                            delta_threshold=float(kwargs.get("delta_threshold", table_policy.get("delta_threshold", 0.0)))
                            if mode == 1 else 0.0)
    barrier_worker()
    return n


def save_model(model_dir: str | None = None, mode: int = 0, **kwargs) -> int:
    return save_persistables(None, model_dir, mode=mode)


def save_one_table(table_id, model_dir: str, mode: int = 0, **kwargs) -> int:
    return save_persistables(None, model_dir, mode=mode)


def save_multi_table_one_path(table_ids, model_dir: str, mode: int = 0, **kwargs) -> int:
    return save_persistables(None, model_dir, mode=mode)


def save_cache_model(executor=None, dirname: str = "", main_program=None, mode: int = 1, table_id: int = 0,
                     cache_threshold: float = 1.0, **kwargs) -> int:
    """Hot rows in serving format; returns the key count over all workers (pslib :429-469)."""
    box = _box()
    box._sync_to_host()
    n = box.host_table.save(dirname, mode=2, shard_base=box.rank, base_threshold=cache_threshold, clear_dirty=False)
    return int(util.all_reduce(int(n), "sum"))


def load_model(model_dir: str | None = None, mode: int = 0, **kwargs) -> int:
    n = _box().load_model(model_dir, serving=mode in (1, 2))
    barrier_worker()
    return n


def load_one_table(table_id, model_path: str, mode: int = 0, **kwargs) -> int:
    return load_model(model_path, mode=mode)


def shrink_sparse_table(decay: float | None = None, delete_threshold: float | None = None,
                        delete_after_unseen_days: float | None = None) -> int:
    """Day-end decay of show / click and removal of stale rows (pslib :471-486); returns the rows removed here.
    Arguments left out take the job strategy's values (``distributed_optimizer``), else 0.98 / 0 / 30."""
    decay = table_policy.get("show_click_decay_rate", 0.98) if decay is None else decay
    delete_threshold = table_policy.get("delete_threshold", 0.0) if delete_threshold is None else delete_threshold
    delete_after_unseen_days = table_policy.get("delete_after_unseen_days", 30.0) \
        if delete_after_unseen_days is None else delete_after_unseen_days
    box = _box()
    box._sync_to_host()
    before = box.host_table.size()
    kept = box.shrink_table(decay, delete_threshold, delete_after_unseen_days)
    barrier_worker()
    return before - kept


def shrink_dense_table(decay: float, emb_dim: int = 11, scope=None, table_id=None, trainer=None) -> None:
    """Decay the data-norm summaries of the dense part (pslib :488-521 multiplies the ``batch_size`` /
    ``batch_sum`` / ``batch_square_sum`` parameters by ``decay``).  ``trainer``: the ``CTRTrainer`` whose model
    holds them."""
    if trainer is None:
        return
    import torch
    with torch.no_grad():
        for m in trainer.model.modules():
            for name in ("batch_size", "batch_sum", "batch_square_sum"):
                t = getattr(m, name, None)
                if isinstance(t, torch.Tensor):
                    t.mul_(decay)
    barrier_worker()


def clear_one_table(table_id=None) -> None:
    clear_model()


def clear_model() -> None:
    """Drop every sparse row (host tier and HBM shard) — a job that restarts training from scratch (pslib :537)."""
    box = _box()
    box.host_table.clear()
    if getattr(box, "table", None) is not None:
        box.table.clear()
    barrier_worker()


def print_table_stat(table_id=None) -> dict:
    """Rows in memory / on the SSD tier of this worker's shard (pslib :318-331 prints the server's counters)."""
    box = _box()
    stat = {"rank": worker_index(), "mem_rows": box.host_table.mem_rows() if hasattr(box.host_table, "mem_rows")
            else box.host_table.size(), "ssd_rows": box.host_table.ssd_rows() if box._ssd else 0,
            "rows": box.host_table.size()}
    print(f"[table stat] {stat}")
    return stat


def set_date(table_id, date) -> None:
    _box().set_date(str(date))


def set_file_num_one_shard(table_id, file_num: int) -> None:
    """Accepted: a shard is always written as one ``part-<rank>-<shard>`` file here."""


def end_pass(scope=None) -> None:
    _box().end_pass()


table_policy: dict = {}       # thresholds of the day-end calls, from the job's strategy (see distributed_optimizer)


def distributed_optimizer(optimizer, strategy=None):
    """The reference wraps the optimizer into a ``DownpourOptimizer`` that rewrites the program for the PS
    (pslib :281-294) and turns ``strategy`` into the fleet description (pslib/node.py).  Here the sparse part is
    optimized inside the table and the dense part by ``CTRTrainer``'s flat optimizer configured from
    ``optimizer``; a ``strategy`` dict with the pslib keys (``sparse_learning_rate``, ``sparse_embedx_threshold``,
    ``sparse_delete_threshold`` ...) configures the wrapper's sparse optimizer (if the wrapper is initialised
    already) and the thresholds ``shrink_sparse_table`` / ``save_persistables`` use by default."""
    from .optimizer import BoxPSOptimizer
    from .distributed_strategy import DistributedStrategy as _DS
    if strategy is None:
        strategy = _init_strategy
    ds = strategy if isinstance(strategy, _DS) else None
    if ds is not None:
        strategy = ds.sparse_strategy()
    if strategy:
        from .fleet_desc import sparse_config_from_strategy
        conf, kind, _, policy = sparse_config_from_strategy(dict(strategy))
        table_policy.clear()
        table_policy.update(policy)
        table_policy["sparse_conf"] = {**conf, "optimizer": kind}
        from .box_wrapper import BoxWrapper
        box = BoxWrapper._instance
        if box is not None and getattr(box, "host_table", None) is not None and kind == box.optimizer:
            from . import _native as N
            cfg = N.SparseOptConfig(**{**box.host_table.config.as_dict(), **conf})
            box.host_table.set_config(cfg)
            if getattr(box, "table", None) is not None:
                box.table.set_config(cfg)
    global _last_optimizer
    if isinstance(optimizer, BoxPSOptimizer):
        _last_optimizer = optimizer
    elif ds is not None:
        mode, k, is_async = ds.dense_mode()
        _last_optimizer = BoxPSOptimizer(optimizer, sync_steps=k, sync_dense_mode=mode, async_mode=is_async)
    else:
        _last_optimizer = BoxPSOptimizer(optimizer)
    _last_optimizer._strategy = ds
    return _last_optimizer


# ---- the remaining role / topology queries of fleet_base.py -------------------------------------------------
def node_num() -> int:
    """Number of nodes (machines) of the job: world size / processes per node (torchrun's LOCAL_WORLD_SIZE)."""
    lw = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    return max(1, worker_num() // lw)


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", "0"))


def local_device_ids() -> list:
    """The device this process drives (one process per GPU)."""
    return [torch.cuda.current_device()] if torch.cuda.is_available() else []


def world_device_ids() -> list:
    lw = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    return [[r % lw] for r in range(worker_num())]


def worker_endpoints(to_string: bool = False):
    """``host:port`` of every worker as far as this process knows them: the rendezvous address for rank 0 and
    PADDLE_TRAINER_ENDPOINTS when the launcher exported it."""
    eps = [e for e in os.environ.get("PADDLE_TRAINER_ENDPOINTS", "").split(",") if e]
    if not eps:
        eps = [f"{os.environ.get('MASTER_ADDR', '127.0.0.1')}:{os.environ.get('MASTER_PORT', '0')}"]
    return ",".join(eps) if to_string else eps


def server_num() -> int:
    return 0               # no separate server processes: the table lives in the workers' HBM


def server_index() -> int:
    return 0


def server_endpoints(to_string: bool = False):
    return "" if to_string else []


def is_coordinator() -> bool:
    return False


def run_server() -> None:
    """See ``is_server``: nothing to run."""


def distributed_model(model):
    """Dygraph ``fleet.distributed_model``: the dense data-parallel wrapper.  Dense synchronisation is done by
    ``CTRTrainer`` (flat gradient buffer, in-kernel or symmetric-memory allreduce), so the model is returned as is."""
    return model


def shrink(threshold=None) -> int:
    """``fleet.shrink``: day-end table shrink with the table's configured thresholds."""
    return shrink_sparse_table() if threshold is None else shrink_sparse_table(delete_after_unseen_days=float(threshold))


def save(dirname: str, feed=(), fetch=(), mode: int = 0, trainer=None, **configs) -> int:
    """``fleet.save``: the sparse table (``mode`` as in ``save_persistables``) and, with ``trainer``, its dense
    persistables under ``<dirname>/dense``."""
    n = save_persistables(None, dirname, mode=mode)
    if trainer is not None and worker_index() == 0:
        from . import checkpoint
        checkpoint.save_persistables(trainer.model, os.path.join(dirname, "dense"), trainer.store)
    barrier_worker()
    return n


def save_inference_model(executor=None, dirname: str = "", feeded_var_names=None, target_vars=None, main_program=None,
                         export_for_deployment: bool = True, mode: int = 0, trainer=None) -> int:
    """Serving model: the serving-format sparse base (xbox layout) plus the dense parameters — what
    ``serving.Predictor`` loads."""
    n = save_persistables(None, dirname, mode=2)
    if trainer is not None and worker_index() == 0:
        from . import checkpoint
        checkpoint.save_persistables(trainer.model, os.path.join(dirname, "dense"))
    barrier_worker()
    return n


def _get_heter_file_shard(self, files):
    return self.get_file_shard(files)


_Util.get_heter_file_shard = _get_heter_file_shard


# ---- the rest of the pslib / fleet-base surface ---------------------------------------------------------------
_whitelist: np.ndarray | None = None
_client_config: dict = {}
_last_optimizer = None


def split_files(files) -> list:
    """This worker's block of ``files`` (fleet_base.py:154-191): as ``util.get_file_shard``; a list is required."""
    if not isinstance(files, list):
        raise TypeError("files should be a list of file need to be read.")
    return util.get_file_shard(files)


def all_reduce_worker(input, output) -> None:
    """``output[:] = sum over workers of input`` (role-maker collective, fleet_base.py:203-213)."""
    np.asarray(output)[...] = util.all_reduce(np.asarray(input), "sum")


def is_xpu() -> bool:
    return False


def set_pull_local_thread_num(thread_num: int) -> None:
    """Threads of the host side of a sparse pull (pslib :231-235): here the threads that fetch host-table rows into
    the staging slots when a pass is built (``FLAGS.pbx_pass_fetch_threads``; takes effect at the next wrapper)."""
    _client_config["pull_local_thread_num"] = int(thread_num)
    from .flags import FLAGS
    FLAGS.pbx_pass_fetch_threads = max(1, int(thread_num))


def _set_client_communication_config(request_timeout_ms: int, connect_timeout_ms: int, max_retry: int) -> None:
    """Kept for scripts that tune the PS client (pslib :221-229); there is no RPC client here, the values are stored."""
    _client_config.update(request_timeout_ms=request_timeout_ms, connect_timeout_ms=connect_timeout_ms, max_retry=max_retry)


def confirm() -> None:
    """``confirm`` / ``revert`` bracket a save on the reference's servers (pslib :520-548: keep or undo the delta
    flags of the last save).  Saves here move the delta baseline only after every part file is written, so there
    is nothing to undo; both are barriers."""
    barrier_worker()


def revert() -> None:
    barrier_worker()


def load_pslib_whitelist(table_id, model_path: str, **kwargs) -> int:
    """Read a key whitelist (one unsigned key per line, first column) for ``save_model_with_whitelist``
    (pslib :424-441).  ``model_path`` is a file or a directory of ``part-*`` files."""
    global _whitelist
    paths = [os.path.join(model_path, n) for n in sorted(os.listdir(model_path))] if os.path.isdir(model_path) else [model_path]
    keys = []
    for p in paths:
        with open(p) as f:
            keys += [int(ln.split()[0]) for ln in f if ln.strip()]
    _whitelist = np.unique(np.asarray(keys, dtype=np.uint64))
    return int(_whitelist.size)


def save_model_with_whitelist(executor=None, dirname: str = "", whitelist_path: str = "", main_program=None, **kwargs) -> int:
    """``save_persistables`` restricted to the keys of a whitelist (pslib :372-422); returns the rows written here."""
    from .host import HostTable
    if whitelist_path:
        load_pslib_whitelist(kwargs.get("table_id", 0), whitelist_path)
    if _whitelist is None:
        raise ValueError("save_model_with_whitelist: no whitelist (pass whitelist_path or call load_pslib_whitelist)")
    mode = int(kwargs.get("mode", 0))
    box = _box()
    box._sync_to_host()
    src = box.host_table
    present = np.intersect1d(src.keys(), _whitelist)
    part = HostTable(src.embedx_dim, src.optimizer, src.config)
    if present.size:
        part.store_rows(present, src.fetch_rows(present, create=False))
    n = part.save(dirname, mode=0 if mode == 3 else mode, shard_base=box.rank, clear_dirty=False)
    barrier_worker()
    return n


def train_from_dataset(model, dataset, trainer=None, **kwargs) -> dict:
    """``fleet.train_from_dataset`` (pslib :236-247): one pass of ``dataset`` through the job's trainer."""
    from .trainer import Executor
    if trainer is not None:
        return trainer.train_from_dataset(dataset)
    return Executor().train_from_dataset(model, dataset, **kwargs)


def _opt():
    if _last_optimizer is None:
        raise RuntimeError("call fleet.distributed_optimizer(...) first")
    return _last_optimizer


def minimize(model, startup_program=None, parameter_list=None, no_grad_set=None):
    """``fleet.minimize(loss)`` of fleet 2.0 (fleet_base.py:1400-1460) on the optimizer ``distributed_optimizer``
    returned: builds the job's trainer for ``model``."""
    return _opt().minimize(model, startup_program, parameter_list, no_grad_set)


def get_dist_env() -> dict:
    """Rank / world size / endpoints of this trainer (pslib :1190-1218)."""
    eps = worker_endpoints()
    r = worker_index()
    return {"trainer_id": r, "num_trainers": worker_num(), "current_endpoint": eps[r] if r < len(eps) else "",
            "trainer_endpoints": eps}
