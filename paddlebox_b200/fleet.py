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


def init(role_maker=None, is_collective: bool = True, strategy=None, backend: str | None = None) -> None:
    """Join the world described by the launcher's environment (RANK / WORLD_SIZE / MASTER_ADDR /
    MASTER_PORT, as torchrun sets them); a single process needs no environment."""
    global _cluster
    import torch.distributed as dist
    if not dist.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0"))))
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"))
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
    return True


def is_server() -> bool:
    return False           # the parameter server lives inside the workers' GPUs


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
