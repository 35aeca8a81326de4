"""Host-side world object: rank / size / barrier / scalar and vector allreduce.

Stands in for the closed library's ``boxps::MPICluster`` (call sites: fleet/box_wrapper.h:444-450,
data_set.cc:2708-2712 batch-count equalisation, fleet/metrics.cc:288-332 metric reduction) and
for ``GlooWrapper`` (fleet/gloo_wrapper.h), over ``torch.distributed``.  Host values travel over
a gloo group when one exists (no GPU sync, usable from loader threads); with an nccl-only
world they are staged through the current device.
"""
from __future__ import annotations

import os
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Sequence

import numpy as np
import torch
import torch.distributed as dist

OP_MAX, OP_MIN, OP_SUM = 0, 1, 2       # the integer codes the reference passes to MPICluster::allreduce


class Cluster:
    _ins: "Cluster | None" = None
    _lock = threading.Lock()

    def __init__(self, group=None):
        self.group = group
        self._host_group = None

    @classmethod
    def ins(cls) -> "Cluster":
        with cls._lock:
            if cls._ins is None:
                cls._ins = Cluster()
            return cls._ins

    # -- identity ------------------------------------------------------------------------------
    def initialized(self) -> bool:
        return dist.is_available() and dist.is_initialized()

    def size(self) -> int:
        return dist.get_world_size(self.group) if self.initialized() else 1

    def rank(self) -> int:
        return dist.get_rank(self.group) if self.initialized() else 0

    def nccl_rank(self, device_id: int, ngpu: int) -> int:
        """``mpi_rank * ngpu + dev`` (fleet/box_wrapper.h:448-450)."""
        return self.rank() * ngpu + device_id

    # -- collectives on host values -------------------------------------------------------------
    def _staging_device(self):
        if self.initialized() and dist.get_backend(self.group) == "nccl":
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")

    def barrier(self) -> None:
        if self.size() > 1:
            dist.barrier(self.group)

    def allreduce(self, value, op: int = OP_SUM):
        """Scalar allreduce; ``op``: 0 max, 1 min, 2 sum."""
        if self.size() == 1:
            return value
        rop = {OP_MAX: dist.ReduceOp.MAX, OP_MIN: dist.ReduceOp.MIN, OP_SUM: dist.ReduceOp.SUM}[op]
        is_int = isinstance(value, (int, np.integer))
        t = torch.tensor([value], dtype=torch.int64 if is_int else torch.float64, device=self._staging_device())
        dist.all_reduce(t, op=rop, group=self.group)
        return int(t.item()) if is_int else float(t.item())

    def allreduce_sum(self, values: "np.ndarray | Sequence[float]") -> np.ndarray:
        a = np.ascontiguousarray(values, dtype=np.float64)
        if self.size() == 1:
            return a
        t = torch.from_numpy(a.copy()).to(self._staging_device())
        dist.all_reduce(t, group=self.group)
        return t.cpu().numpy()

    def all_to_all_arrays(self, parts: Sequence[np.ndarray]) -> list:
        """``parts[d]`` (1-D array, same dtype on every rank) goes to rank d; returns what every rank
        sent here, by source rank.  One size exchange + one ``all_to_all`` of the raw data — no pickling."""
        if self.size() == 1:
            return [np.asarray(parts[0])]
        dev = self._staging_device()
        dtype = np.asarray(parts[0]).dtype
        send_counts = torch.tensor([int(np.asarray(p).size) for p in parts], dtype=torch.int64, device=dev)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts, group=self.group)
        sc, rc = send_counts.tolist(), recv_counts.tolist()
        flat = np.ascontiguousarray(np.concatenate([np.asarray(p).reshape(-1) for p in parts])).view(np.uint8)
        send = torch.from_numpy(flat.copy()).to(dev)
        item = dtype.itemsize
        recv = torch.empty(sum(rc) * item, dtype=torch.uint8, device=dev)
        dist.all_to_all_single(recv, send, output_split_sizes=[c * item for c in rc],
                               input_split_sizes=[c * item for c in sc], group=self.group)
        got = recv.cpu().numpy().view(dtype)
        out, off = [], 0
        for c in rc:
            out.append(got[off:off + c])
            off += c
        return out

    def allgather_object(self, obj) -> list:
        if self.size() == 1:
            return [obj]
        out = [None] * self.size()
        dist.all_gather_object(out, obj, group=self.group)
        return out


# ---- per-device CPU parallel-for (boxps ExecuteFunc / ExecRangeFunc, fleet/box_wrapper.h:834-844)
_POOL: ThreadPoolExecutor | None = None


def _pool() -> ThreadPoolExecutor:
    global _POOL
    if _POOL is None:
        _POOL = ThreadPoolExecutor(max_workers=max(2, min(32, (os.cpu_count() or 2))))
    return _POOL


def execute_func(n: int, fn: Callable[[int], None]) -> None:
    """Run ``fn(i)`` for i in [0, n) on the host pool and wait."""
    list(_pool().map(fn, range(n)))


def exec_range_func(n: int, fn: Callable[[int, int], None], parts: int | None = None) -> None:
    """Run ``fn(begin, end)`` over a partition of [0, n) on the host pool and wait."""
    parts = max(1, min(parts or _pool()._max_workers, n))
    step = (n + parts - 1) // parts
    futs = [_pool().submit(fn, b, min(n, b + step)) for b in range(0, n, step)]
    for f in futs:
        f.result()


# cores of this process as they were when the module was imported: pinning one thread later must not
# shrink what the other helpers see
_PROCESS_CORES = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))


def get_train_cores() -> list:
    """Cores the trainer threads may be pinned to (boxps::get_train_cores, boxps_trainer.cc:179):
    the first half of this process's cores."""
    return _PROCESS_CORES[:max(1, len(_PROCESS_CORES) // 2)]


def get_readins_cores() -> list:
    """Cores for the reader threads (boxps::get_readins_cores, data_set.cc:2028): the other half."""
    return _PROCESS_CORES[len(_PROCESS_CORES) // 2:] or _PROCESS_CORES


def set_cpu_affinity(cores: Sequence[int]) -> bool:
    """Pin the calling thread (``ThreadPool::SetCPUAffinity``, threadpool.h:104)."""
    if not hasattr(os, "sched_setaffinity") or not cores:
        return False
    try:
        os.sched_setaffinity(threading.get_native_id(), set(cores))
        return True
    except OSError:
        return False
