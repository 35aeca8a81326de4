"""Sharded sparse pull / push over ``torch.distributed`` all-to-all.

Where the peer-memory engine (``parallel/sparse_engine.py``) cannot reach — ranks on other
nodes, or a CPU-only world — the sparse table is still sharded by key over every rank, and a
pull / push is: dedup, bucket by owner, exchange counts, all-to-all the keys, serve them from
the owner's shard, all-to-all the rows back (push: all-to-all merged gradients, the owner
merges what several ranks sent for one key and runs the optimizer).

Reference role: the multi-node branch of HeterComm — gather keys / gradients from the other
nodes, select the local shard, merge, update (heter_ps/heter_comm_inl.h:1996-2198) — and, in
BoxPS proper, whatever the closed library does between nodes behind ``PullSparseGPU`` /
``PushSparseGPU``.  The owner of a key is ``pbx_owner_of(key, world)`` — the same function the
one-node engine uses, so a table saved by one can be loaded by the other.

The shard behind the router is either the host table (CPU worlds; the rows live in
``host.HostTable``) or the HBM shard (``ops.sparse_table.SparseTable`` lookup / update
kernels) — the exchange itself is NCCL on GPUs, gloo on CPU.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def _mix64(k: torch.Tensor) -> torch.Tensor:
    """splitmix finaliser on int64 with wrap-around (logical shifts emulated)."""
    def c(v):
        return torch.tensor(v - (1 << 64) if v >= (1 << 63) else v, dtype=torch.int64, device=k.device)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xff51afd7ed558ccd)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xc4ceb9fe1a85ec53)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    return k


def owner_of_t(keys: torch.Tensor, nranks: int) -> torch.Tensor:
    """Tensor twin of ``box_wrapper.owner_of`` / ``pbx_owner_of`` (csrc/cuda/common.cuh)."""
    return ((_mix64(keys) >> 40) & 0xFFFFFF) % nranks


class DistSparseRouter:
    """``pull(keys) -> (values [U, hidden], restore int32 [N])`` and ``push(merged [U, P])`` with the
    same contract as ``BoxWrapper.pull_sparse`` / the merged-record push, over any process group."""

    def __init__(self, rank: int, world: int, hidden: int, group=None, host_table=None, device_table=None):
        assert (host_table is None) != (device_table is None), "exactly one shard backend"
        self.rank, self.world, self.hidden, self.group = rank, world, hidden, group
        self.host_table, self.device_table = host_table, device_table
        self._last = None

    # ---- owner side ----------------------------------------------------------------------------
    def _serve(self, keys: torch.Tensor) -> torch.Tensor:
        if self.device_table is not None:
            return self.device_table.lookup(keys)[:, :self.hidden]
        if keys.numel() == 0:
            return torch.zeros(0, self.hidden)
        return self.host_table.pull(keys.numpy().view(np.uint64))[:, :self.hidden].contiguous()

    def _apply(self, keys: torch.Tensor, recs: torch.Tensor) -> None:
        """Several ranks may have sent the same key: merge (sum every column but the slot id)."""
        if keys.numel() == 0:
            return
        uk, inv = torch.unique(keys, return_inverse=True)
        merged = torch.zeros(uk.numel(), recs.shape[1], dtype=recs.dtype, device=recs.device)
        merged.index_add_(0, inv, recs)
        merged[inv, 0] = recs[:, 0]                         # slot id is not additive
        if self.device_table is not None:
            self.device_table.update(uk, merged)
        else:
            self.host_table.push(uk.numpy().view(np.uint64), merged)

    # ---- exchange helpers ------------------------------------------------------------------------
    def _counts(self, send_counts: torch.Tensor) -> list:
        recv = torch.empty_like(send_counts)
        dist.all_to_all_single(recv, send_counts, group=self.group)
        return recv.tolist()

    def _a2a(self, send: torch.Tensor, send_counts: list, recv_counts: list) -> torch.Tensor:
        out = send.new_empty((sum(recv_counts),) + tuple(send.shape[1:]))
        dist.all_to_all_single(out, send.contiguous(), output_split_sizes=recv_counts,
                               input_split_sizes=send_counts, group=self.group)
        return out

    # ---- requester side -----------------------------------------------------------------------------
    def pull(self, keys: torch.Tensor):
        uk, inv = torch.unique(keys, return_inverse=True)
        own = owner_of_t(uk, self.world)
        order = torch.argsort(own, stable=True)
        sorted_keys = uk[order]
        send_counts = torch.bincount(own, minlength=self.world)
        sc = send_counts.tolist()
        rc = self._counts(send_counts)
        asked = self._a2a(sorted_keys, sc, rc)              # keys other ranks want from my shard
        rows = self._serve(asked)
        back = self._a2a(rows, rc, sc)                      # rows of my keys, in `sorted_keys` order
        vals = torch.empty(uk.numel(), self.hidden, dtype=torch.float32, device=keys.device)
        vals[order] = back
        self._last = (sorted_keys, order, sc, rc)
        return vals, inv.to(torch.int32)

    def push(self, merged: torch.Tensor) -> None:
        """``merged`` [U, P]: one push record per unique key of the LAST pull (row = restore idx)."""
        sorted_keys, order, sc, rc = self._last
        recs = self._a2a(merged[order], sc, rc)
        keys = self._a2a(sorted_keys, sc, rc)
        self._apply(keys, recs)
