"""Collective ops of the reference on lists of dense tensors.

* ``c_allreduce_xsum``  — grouped allreduce of several tensors in one go
  (operators/collective/c_allreduce_x_op.cc:44-163: one ncclGroupStart/End around N allreduces).
* ``c_mixallgather``    — fuse N dense grads into one buffer, then mode 0 allreduce /
  1 reduce-scatter + all-gather ("mixallgather") / 2 all-gather
  (operators/collective/c_mixallgather_op.cc:121-350).

Product path: the tensors are packed into a ``SymmetricBuffer`` and reduced by the one-kernel
peer-memory collective (``csrc/cuda/symm.cu``) with the scale fused.  ``backend="nccl"``
(or CPU tensors) uses ``torch.distributed`` — the labelled baseline, and the multi-node
fallback.  The trainer itself avoids the packing copy altogether by aliasing every gradient
into the symmetric buffer (``parallel/dense.py``); these functions serve ad-hoc tensor lists.
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.distributed as dist

_BUFS: dict = {}


def _world(group=None):
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _symm(numel: int, device, group):
    from .symm import SymmetricBuffer
    rank, world = _world(group)
    key = (str(device), id(group))
    buf = _BUFS.get(key)
    if buf is None or buf.numel < numel:
        cap = max(numel, 1 << 20)
        buf = _BUFS[key] = SymmetricBuffer(cap, rank, world, group=group, device=device)
    return buf


def _pack(tensors, buf):
    offs, off = [], 0
    for t in tensors:
        n = t.numel()
        buf.tensor[off:off + n].copy_(t.reshape(-1))
        offs.append(off)
        off += (n + 3) // 4 * 4
    return offs, off


def _unpack(tensors, buf, offs):
    for t, o in zip(tensors, offs):
        t.copy_(buf.tensor[o:o + t.numel()].view_as(t))


def c_allreduce_xsum(tensors: Sequence[torch.Tensor], scale: float = 1.0, group=None, backend: str = "auto"):
    """In-place sum of every tensor over the group (then ``* scale``)."""
    rank, world = _world(group)
    if world == 1:
        if scale != 1.0:
            for t in tensors:
                t.mul_(scale)
        return tensors
    cuda = all(t.is_cuda for t in tensors)
    if backend == "nccl" or not cuda:
        for t in tensors:                       # NCCL coalesces these on one stream
            dist.all_reduce(t, group=group)
            if scale != 1.0:
                t.mul_(scale)
        return tensors
    total = sum((t.numel() + 3) // 4 * 4 for t in tensors)
    buf = _symm(total, tensors[0].device, group)
    offs, used = _pack(tensors, buf)
    buf.allreduce(0, used, scale)
    _unpack(tensors, buf, offs)
    return tensors


def c_mixallgather(tensors: Sequence[torch.Tensor], mode: int = 0, scale: float | None = None, group=None,
                   backend: str = "auto", topo=None):
    """mode 0: allreduce; 1: reduce-scatter then all-gather (same result, the two halves are
    what a sharded optimizer interleaves with its update); 2: all-gather of every rank's shard.

    ``topo`` (``parallel.topology.NodeTopology``) with several nodes: modes 0 / 1 run the
    multi-NIC path of the reference (c_mixallgather_op.cc:231-328) — reduce-scatter inside the
    node, ``SyncDense`` across nodes on each rank's shard, all-gather inside the node."""
    rank, world = _world(group)
    scale = (1.0 / world) if scale is None else scale
    if topo is not None and topo.multi_node and mode in (0, 1):
        flat = torch.cat([t.reshape(-1) for t in tensors])
        topo.hierarchical_allreduce(flat, scale)
        off = 0
        for t in tensors:
            t.copy_(flat[off:off + t.numel()].view_as(t))
            off += t.numel()
        return tensors
    if mode == 0 or world == 1:
        return c_allreduce_xsum(tensors, scale, group, backend)
    cuda = all(t.is_cuda for t in tensors)
    if backend == "nccl" or not cuda:
        if mode == 1:
            return c_allreduce_xsum(tensors, scale, group, "nccl")
        for t in tensors:                                   # mode 2: each rank owns a contiguous shard
            flat = t.reshape(-1)
            chunk = (flat.numel() + world - 1) // world
            shard = torch.zeros(chunk, dtype=t.dtype, device=t.device)
            b = min(flat.numel(), rank * chunk)
            e = min(flat.numel(), b + chunk)
            shard[:e - b] = flat[b:e]
            out = [torch.empty_like(shard) for _ in range(world)]
            dist.all_gather(out, shard, group=group)
            flat.copy_(torch.cat(out)[:flat.numel()])
        return tensors
    total = sum((t.numel() + 3) // 4 * 4 for t in tensors)
    buf = _symm(total, tensors[0].device, group)
    offs, used = _pack(tensors, buf)
    if mode == 1:
        buf.collective("reduce_scatter", 0, used, scale)
    buf.collective("all_gather", 0, used, 1.0)
    _unpack(tensors, buf, offs)
    return tensors


def c_broadcast(tensors: Sequence[torch.Tensor], root: int = 0, group=None):
    rank, world = _world(group)
    if world > 1:
        for t in tensors:
            dist.broadcast(t, src=root, group=group)
    return tensors


def barrier(group=None):
    if dist.is_available() and dist.is_initialized():
        dist.barrier(group=group)
