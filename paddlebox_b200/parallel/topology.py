"""Node topology and the inter-node leg of the dense collectives.

The reference runs its dense sync hierarchically when there is more than one node
(boxps_worker.cc:1220-1233): ``ncclReduceScatter`` inside the node, the closed
``BoxPSBase::SyncDense(stream, size, send, recv, device_id, allgather)`` across nodes on each
GPU's shard (fleet/box_wrapper.h:686-698), ``ncclAllGather`` inside the node, then one scale by
``1/(ndev*nnode)``.  ``c_mixallgather`` has the same three legs
(operators/collective/c_mixallgather_op.cc:231-328).

Here the intra-node legs are the one-kernel peer-memory collectives over NVLink / NVSwitch
(``parallel/symm.py``); only the inter-node leg goes through ``torch.distributed`` (NCCL over the
NICs on GPUs, gloo on CPU), on a group of the ranks that share a local rank — so every NIC
carries ``1/ndev`` of the buffer, like the reference's multi-NIC path.

World rank layout is the reference's: ``rank = node * gpus_per_node + device``
(fleet/box_wrapper.h:448-450).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class NodeTopology:
    """Splits a world into nodes.  ``gpus_per_node`` defaults to LOCAL_WORLD_SIZE (torchrun) or the
    whole world (one node).  Creating it is a collective call: every rank of ``group`` must do it."""

    def __init__(self, gpus_per_node: int | None = None, group=None):
        self.group = group
        ok = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if ok else 0
        self.world = dist.get_world_size(group) if ok else 1
        if gpus_per_node is None:
            gpus_per_node = int(os.environ.get("PBX_GPUS_PER_NODE", os.environ.get("LOCAL_WORLD_SIZE", self.world)))
        gpus_per_node = max(1, min(int(gpus_per_node), self.world))
        if self.world % gpus_per_node:
            raise ValueError(f"world size {self.world} is not a multiple of gpus_per_node {gpus_per_node}")
        self.local_size = gpus_per_node
        self.nnodes = self.world // gpus_per_node
        self.node_rank, self.local_rank = divmod(self.rank, gpus_per_node)
        self.intra_group = group
        self.inter_group = None
        if self.nnodes > 1:
            # new_group is collective over the whole world: every rank creates every group
            for n in range(self.nnodes):
                ranks = [n * gpus_per_node + d for d in range(gpus_per_node)]
                g = dist.new_group(self._global(ranks))
                if n == self.node_rank:
                    self.intra_group = g
            for d in range(gpus_per_node):
                ranks = [n * gpus_per_node + d for n in range(self.nnodes)]
                g = dist.new_group(self._global(ranks))
                if d == self.local_rank:
                    self.inter_group = g

    def _global(self, ranks):
        if self.group is None:
            return ranks
        return [dist.get_global_rank(self.group, r) for r in ranks]

    @property
    def multi_node(self) -> bool:
        return self.nnodes > 1

    def shard_range(self, numel: int, align: int = 4):
        """Slice of a flat buffer this rank owns inside its node (what it sends across nodes)."""
        chunk = ((numel + self.local_size - 1) // self.local_size + align - 1) // align * align
        b = min(numel, self.local_rank * chunk)
        return b, min(numel, b + chunk)

    # ---- the closed library's SyncDense --------------------------------------------------------
    def sync_dense(self, send: torch.Tensor, recv: torch.Tensor | None = None, allgather: bool = False,
                   scale: float = 1.0) -> torch.Tensor:
        """Inter-node leg on this rank's shard.  ``allgather=False``: sum over the nodes (in place when
        ``recv`` is None).  ``allgather=True``: ``recv`` (``nnodes * send.numel()``) receives every
        node's ``send`` in node order."""
        if allgather:
            if recv is None:
                recv = torch.empty(self.nnodes * send.numel(), dtype=send.dtype, device=send.device)
            if self.nnodes == 1:
                recv.copy_(send.reshape(-1))
            else:
                dist.all_gather_into_tensor(recv, send.contiguous().reshape(-1), group=self.inter_group) \
                    if send.is_cuda else self._cpu_all_gather(recv, send)
            return recv
        out = send if recv is None else recv
        if recv is not None:
            recv.copy_(send)
        if self.nnodes > 1:
            dist.all_reduce(out, group=self.inter_group)
        if scale != 1.0:
            out.mul_(scale)
        return out

    def _cpu_all_gather(self, recv, send):
        parts = list(recv.reshape(self.nnodes, -1).unbind(0))
        dist.all_gather(parts, send.contiguous().reshape(-1), group=self.inter_group)
        return recv

    # ---- three-leg allreduce of a flat buffer --------------------------------------------------
    def hierarchical_allreduce(self, flat: torch.Tensor, scale: float = 1.0, symm=None, offset: int = 0) -> None:
        """``flat`` <- sum over every rank of the world * ``scale``.

        ``symm``: the node-local ``SymmetricBuffer`` that ``flat`` is a view of (at ``offset``
        elements) — then legs 1 and 3 are the peer-memory kernels; otherwise ``torch.distributed``
        on the intra-node group (gloo has no reduce-scatter: the first leg is an allreduce there)."""
        n = flat.numel()
        if self.world == 1:
            if scale != 1.0:
                flat.mul_(scale)
            return
        if not self.multi_node:
            if symm is not None:
                symm.allreduce(offset, n, scale)
            else:
                dist.all_reduce(flat, group=self.intra_group)
                if scale != 1.0:
                    flat.mul_(scale)
            return
        b, e = self.shard_range(n)
        if symm is not None:
            symm.collective("reduce_scatter", offset, n, scale)          # leg 1, scale fused
            if e > b:
                dist.all_reduce(flat[b:e], group=self.inter_group)       # leg 2: 1/ndev of the buffer per NIC
            else:
                dist.all_reduce(flat[:0], group=self.inter_group)
            symm.collective("all_gather", offset, n, 1.0)                # leg 3
            return
        if flat.is_cuda:
            chunk =((n + self.local_size - 1) // self.local_size + 3) // 4 * 4
            padded = torch.zeros(chunk * self.local_size, dtype=flat.dtype, device=flat.device)
            padded[:n] = flat
            shard = torch.empty(chunk, dtype=flat.dtype, device=flat.device)
            dist.reduce_scatter_tensor(shard, padded, group=self.intra_group)
            dist.all_reduce(shard, group=self.inter_group)
            if scale != 1.0:
                shard.mul_(scale)
            dist.all_gather_into_tensor(padded, shard, group=self.intra_group)
            flat.copy_(padded[:n])
            return
        dist.all_reduce(flat, group=self.intra_group)                    # CPU: gloo
        mine = flat[b:e].clone()
        dist.all_reduce(mine, group=self.inter_group)
        if scale != 1.0:
            mine.mul_(scale)
        chunk = ((n + self.local_size - 1) // self.local_size + 3) // 4 * 4
        send = torch.zeros(chunk, dtype=flat.dtype)
        send[:e - b] = mine
        parts = [torch.empty(chunk, dtype=flat.dtype) for _ in range(self.local_size)]
        dist.all_gather(parts, send, group=self.intra_group)
        flat.copy_(torch.cat(parts)[:n])
