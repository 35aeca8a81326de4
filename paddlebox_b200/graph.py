"""Graph engine: typed edge tables, neighbour sampling, random walks, walk-pair batches.

Reference role: ``GraphGpuWrapper`` / ``GpuPsGraphTable`` (heter_ps/graph_gpu_wrapper.h:34-99,
graph_gpu_ps_table_inl.cu) and ``GraphDataGenerator`` (data_feed.h:892-975, data_feed.cu:200-1206):
the graph lives next to the parameter server, walks over it produce (centre, context) node pairs,
and the node ids are feasigns whose embeddings are pulled / pushed like any other slot.

Here

* ``GraphTable`` is the native host store (csrc/host/graph.cc): edge lists -> CSR per edge type,
  multi-threaded sampling and (metapath) walks on the host;
* ``DeviceGraph`` is the same CSR resident in HBM: a B200 holds ~10 G edges in its 180 GB, so the
  whole graph of one rank stays on the device and sampling / walking is a handful of gather ops per
  step with no host round trip (the reference pages partitions through ``load_next_partition``);
* ``GraphDataGenerator`` turns walks into skip-gram batches in the ``SlotBatch`` layout (slot 0 =
  centre node, slot 1 = context node, one key per bag), so the sparse path is the stock one;
* ``GraphGpuWrapper`` keeps the reference's method names.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, Sequence

import numpy as np
import torch

from . import _native as N


def _u64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint64))


class GraphTable:
    """Host graph store; edge type = small integer."""

    def __init__(self):
        self._h = C.c_void_p()
        lib = N.host()
        N.hcheck(lib.pbh_graph_create(C.byref(self._h)))
        lib.pbh_graph_num_nodes.restype = C.c_int64
        lib.pbh_graph_num_edges.restype = C.c_int64

    def __del__(self):
        try:
            if self._h:
                N.host().pbh_graph_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ---- construction -------------------------------------------------------------------------
    def add_edges(self, src, dst, etype: int = 0, reverse: bool = False) -> None:
        s = np.ascontiguousarray(src, dtype=np.uint64)
        d = np.ascontiguousarray(dst, dtype=np.uint64)
        assert s.shape == d.shape
        N.hcheck(N.host().pbh_graph_add_edges(self._h, etype, _u64p(s), _u64p(d), C.c_int64(s.size), int(reverse)))

    def load_edge_file(self, path: str, etype: int = 0, reverse: bool = False) -> int:
        """Text file ``src<TAB>dst`` per line; ``reverse`` stores dst -> src."""
        n = C.c_int64()
        N.hcheck(N.host().pbh_graph_load_edge_file(self._h, etype, path.encode(), int(reverse), C.byref(n)))
        return n.value

    def build(self, etype: int = 0) -> None:
        N.hcheck(N.host().pbh_graph_build(self._h, etype))

    def num_nodes(self, etype: int = 0) -> int:
        return N.host().pbh_graph_num_nodes(self._h, etype)

    def num_edges(self, etype: int = 0) -> int:
        return N.host().pbh_graph_num_edges(self._h, etype)

    def csr(self, etype: int = 0):
        """(node_ids uint64 [n] sorted, row_ptr int64 [n+1], col int64 [m] neighbour indices)."""
        n, m = self.num_nodes(etype), self.num_edges(etype)
        ids = np.empty(n, np.uint64)
        rp = np.empty(n + 1, np.int64)
        col = np.empty(m, np.int64)
        N.hcheck(N.host().pbh_graph_export_csr(self._h, etype, _u64p(ids), rp.ctypes.data_as(C.POINTER(C.c_int64)),
                                               col.ctypes.data_as(C.POINTER(C.c_int64))))
        return ids, rp, col

    def all_ids(self, etype: int = 0) -> np.ndarray:
        return self.csr(etype)[0]

    # ---- host sampling --------------------------------------------------------------------------
    def sample_neighbors(self, nodes, k: int, etype: int = 0, seed: int = 0):
        """(neighbours uint64 [n, k], counts int32 [n]); without replacement, unused entries 0."""
        q = np.ascontiguousarray(nodes, dtype=np.uint64)
        out = np.empty((q.size, k), np.uint64)
        cnt = np.empty(q.size, np.int32)
        N.hcheck(N.host().pbh_graph_sample_neighbors(self._h, etype, _u64p(q), C.c_int64(q.size), k, C.c_uint64(seed),
                                                     _u64p(out), cnt.ctypes.data_as(C.POINTER(C.c_int32))))
        return out, cnt

    def random_walk(self, starts, walk_len: int, metapath: Sequence[int] = (0,), seed: int = 0):
        """(walks uint64 [n, walk_len], lengths int32 [n]); step s follows edge type
        ``metapath[s % len]``; a walk ends early at a node without out-edges (rest of the row 0)."""
        q = np.ascontiguousarray(starts, dtype=np.uint64)
        mp = (C.c_int * len(metapath))(*metapath)
        out = np.empty((q.size, walk_len), np.uint64)
        ln = np.empty(q.size, np.int32)
        N.hcheck(N.host().pbh_graph_random_walk(self._h, mp, len(metapath), _u64p(q), C.c_int64(q.size), walk_len,
                                                C.c_uint64(seed), _u64p(out), ln.ctypes.data_as(C.POINTER(C.c_int32))))
        return out, ln

    def to_device(self, device=None, etypes: Sequence[int] | None = None) -> "DeviceGraph":
        n_types = N.host().pbh_graph_num_types(self._h)
        return DeviceGraph({t: self.csr(t) for t in (etypes if etypes is not None else range(n_types))}, device)

    # ---- weighted edges ------------------------------------------------------------------------------
    def add_weighted_edges(self, src, dst, weight, etype: int = 0, reverse: bool = False) -> None:
        s = np.ascontiguousarray(src, dtype=np.uint64)
        d = np.ascontiguousarray(dst, dtype=np.uint64)
        w = np.ascontiguousarray(weight, dtype=np.float32)
        assert s.shape == d.shape == w.shape
        N.hcheck(N.host().pbh_graph_add_edges_weighted(self._h, etype, _u64p(s), _u64p(d),
                                                       w.ctypes.data_as(C.POINTER(C.c_float)), C.c_int64(s.size),
                                                       int(reverse)))

    def load_weighted_edge_file(self, path: str, etype: int = 0, reverse: bool = False) -> int:
        """Text file ``src<TAB>dst<TAB>weight`` per line (a missing weight counts as 1)."""
        n = C.c_int64()
        N.hcheck(N.host().pbh_graph_load_weighted_edge_file(self._h, etype, path.encode(), int(reverse), C.byref(n)))
        return n.value

    def is_weighted(self, etype: int = 0) -> bool:
        return bool(N.host().pbh_graph_is_weighted(self._h, etype))

    def edge_weights(self, etype: int = 0) -> np.ndarray:
        """Weight of every CSR edge, aligned with ``csr(etype)[2]``."""
        w = np.empty(self.num_edges(etype), np.float32)
        N.hcheck(N.host().pbh_graph_export_weights(self._h, etype, w.ctypes.data_as(C.POINTER(C.c_float))))
        return w

    def sample_neighbors_weighted(self, nodes, k: int, etype: int = 0, seed: int = 0):
        """Like ``sample_neighbors`` with probability proportional to the edge weight (without replacement)."""
        q = np.ascontiguousarray(nodes, dtype=np.uint64)
        out = np.empty((q.size, k), np.uint64)
        cnt = np.empty(q.size, np.int32)
        N.hcheck(N.host().pbh_graph_sample_neighbors_weighted(self._h, etype, _u64p(q), C.c_int64(q.size), k,
                                                              C.c_uint64(seed), _u64p(out),
                                                              cnt.ctypes.data_as(C.POINTER(C.c_int32))))
        return out, cnt

    # ---- node types with feature slots ---------------------------------------------------------------
    def add_nodes(self, ids, features: Sequence[Sequence] | None = None, ntype: int = 0) -> None:
        """``features[i]`` = list of ``(slot, feasign)`` pairs of node ``ids[i]``."""
        q = np.ascontiguousarray(ids, dtype=np.uint64)
        if features is None:
            N.hcheck(N.host().pbh_graph_add_nodes(self._h, ntype, _u64p(q), C.c_int64(q.size), None, None, None))
            return
        off = np.zeros(q.size + 1, np.int64)
        off[1:] = np.cumsum([len(f) for f in features])
        sl = np.ascontiguousarray([p[0] for f in features for p in f], dtype=np.int32)
        va = np.ascontiguousarray([p[1] for f in features for p in f], dtype=np.uint64)
        N.hcheck(N.host().pbh_graph_add_nodes(self._h, ntype, _u64p(q), C.c_int64(q.size),
                                              off.ctypes.data_as(C.POINTER(C.c_int64)),
                                              sl.ctypes.data_as(C.POINTER(C.c_int32)), _u64p(va)))

    def load_node_file(self, path: str, ntype: int = 0, type_name: str = "", sep: str = " ") -> int:
        """Lines ``<type name>\t<id>[\t<slot>:<v>[<sep><v>...]]...``; only lines of ``type_name`` when given."""
        n = C.c_int64()
        N.hcheck(N.host().pbh_graph_load_node_file(self._h, ntype, path.encode(), type_name.encode(),
                                                   C.c_char(sep.encode()[:1]), C.byref(n)))
        return n.value

    def typed_nodes(self, ntype: int = 0) -> np.ndarray:
        N.host().pbh_graph_num_typed_nodes.restype = C.c_int64
        out = np.empty(N.host().pbh_graph_num_typed_nodes(self._h, ntype), np.uint64)
        N.hcheck(N.host().pbh_graph_export_typed_nodes(self._h, ntype, _u64p(out)))
        return out

    def node_features(self, nodes, slots: Sequence[int], ntype: int = 0) -> np.ndarray:
        """``get_feature_of_nodes``: uint64 [n, len(slots)], the first feasign of every slot (0 = none)."""
        q = np.ascontiguousarray(nodes, dtype=np.uint64)
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        out = np.empty((q.size, sl.size), np.uint64)
        N.hcheck(N.host().pbh_graph_node_features(self._h, ntype, _u64p(q), C.c_int64(q.size),
                                                  sl.ctypes.data_as(C.POINTER(C.c_int32)), int(sl.size), _u64p(out)))
        return out

    def all_feature_ids(self, ntype: int = 0) -> np.ndarray:
        n = C.c_int64()
        N.hcheck(N.host().pbh_graph_all_feature_ids(self._h, ntype, None, C.byref(n)))
        out = np.empty(n.value, np.uint64)
        N.hcheck(N.host().pbh_graph_all_feature_ids(self._h, ntype, _u64p(out), C.byref(n)))
        return out

    def load_node_weight(self, path: str, ntype: int = 0) -> int:
        n = C.c_int64()
        N.hcheck(N.host().pbh_graph_load_node_weight(self._h, ntype, path.encode(), C.byref(n)))
        return n.value

    def node_weights(self, nodes, ntype: int = 0) -> np.ndarray:
        q = np.ascontiguousarray(nodes, dtype=np.uint64)
        out = np.empty(q.size, np.float32)
        N.hcheck(N.host().pbh_graph_node_weights(self._h, ntype, _u64p(q), C.c_int64(q.size),
                                                 out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    # ---- partitions (paging a graph that does not fit in HBM) ---------------------------------------
    def make_partitions(self, etype: int, byte_size: int, device_len: int = 1) -> int:
        n = C.c_int()
        N.hcheck(N.host().pbh_graph_make_partitions(self._h, etype, C.c_int64(byte_size), device_len, C.byref(n)))
        return n.value

    def partition_num(self, etype: int = 0) -> int:
        return N.host().pbh_graph_partition_num(self._h, etype)

    def next_partition(self, etype: int = 0) -> int:
        """Index of the partition to page in next; -1 once after the last (then the cycle restarts)."""
        return N.host().pbh_graph_next_partition(self._h, etype)

    def partition(self, etype: int, part: int):
        """(node_ids uint64 [n], row_ptr int64 [n+1], neighbour ids uint64 [m], weights fp32 [m] | None)."""
        nn, ne = C.c_int64(), C.c_int64()
        N.hcheck(N.host().pbh_graph_partition_shape(self._h, etype, part, C.byref(nn), C.byref(ne)))
        ids = np.empty(nn.value, np.uint64)
        rp = np.empty(nn.value + 1, np.int64)
        col = np.empty(ne.value, np.uint64)
        w = np.empty(ne.value, np.float32) if self.is_weighted(etype) else None
        N.hcheck(N.host().pbh_graph_export_partition(self._h, etype, part, _u64p(ids),
                                                     rp.ctypes.data_as(C.POINTER(C.c_int64)), _u64p(col),
                                                     w.ctypes.data_as(C.POINTER(C.c_float)) if w is not None else None))
        return ids, rp, col, w

    def export_partition_files(self, etype: int, directory: str) -> None:
        import os
        os.makedirs(directory, exist_ok=True)
        N.hcheck(N.host().pbh_graph_export_partition_files(self._h, etype, directory.encode()))


class DeviceGraph:
    """CSR of every edge type as tensors on one device; sampling and walking are gathers there.
    Node ids are int64 on the device (the uint64 bit pattern, as for every key)."""

    def __init__(self, csr_by_type: dict, device=None):
        self.device = torch.device(device) if device is not None else torch.device(
            "cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.types = {}
        for t, (ids, rp, col) in csr_by_type.items():
            self.types[t] = (torch.from_numpy(ids.view(np.int64)).to(self.device),
                             torch.from_numpy(rp).to(self.device), torch.from_numpy(col).to(self.device))

    def index_of(self, nodes: torch.Tensor, etype: int = 0):
        """(index int64 [n], known bool [n]).  Ids are sorted as uint64; int64 order differs only for
        ids >= 2^63, which are mapped through an order-preserving flip of the sign bit."""
        ids = self.types[etype][0]
        flip = torch.iinfo(torch.int64).min
        idx = torch.searchsorted(ids ^ flip, nodes ^ flip).clamp_(max=max(ids.numel() - 1, 0))
        known = ids[idx] == nodes if ids.numel() else torch.zeros_like(nodes, dtype=torch.bool)
        return idx, known

    def degree(self, nodes: torch.Tensor, etype: int = 0) -> torch.Tensor:
        _, rp, _ = self.types[etype]
        idx, known = self.index_of(nodes, etype)
        return torch.where(known, rp[idx + 1] - rp[idx], torch.zeros_like(idx))

    def sample_neighbors(self, nodes: torch.Tensor, k: int, etype: int = 0, generator=None):
        """(neighbours int64 [n, k], counts int64 [n]).  Nodes with <= k neighbours return all of them
        (in id order); larger neighbourhoods are sampled uniformly WITH replacement (one gather per
        sample instead of a per-node shuffle).  Unused entries are 0."""
        ids, rp, col = self.types[etype]
        idx, known = self.index_of(nodes, etype)
        deg = torch.where(known, rp[idx + 1] - rp[idx], torch.zeros_like(idx))
        n = nodes.numel()
        j = torch.arange(k, device=self.device).expand(n, k)
        r = torch.rand(n, k, device=self.device, generator=generator)
        pick = torch.where((deg > k)[:, None], (r * deg[:, None]).long(), j)
        cnt = deg.clamp(max=k)
        valid = j < cnt[:, None]
        pos = (rp[idx][:, None] + torch.minimum(pick, (deg - 1).clamp(min=0)[:, None])).clamp_(max=max(col.numel() - 1, 0))
        nb = ids[col[pos]] if col.numel() else torch.zeros(n, k, dtype=torch.int64, device=self.device)
        return torch.where(valid, nb, torch.zeros_like(nb)), cnt

    def random_walk(self, starts: torch.Tensor, walk_len: int, metapath: Sequence[int] = (0,), generator=None):
        """(walks int64 [n, walk_len], lengths int64 [n]); same contract as ``GraphTable.random_walk``."""
        n = starts.numel()
        walks = torch.zeros(n, walk_len, dtype=torch.int64, device=self.device)
        walks[:, 0] = starts
        alive = torch.ones(n, dtype=torch.bool, device=self.device)
        length = torch.ones(n, dtype=torch.int64, device=self.device)
        cur = starts
        for s in range(1, walk_len):
            ids, rp, col = self.types[metapath[(s - 1) % len(metapath)]]
            idx, known = self.index_of(cur, metapath[(s - 1) % len(metapath)])
            deg = torch.where(known, rp[idx + 1] - rp[idx], torch.zeros_like(idx))
            alive = alive & (deg > 0)
            r = torch.rand(n, device=self.device, generator=generator)
            pos = (rp[idx] + torch.minimum((r * deg).long(), (deg - 1).clamp(min=0))).clamp_(max=max(col.numel() - 1, 0))
            nxt = ids[col[pos]] if col.numel() else cur
            cur = torch.where(alive, nxt, cur)
            walks[:, s] = torch.where(alive, cur, torch.zeros_like(cur))
            length += alive.long()
        return walks, length


def walk_pairs(walks: torch.Tensor, lengths: torch.Tensor, window: int):
    """Skip-gram pairs of every walk: (centre, context) for |i - j| <= window, i != j, both inside
    the walk's valid prefix (GraphDataGenerator's pair kernel, data_feed.cu:260-330)."""
    n, L = walks.shape
    i = torch.arange(L, device=walks.device)
    off = torch.cat([torch.arange(-window, 0, device=walks.device), torch.arange(1, window + 1, device=walks.device)])
    ci = i[:, None].expand(L, off.numel())
    cj = ci + off[None, :]
    ok = (cj >= 0) & (cj < L)
    ci, cj = ci[ok], cj[ok]                                  # [P] position pairs shared by all walks
    valid = (ci[None, :] < lengths[:, None]) & (cj[None, :] < lengths[:, None])
    return walks[:, ci][valid], walks[:, cj][valid]


class GraphDataGenerator:
    """Walks -> skip-gram mini-batches with sampled negatives.

    Every batch holds ``batch_size`` rows: ``1/(1+neg)`` of them positive pairs from the walks,
    the rest the same centres with a uniformly drawn node as context (label 0).  Rows are shuffled.
    ``batches()`` yields ``SlotBatch`` objects (2 slots, one key per bag, ``dense[:, 0]`` = label)."""

    def __init__(self, graph: "DeviceGraph | GraphTable", walk_len: int = 8, window: int = 2, batch_size: int = 1024,
                 neg: int = 2, metapath: Sequence[int] = (0,), walks_per_node: int = 1, seed: int = 0, device=None):
        self.graph = graph if isinstance(graph, DeviceGraph) else graph.to_device(device)
        self.walk_len, self.window, self.batch_size, self.neg = walk_len, window, batch_size, neg
        self.metapath, self.walks_per_node = tuple(metapath), walks_per_node
        self.gen = torch.Generator(device=self.graph.device).manual_seed(seed)

    def start_nodes(self) -> torch.Tensor:
        ids = self.graph.types[self.metapath[0]][0]
        perm = torch.randperm(ids.numel(), device=ids.device, generator=self.gen)
        return ids[perm].repeat(self.walks_per_node)

    def pairs(self):
        walks, lens = self.graph.random_walk(self.start_nodes(), self.walk_len, self.metapath, generator=self.gen)
        return walk_pairs(walks, lens, self.window)

    def all_keys(self) -> torch.Tensor:
        """Every node id (for the feed pass)."""
        return torch.unique(torch.cat([t[0] for t in self.graph.types.values()]))

    def batches(self) -> Iterator:
        from .ops.pull_box_sparse import SlotBatch
        dev = self.graph.device
        c, x = self.pairs()
        ids = self.graph.types[self.metapath[-1]][0]
        P = c.numel()
        if self.neg:
            nc = c.repeat(self.neg)
            nx = ids[torch.randint(ids.numel(), (P * self.neg,), device=dev, generator=self.gen)]
            c, x = torch.cat([c, nc]), torch.cat([x, nx])
        label = torch.zeros(c.numel(), device=dev)
        label[:P] = 1.0
        perm = torch.randperm(c.numel(), device=dev, generator=self.gen)
        c, x, label = c[perm], x[perm], label[perm]
        B = self.batch_size
        for b in range(0, c.numel() - B + 1, B):
            keys = torch.cat([c[b:b + B], x[b:b + B]])             # slot-major: slot 0 bags, then slot 1
            lod = torch.arange(2 * B + 1, dtype=torch.int32, device=dev)
            yield SlotBatch(keys, lod, B, 2, label[b:b + B, None].contiguous(), n_keys=2 * B)


class GraphGpuWrapper:
    """The reference's method names over ``GraphTable`` / ``DeviceGraph``
    (graph_gpu_wrapper.h:34-99)."""
    _ins = None

    @classmethod
    def get_instance(cls) -> "GraphGpuWrapper":
        if cls._ins is None:
            cls._ins = cls()
        return cls._ins

    def __init__(self):
        self.table = GraphTable()
        self.edge_types: dict = {}
        self.node_types: dict = {}
        self.feature_separator = " "
        self.resident_partition = -1
        self.device_graph: DeviceGraph | None = None

    def set_up_types(self, edge_types: Sequence[str], node_types: Sequence[str] = ()) -> None:
        self.edge_types = {name: i for i, name in enumerate(edge_types)}
        self.node_types = {name: i for i, name in enumerate(node_types)}

    def load_edge_file(self, name: str, filepath: str, reverse: bool = False) -> int:
        t = self.edge_types.setdefault(name, len(self.edge_types))
        return self.table.load_edge_file(filepath, t, reverse)

    def upload_batch(self, device=None) -> DeviceGraph:
        """Build every edge type and place it in HBM."""
        for t in self.edge_types.values():
            self.table.build(t)
        self.device_graph = self.table.to_device(device, list(self.edge_types.values()))
        return self.device_graph

    def get_all_id(self, name: str) -> np.ndarray:
        return self.table.all_ids(self.edge_types[name])

    # ---- node files / features / partitions: the rest of graph_gpu_wrapper.h:50-99 ------------------------
    def load_node_file(self, name: str, filepath: str) -> int:
        t = self.node_types.setdefault(name, len(self.node_types))
        return self.table.load_node_file(filepath, t, name, self.feature_separator)

    def set_feature_separator(self, ch: str) -> None:
        self.feature_separator = ch

    def get_feature_of_nodes(self, name: str, nodes, slots: Sequence[int]) -> np.ndarray:
        return self.table.node_features(nodes, slots, self.node_types[name])

    def get_all_feature_ids(self, name: str) -> np.ndarray:
        return self.table.all_feature_ids(self.node_types[name])

    def load_node_weight(self, name: str, path: str) -> int:
        return self.table.load_node_weight(path, self.node_types[name])

    def make_partitions(self, name: str, byte_size: int, device_len: int = 1) -> int:
        t = self.edge_types[name]
        self.table.build(t) if self.table.num_nodes(t) == 0 else None
        return self.table.make_partitions(t, byte_size, device_len)

    def get_partition_num(self, name: str) -> int:
        return self.table.partition_num(self.edge_types[name])

    def get_partition(self, name: str, num: int) -> np.ndarray:
        return self.table.partition(self.edge_types[name], num)[0]

    def load_next_partition(self, name: str, device=None) -> int:
        """Pages the next partition of the edge type into HBM (it replaces the resident one); returns its index,
        -1 when the cycle is complete.  Neighbour ids that live in another partition stay valid ids: a walk that
        leaves the resident partition ends there, as with the reference's paged graph."""
        t = self.edge_types[name]
        p = self.table.next_partition(t)
        if p < 0:
            return -1
        ids, rp, col_ids, _ = self.table.partition(t, p)
        # resident CSR in the partition's own index space; edges to neighbours outside the partition are dropped
        if ids.size and col_ids.size:
            pos = np.minimum(np.searchsorted(ids, col_ids), ids.size - 1)
            keep = ids[pos] == col_ids
            row_of_edge = np.repeat(np.arange(ids.size), np.diff(rp))
            new_rp = np.zeros(ids.size + 1, np.int64)
            new_rp[1:] = np.cumsum(np.bincount(row_of_edge[keep], minlength=ids.size))
            col = pos.astype(np.int64)
        else:
            keep = np.zeros(col_ids.size, bool)
            new_rp = np.zeros(ids.size + 1, np.int64)
            col = np.zeros(col_ids.size, np.int64)
        self.device_graph = DeviceGraph({t: (ids, new_rp, col[keep])}, device)
        self.resident_partition = p
        return p

    def export_partition_files(self, name: str, file_path: str) -> None:
        self.table.export_partition_files(self.edge_types[name], file_path)

    def graph_neighbor_sample(self, name: str, nodes, sample_size: int):
        t = self.edge_types[name]
        if self.device_graph is not None and torch.is_tensor(nodes):
            return self.device_graph.sample_neighbors(nodes, sample_size, t)
        return self.table.sample_neighbors(nodes, sample_size, t)

    # ---- service set-up names (graph_gpu_wrapper.cu:107-260) -----------------------------------------------
    def set_device(self, device_ids: Sequence[int]) -> None:
        """Devices of this process (one per process here: the first entry is used by ``upload_batch``)."""
        self.device_ids = list(device_ids)

    def init_search_level(self, level: int) -> None:
        """0 = the whole graph resident in HBM, 1 = partitions paged from host memory, 2 = additionally the
        complementary graph of hot neighbours (``make_complementary_graph``)."""
        assert level in (0, 1, 2), level
        self.search_level = int(level)

    set_search_level = init_search_level

    def init_service(self) -> None:
        """The reference builds its table proto and shard pool here; the native store needs no set-up — this
        validates what was declared (every feature table names a declared node type)."""
        for tname in getattr(self, "feat_conf", {}):
            if tname not in self.node_types:
                raise ValueError(f"add_table_feat_conf names node type {tname!r}, set_up_types did not")
        self.service_ready = True

    def add_table_feat_conf(self, table_name: str, feat_name: str, feat_dtype: str, feat_shape: int) -> int:
        """Declare feature ``feat_name`` of node type ``table_name``; returns its slot index (first come, first
        numbered; declaring a name again overrides its dtype / shape, graph_gpu_wrapper.cu:163-190)."""
        if table_name not in self.node_types:
            return -1
        conf = self.__dict__.setdefault("feat_conf", {}).setdefault(table_name, {})
        slot = conf[feat_name][0] if feat_name in conf else len(conf)
        conf[feat_name] = (slot, feat_dtype, int(feat_shape))
        return slot

    def feature_slot(self, table_name: str, feat_name: str) -> int:
        return self.feat_conf[table_name][feat_name][0]

    @staticmethod
    def inverse_etype(etype: str) -> str:
        a, _, b = etype.partition("2")
        return f"{b}2{a}" if b else etype

    def load_node_and_edge(self, etype: str, ntype: str, epath: str, npath: str, part_num: int = 0,
                           reverse: bool = False) -> dict:
        """Edge types ``"u2i,i2i"`` from ``epath/<etype>/*`` and node types ``"u,i"`` from the files of ``npath``,
        at most ``part_num`` files of each directory (0 = all); ``reverse`` also loads every edge file backwards
        as the inverse type (``u2i`` -> ``i2u``) (common_graph_table.cc:1081-1145)."""
        import os
        out = {}

        def files_of(d):
            fs = sorted(os.path.join(d, f) for f in os.listdir(d) if os.path.isfile(os.path.join(d, f)))
            return fs[:part_num] if 0 < part_num < len(fs) else fs

        for et in [e for e in etype.split(",") if e]:
            n = 0
            for f in files_of(os.path.join(epath, et)):
                n += self.load_edge_file(et, f, False)
                if reverse:
                    self.load_edge_file(self.inverse_etype(et), f, True)
            out[et] = n
        node_files = files_of(npath) if ntype else []
        for nt in [t for t in ntype.split(",") if t]:
            out[nt] = sum(self.load_node_file(nt, f) for f in node_files)
        return out

    def query_node_list(self, gpu_id: int, name: str, start: int, query_size: int) -> np.ndarray:
        """``query_size`` node ids of the type from position ``start`` of its id list (the reference pages
        through a device's node list this way, graph_gpu_wrapper.h:84-87)."""
        ids = self.get_all_id(name) if name in self.edge_types else self.table.typed_nodes(self.node_types[name])
        return ids[int(start):int(start) + int(query_size)]

    def neighbor_sample(self, name: str, nodes, sample_size: int):
        return self.graph_neighbor_sample(name, nodes, sample_size)

    def make_complementary_graph(self, name: str, byte_size: int, device=None) -> np.ndarray:
        """Search level 2: next to the resident partition keep the adjacency lists of the nodes that the
        resident edges point to most often, as many as fit ``byte_size`` (8 bytes per edge) — walks that leave
        the partition mostly land there.  Returns the chosen node ids, most referenced first, and makes
        ``complementary_graph`` a ``DeviceGraph`` of their rows (ids stay global; rows of other nodes are absent)
        (reference: GraphTable::make_complementary_graph, common_graph_table.cc:556-620)."""
        t = self.edge_types[name]
        ids, rp, col = self.table.csr(t)
        if self.resident_partition >= 0:
            pids, prp, pcol_ids, _ = self.table.partition(t, self.resident_partition)
            targets = pcol_ids
        else:
            pids, targets = ids, ids[col]
        if targets.size == 0:
            self.complementary_graph = None
            return np.zeros(0, np.uint64)
        uniq, cnt = np.unique(targets, return_counts=True)
        outside = ~np.isin(uniq, pids) if self.resident_partition >= 0 else np.ones(uniq.size, bool)
        uniq, cnt = uniq[outside], cnt[outside]
        order = np.lexsort((uniq, -cnt))                       # most referenced first, ties by id
        uniq = uniq[order]
        pos = np.searchsorted(ids, uniq)
        ok = (pos < ids.size) & (ids[np.minimum(pos, ids.size - 1)] == uniq)
        uniq, pos = uniq[ok], pos[ok]
        deg = (rp[pos + 1] - rp[pos]).astype(np.int64)
        fits = np.cumsum(deg * 8) <= int(byte_size)
        chosen, cpos = uniq[fits], pos[fits]
        srt = np.argsort(chosen)
        c_ids, c_pos = chosen[srt], cpos[srt]
        c_rp = np.zeros(c_ids.size + 1, np.int64)
        c_rp[1:] = np.cumsum(rp[c_pos + 1] - rp[c_pos])
        c_col_ids = np.concatenate([ids[col[rp[p]:rp[p + 1]]] for p in c_pos]) if c_ids.size else np.zeros(0, np.uint64)
        # neighbours that are themselves in the complementary set are indexable; the rest end a walk here
        npos = np.minimum(np.searchsorted(c_ids, c_col_ids), max(c_ids.size - 1, 0))
        keep = c_ids[npos] == c_col_ids if c_ids.size else np.zeros(0, bool)
        row = np.repeat(np.arange(c_ids.size), np.diff(c_rp))
        k_rp = np.zeros(c_ids.size + 1, np.int64)
        k_rp[1:] = np.cumsum(np.bincount(row[keep], minlength=c_ids.size))
        self.complementary_graph = DeviceGraph({t: (c_ids, k_rp, npos[keep].astype(np.int64))}, device)
        self.complementary_ids = chosen
        return chosen

    def finalize(self) -> None:
        self.device_graph = None
        self.table = GraphTable()
        self.edge_types = {}
