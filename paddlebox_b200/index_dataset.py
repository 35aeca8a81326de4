"""Tree index for tree-based deep models (TDM): the index file, its queries and the layer-wise sampler.

Reference: ``paddle/fluid/distributed/index_dataset/`` (``TreeIndex`` / ``IndexWrapper`` index_wrapper.{h,cc},
``LayerWiseSampler`` / ``BeamSearchSampler`` index_sampler.{h,cc}, ``index_dataset.proto``), the Python wrapper
``python/paddle/distributed/fleet/dataset/index_dataset.py`` and ``Dataset.tdm_sample`` (data_set.cc ``TDMSample``).
A tree file is a sequence of ``[int32 length][KVItem]`` records; ``KVItem {bytes key = 1; bytes value = 2}``; the key
``".tree_meta"`` holds ``TreeMeta {int32 height = 1; int32 branch = 2}``, every other key is a node code in decimal
with ``IndexNode {uint64 id = 1; bool is_leaf = 2; float probability = 3; string item_name = 4}``.  Codes number a
complete ``branch``-ary tree in breadth-first order (root 0, children of c are ``c * branch + 1 ... + branch``).
The file is read and written here without protobuf (the four messages are a dozen lines of wire format).  Queries
run on sorted numpy arrays; the sampler draws all negatives of a layer for the whole batch at once.  The tensor
ops over a tree (``tdm_child`` / ``tdm_sampler``) are in ops/tdm.py.
"""
from __future__ import annotations

import struct

import numpy as np

__all__ = ["IndexNode", "TreeIndex", "IndexWrapper", "IndexSampler", "save_tree_index"]


# ---- protobuf wire format of the four messages -------------------------------------------------------------------
def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def _read_varint(buf: bytes, pos: int):
    shift = val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _fields(buf: bytes):
    """(field number, wire type, value) of one message; value is int (varint / fixed) or bytes."""
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        num, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            v, pos = buf[pos:pos + n], pos + n
            if len(v) != n:
                raise ValueError("tree index: truncated field")
        elif wt == 5:
            v, pos = buf[pos:pos + 4], pos + 4
        elif wt == 1:
            v, pos = buf[pos:pos + 8], pos + 8
        else:
            raise ValueError(f"tree index: unsupported wire type {wt}")
        yield num, wt, v


def _ld(num: int, payload: bytes) -> bytes:
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


class IndexNode:
    """One tree node; the accessors are methods as in the reference's binding."""
    __slots__ = ("_id", "_is_leaf", "_probability", "_item_name")

    def __init__(self, id: int = 0, is_leaf: bool = False, probability: float = 0.0, item_name: str = ""):
        self._id, self._is_leaf, self._probability, self._item_name = int(id), bool(is_leaf), float(probability), item_name

    def id(self) -> int:
        return self._id

    def is_leaf(self) -> bool:
        return self._is_leaf

    def probability(self) -> float:
        return self._probability

    def item_name(self) -> str:
        return self._item_name

    def __repr__(self):
        return f"IndexNode(id={self._id}, is_leaf={self._is_leaf}, probability={self._probability})"

    def _encode(self) -> bytes:
        out = _varint(1 << 3) + _varint(self._id) + _varint(2 << 3) + _varint(int(self._is_leaf)) + \
            _varint((3 << 3) | 5) + struct.pack("<f", self._probability)
        if self._item_name:
            out += _ld(4, self._item_name.encode())
        return out

    @classmethod
    def _decode(cls, buf: bytes) -> "IndexNode":
        n = cls()
        want = {1: 0, 2: 0, 3: 5, 4: 2}
        for num, wt, v in _fields(buf):
            if num in want and wt != want[num] or (wt == 5 and len(v) != 4):
                raise ValueError(f"tree index: field {num} of a node has wire type {wt}")
            if num == 1:
                n._id = v
            elif num == 2:
                n._is_leaf = bool(v)
            elif num == 3:
                n._probability = struct.unpack("<f", v)[0]
            elif num == 4:
                n._item_name = v.decode()
        return n


def save_tree_index(path: str, nodes: dict, height: int, branch: int) -> None:
    """Write a tree file: ``nodes`` maps code -> ``IndexNode`` (or ``(id, is_leaf[, probability[, item_name]])``)."""
    with open(path, "wb") as f:
        def put(key: bytes, value: bytes):
            item = _ld(1, key) + _ld(2, value)
            f.write(struct.pack("<i", len(item)) + item)
        put(b".tree_meta", _varint(1 << 3) + _varint(height) + _varint(2 << 3) + _varint(branch))
        for code in sorted(nodes):
            n = nodes[code]
            put(str(int(code)).encode(), (n if isinstance(n, IndexNode) else IndexNode(*n))._encode())


class _Tree:
    """The loaded index (``TreeIndex`` of index_wrapper.h:33-90)."""

    def __init__(self, path: str):
        try:
            self._load(path)
        except (IndexError, struct.error, UnicodeDecodeError, OverflowError) as e:
            raise ValueError(f"tree index {path}: damaged file ({type(e).__name__}: {e})") from None

    def _load(self, path: str) -> None:
        codes, nodes = [], []
        self._height = self._branch = 0
        with open(path, "rb") as f:
            buf = f.read()
        pos = 0
        while pos + 4 <= len(buf):
            n, = struct.unpack_from("<i", buf, pos)
            pos += 4
            if n <= 0:
                break
            if pos + n > len(buf):
                raise ValueError(f"tree index {path}: record of {n} bytes runs past the end of the file")
            key = value = b""
            for num, wt, v in _fields(buf[pos:pos + n]):
                if num in (1, 2) and wt != 2:
                    raise ValueError(f"tree index {path}: key / value field with wire type {wt}")
                if num == 1:
                    key = v
                elif num == 2:
                    value = v
            pos += n
            if key == b".tree_meta":
                for num, wt, v in _fields(value):
                    if wt != 0:
                        raise ValueError(f"tree index {path}: .tree_meta field {num} has wire type {wt}")
                    if num == 1:
                        self._height = v
                    elif num == 2:
                        self._branch = v
            else:
                codes.append(int(key))
                nodes.append(IndexNode._decode(value))
        if not 2 <= self._branch <= 1 << 16 or not 1 <= self._height <= 64:
            raise ValueError(f"tree index {path}: no valid .tree_meta (height {self._height}, branch {self._branch})")
        if codes and (min(codes) < 0 or max(codes) >= 1 << 63):
            raise ValueError(f"tree index {path}: node code out of range")
        order = np.argsort(np.asarray(codes, dtype=np.uint64), kind="stable")
        self._codes = np.asarray(codes, dtype=np.uint64)[order]
        self._nodes = [nodes[i] for i in order]
        self._node_ids = np.asarray([n._id for n in self._nodes], dtype=np.uint64)
        leaf = np.asarray([n._is_leaf for n in self._nodes], dtype=bool)
        lid, lcode = self._node_ids[leaf], self._codes[leaf]
        o = np.argsort(lid, kind="stable")
        self._leaf_ids, self._leaf_codes = lid[o], lcode[o]
        self._max_code = int(self._codes.max()) + 1 if self._codes.size else 1
        self._fake = IndexNode(0, False, 0.0)

    # ---- shape ---------------------------------------------------------------------------------------------------
    def height(self) -> int:
        return self._height

    def branch(self) -> int:
        return self._branch

    def total_node_nums(self) -> int:
        return int(self._codes.size)

    def emb_size(self) -> int:
        """Rows an embedding table over node ids needs (largest id + 1)."""
        return int(self._node_ids.max()) + 1 if self._node_ids.size else 1

    # ---- queries -------------------------------------------------------------------------------------------------
    def _find(self, codes: np.ndarray):
        codes = np.asarray(codes, dtype=np.uint64).reshape(-1)
        at = np.searchsorted(self._codes, codes)
        ok = (at < self._codes.size) & (self._codes[np.minimum(at, max(self._codes.size - 1, 0))] == codes) \
            if self._codes.size else np.zeros(codes.size, bool)
        return at, ok

    def node_ids_of(self, codes) -> np.ndarray:
        """Node ids of ``codes`` (0 where there is no such node) — the array form of ``get_nodes``."""
        at, ok = self._find(codes)
        return np.where(ok, self._node_ids[np.minimum(at, max(self._codes.size - 1, 0))], 0).astype(np.uint64)

    def get_nodes(self, codes) -> list:
        at, ok = self._find(codes)
        return [self._nodes[a] if o else self._fake for a, o in zip(at.tolist(), ok.tolist())]

    def get_all_leafs(self) -> list:
        at, _ = self._find(self._leaf_codes)
        return [self._nodes[a] for a in at.tolist()]

    def get_layer_codes(self, level: int) -> list:
        lo = (self._branch ** level - 1) // (self._branch - 1)
        hi = (self._branch ** (level + 1) - 1) // (self._branch - 1)
        return self._codes[(self._codes >= lo) & (self._codes < hi)].tolist()

    def codes_of_ids(self, ids) -> np.ndarray:
        """Leaf codes of item ids; ``max_code`` (one past the largest code) for an unknown id."""
        ids = np.asarray(ids, dtype=np.uint64).reshape(-1)
        if not self._leaf_ids.size:
            return np.full(ids.size, self._max_code, dtype=np.int64)
        at = np.minimum(np.searchsorted(self._leaf_ids, ids), self._leaf_ids.size - 1)
        return np.where(self._leaf_ids[at] == ids, self._leaf_codes[at].astype(np.int64), self._max_code)

    def ancestor_codes(self, ids, level: int) -> np.ndarray:
        code = self.codes_of_ids(ids)
        known = code != self._max_code
        cur = code.copy()
        for _ in range(max(self._height - 1 - level, 0) if level >= 0 else 0):
            cur = np.where(known, (cur - 1) // self._branch, cur)
        return cur

    def get_ancestor_codes(self, ids, level: int) -> list:
        return self.ancestor_codes(ids, level).tolist()

    def get_children_codes(self, ancestor: int, level: int) -> list:
        """The descendants of ``ancestor`` that sit on ``level``."""
        b = self._branch
        lo = (b ** level - 1) // (b - 1)
        hi = (b ** (level + 1) - 1) // (b - 1)
        first = last = int(ancestor)
        while first < lo:                                  # the code range of the subtree, one level down per step
            first, last = first * b + 1, last * b + b
        if first >= hi:
            return [int(ancestor)] if lo <= ancestor < hi else []
        return self._codes[(self._codes >= first) & (self._codes <= last)].tolist()

    def get_travel_codes(self, id: int, start_level: int = 0) -> list:
        """Leaf-to-``start_level`` path of an item: its code, its parent's ..."""
        code = int(self.codes_of_ids([id])[0])
        if code == self._max_code:
            raise ValueError(f"id = {id} doesn't exist in Tree.")
        out, level = [], self._height - 1
        while level >= start_level:
            out.append(code)
            code = (code - 1) // self._branch
            level -= 1
        return out


class IndexWrapper:
    """Process-wide registry of loaded trees (index_wrapper.h:92-125)."""
    _trees: dict = {}

    def insert_tree_index(self, name: str, tree_path: str) -> None:
        if name not in IndexWrapper._trees:
            IndexWrapper._trees[name] = _Tree(tree_path)

    def get_tree_index(self, name: str) -> _Tree:
        try:
            return IndexWrapper._trees[name]
        except KeyError:
            raise ValueError(f"tree [{name}] doesn't exist. Please insert it firstly by API[insert_tree_index]") from None

    def clear_tree(self) -> None:
        IndexWrapper._trees.clear()


class IndexSampler:
    """``IndexSampler("by_layerwise", tree_name)`` (index_sampler.h:30-139).  Layer-wise sampling: for a target
    item, every node on its path from the leaf up to ``start_sample_layer`` is a positive, and ``count`` nodes
    drawn uniformly from the same layer (never the positive itself) are its negatives.  ``"by_beamsearch"`` is
    declared but empty in the reference (its ``sample`` returns nothing); the same here."""

    def __init__(self, kind: str, name: str):
        if kind not in ("by_layerwise", "by_beamsearch"):
            raise ValueError(f"unsupported IndexSampler type {kind!r}")
        self.kind = kind
        self.tree = IndexWrapper().get_tree_index(name)
        self._layer_counts: list = []
        self._layer_codes: list = []
        self._start = 1
        self._rng = np.random.RandomState(0)

    def init_beamsearch_conf(self, k: int) -> None:
        self._beam = int(k)

    def init_layerwise_conf(self, layer_sample_counts, start_sample_layer: int = 1, seed: int = 0) -> None:
        t = self.tree
        if not 0 < start_sample_layer < t.height():
            raise ValueError(f"start sampler layer = [{start_sample_layer}], it should be greater than 0 and less than "
                             f"max_layer, which is [{t.height()}].")
        self._start = int(start_sample_layer)
        counts = [int(layer_sample_counts[i]) if i < len(layer_sample_counts) else 1
                  for i in range(t.height() - self._start)]
        self._layer_counts = counts[::-1]                                     # index 0 = the leaf layer
        self._layer_codes = [np.asarray(t.get_layer_codes(level), dtype=np.int64)
                             for level in range(t.height() - 1, self._start - 1, -1)]
        for j, (c, codes) in enumerate(zip(self._layer_counts, self._layer_codes)):
            if c > 0 and codes.size < 2:
                raise ValueError(f"layer {t.height() - 1 - j} has {codes.size} node(s): no negative can be drawn from it")
        self._rng = np.random.RandomState(int(seed))

    @property
    def layer_counts_sum(self) -> int:
        return sum(self._layer_counts) + len(self._layer_counts)

    def sample_codes(self, target_ids):
        """``(codes int64 [n, rows_per_target], labels uint8 [n, rows_per_target])``: per target, layer after layer
        from the leaf up, the positive's code then that layer's negatives."""
        t = self.tree
        leaf = t.codes_of_ids(target_ids)
        if bool((leaf == t._max_code).any()):
            bad = np.asarray(target_ids).reshape(-1)[leaf == t._max_code][0]
            raise ValueError(f"id = {int(bad)} doesn't exist in Tree.")
        n = leaf.size
        codes = np.empty((n, self.layer_counts_sum), dtype=np.int64)
        labels = np.zeros((n, self.layer_counts_sum), dtype=np.uint8)
        col, cur = 0, leaf.copy()
        for cnt, layer in zip(self._layer_counts, self._layer_codes):
            codes[:, col] = cur
            labels[:, col] = 1
            if cnt:
                at = np.searchsorted(layer, cur)                              # the positive's place in its layer
                draw = self._rng.randint(0, layer.size - 1, size=(n, cnt))
                draw += draw >= at[:, None]                                   # skip the positive: uniform over the rest
                codes[:, col + 1:col + 1 + cnt] = layer[draw]
            col += 1 + cnt
            cur = (cur - 1) // t.branch()
        return codes, labels

    def sample(self, user_inputs, target_ids, with_hierarchy: bool = False) -> list:
        """Rows ``[user features..., node id, label]`` (index_sampler.cc:22-73); with ``with_hierarchy`` the user
        features (item ids) of the rows above the leaf layer are replaced by their ancestors on that layer."""
        if self.kind != "by_layerwise":
            return []
        t = self.tree
        users = np.asarray(user_inputs, dtype=np.uint64).reshape(len(target_ids), -1)
        n, F = users.shape
        codes, labels = self.sample_codes(target_ids)
        out = np.empty((n, self.layer_counts_sum, F + 2), dtype=np.uint64)
        out[:, :, :F] = users[:, None, :]
        out[:, :, F] = t.node_ids_of(codes.reshape(-1)).reshape(codes.shape)
        out[:, :, F + 1] = labels
        if with_hierarchy:
            col = 0
            for j, cnt in enumerate(self._layer_counts):
                if j > 0:
                    anc = t.ancestor_codes(users.reshape(-1), t.height() - j - 1)
                    out[:, col:col + 1 + cnt, :F] = t.node_ids_of(anc).reshape(n, 1, F)
                col += 1 + cnt
        return out.reshape(n * self.layer_counts_sum, F + 2).tolist()


class Index:
    def __init__(self, name: str):
        self._name = name


class TreeIndex(Index):
    """``paddle.distributed.fleet.dataset.TreeIndex(name, path)`` (index_dataset.py:26-97)."""

    def __init__(self, name: str, path: str):
        super().__init__(name)
        self._wrapper = IndexWrapper()
        self._wrapper.insert_tree_index(name, path)
        self._tree = self._wrapper.get_tree_index(name)
        self._height, self._branch = self._tree.height(), self._tree.branch()
        self._total_node_nums, self._emb_size = self._tree.total_node_nums(), self._tree.emb_size()
        self._layerwise_sampler = None

    def height(self) -> int:
        return self._height

    def branch(self) -> int:
        return self._branch

    def total_node_nums(self) -> int:
        return self._total_node_nums

    def emb_size(self) -> int:
        return self._emb_size

    def get_all_leafs(self) -> list:
        return self._tree.get_all_leafs()

    def get_nodes(self, codes) -> list:
        return self._tree.get_nodes(codes)

    def get_layer_codes(self, level: int) -> list:
        return self._tree.get_layer_codes(level)

    def get_travel_codes(self, id: int, start_level: int = 0) -> list:
        return self._tree.get_travel_codes(id, start_level)

    def get_ancestor_codes(self, ids, level: int) -> list:
        return self._tree.get_ancestor_codes(ids, level)

    def get_children_codes(self, ancestor: int, level: int) -> list:
        return self._tree.get_children_codes(ancestor, level)

    def get_travel_path(self, child: int, ancestor: int) -> list:
        res = []
        while child > ancestor:
            res.append(child)
            child = (child - 1) // self._branch
        return res

    def get_pi_relation(self, ids, level: int) -> dict:
        return dict(zip(ids, self.get_ancestor_codes(ids, level)))

    def init_layerwise_sampler(self, layer_sample_counts, start_sample_layer: int = 1, seed: int = 0) -> None:
        assert self._layerwise_sampler is None
        self._layerwise_sampler = IndexSampler("by_layerwise", self._name)
        self._layerwise_sampler.init_layerwise_conf(layer_sample_counts, start_sample_layer, seed)

    def layerwise_sample(self, user_input, index_input, with_hierarchy: bool = False) -> list:
        if self._layerwise_sampler is None:
            raise ValueError("please init layerwise_sampler first.")
        return self._layerwise_sampler.sample(user_input, index_input, with_hierarchy)

    def to_tdm_tree(self):
        """The tensor form the ``tdm_child`` / ``tdm_sampler`` ops take (ops/tdm.py), built from the leaves in code order."""
        from .ops.tdm import TDMTree
        leafs = sorted(zip(self._tree._leaf_codes.tolist(), self._tree._leaf_ids.tolist()))
        return TDMTree([i for _, i in leafs], branch=self._branch)
