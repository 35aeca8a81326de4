"""Record-based legacy feeds: ``MultiSlotDataFeed`` (streaming), ``MultiSlotInMemoryDataFeed`` +
``MultiSlotDataset`` (in-memory, global shuffle, merge by instance id) and ``PaddleBoxDataFeed`` (Record +
page-view merge).

The production path of this framework is the native SlotRecord pipeline (csrc/host/dataset.cc).  The reference
also keeps its older feeds whose unit is the ``Record`` — two flat lists of ``FeatureItem`` (sign, slot) instead
of per-slot CSR (framework/data_feed.h:150-193) — with its own wire format for the cross-trainer shuffle
(framework/data_feed.h:1536-1583) and its own merge rules (framework/data_set.cc:771-877, :1387-1581).  This module
reproduces those: the same text format in, the same bytes on the wire, the same shuffle / merge semantics, and a
bridge to ``SlotBatch`` so that a Record dataset trains through the same sparse ops.

Wire format of one Record (little endian, ``BinaryArchive``: a vector is ``u64 count`` + elements, a string is
``u64 length`` + bytes; a ``FeatureItem`` is ``u64 sign | f32 float view of the sign | u16 slot`` = 14 bytes):

    vector<FeatureItem> uint64_feasigns | vector<FeatureItem> float_feasigns | string ins_id
"""
from __future__ import annotations

import hashlib
import struct
from dataclasses import dataclass, field
from typing import Iterable, Iterator, Sequence

import numpy as np
import torch

_ITEM = np.dtype([("sign", "<u8"), ("fview", "<f4"), ("slot", "<u2")])       # 14 bytes, packed
assert _ITEM.itemsize == 14


@dataclass
class Record:
    """``framework::Record`` (data_feed.h:184-193)."""
    u64_signs: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint64))
    u64_slots: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint16))
    f32_vals: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    f32_slots: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint16))
    ins_id: str = ""
    content: str = ""
    search_id: int = 0
    rank: int = 0
    cmatch: int = 0
    uid: str = ""


# ------------------------------------------------------------------------------------------- wire format
def _items(signs_u64: np.ndarray, slots: np.ndarray, as_float: bool) -> bytes:
    a = np.zeros(signs_u64.size, _ITEM)
    if as_float:
        # a float feasign lives in the low 4 bytes of the 8-byte union; the float view repeats it
        a["sign"] = signs_u64.astype(np.float32).view(np.uint32).astype(np.uint64)
        a["fview"] = signs_u64.astype(np.float32)
    else:
        a["sign"] = signs_u64
        a["fview"] = (signs_u64 & np.uint64(0xFFFFFFFF)).astype(np.uint32).view(np.float32)
    a["slot"] = slots
    return struct.pack("<Q", a.size) + a.tobytes()


def serialize_record(r: Record) -> bytes:
    ins = r.ins_id.encode()
    return (_items(r.u64_signs, r.u64_slots, False) + _items(r.f32_vals, r.f32_slots, True)
            + struct.pack("<Q", len(ins)) + ins)


def serialize_records(records: Iterable[Record]) -> bytes:
    return b"".join(serialize_record(r) for r in records)


def deserialize_records(buf: bytes) -> list:
    out, pos, n = [], 0, len(buf)
    mv = memoryview(buf)
    while pos < n:
        parts = []
        for _ in range(2):
            (cnt,) = struct.unpack_from("<Q", mv, pos)
            pos += 8
            if pos + cnt * 14 > n:
                raise ValueError("truncated Record archive")
            parts.append(np.frombuffer(mv, _ITEM, cnt, pos).copy())
            pos += cnt * 14
        (ln,) = struct.unpack_from("<Q", mv, pos)
        pos += 8
        if pos + ln > n:
            raise ValueError("truncated Record archive")
        ins = bytes(mv[pos:pos + ln]).decode()
        pos += ln
        u, f = parts
        out.append(Record(u["sign"].copy(), u["slot"].copy(), f["fview"].copy(), f["slot"].copy(), ins))
    return out


# ------------------------------------------------------------------------------------------- text parsing
def parse_multislot_line(line: str, slot_types: Sequence[str], used: Sequence[bool] | None = None,
                         parse_ins_id: bool = False, parse_content: bool = False) -> Record | None:
    """One line of the MultiSlot text format, ``[1 ins_id] [1 content] (n v_1 .. v_n) per slot``
    (MultiSlotInMemoryDataFeed::ParseOneInstanceFromPipe, data_feed.cc:1040-1168): ``slot_types[i]`` is
    ``"uint64"`` or ``"float"``; unused slots are skipped; zero feasigns of sparse slots are dropped."""
    tok = line.split()
    if not tok:
        return None
    p = 0
    r = Record()
    if parse_ins_id:
        if tok[p] != "1":
            raise ValueError("ins_id group must start with 1")
        r.ins_id = tok[p + 1]
        p += 2
    if parse_content:
        if tok[p] != "1":
            raise ValueError("content group must start with 1")
        r.content = tok[p + 1]
        p += 2
    us, uslot, fs, fslot = [], [], [], []
    for i, ty in enumerate(slot_types):
        if p >= len(tok):
            raise ValueError(f"line ends before slot {i}")
        n = int(tok[p])
        if n <= 0:
            raise ValueError("the number of ids can not be zero, you need padding it in data generator")
        vals = tok[p + 1:p + 1 + n]
        if len(vals) != n:
            raise ValueError(f"slot {i} announces {n} values, {len(vals)} present")
        p += 1 + n
        if used is not None and not used[i]:
            continue
        if ty == "float":
            for v in vals:
                fv = float(v)
                if abs(fv) < 1e-6:
                    continue
                fs.append(fv)
                fslot.append(i)
        else:
            for v in vals:
                iv = int(v)
                if iv == 0:
                    continue
                us.append(iv)
                uslot.append(i)
    r.u64_signs = np.array(us, np.uint64)
    r.u64_slots = np.array(uslot, np.uint16)
    r.f32_vals = np.array(fs, np.float32)
    r.f32_slots = np.array(fslot, np.uint16)
    return r


# ------------------------------------------------------------------------------------------- batch bridge
def records_to_slot_batch(records: Sequence[Record], sparse_slots: Sequence[int], dense_slots: Sequence[int],
                          dense_dims: Sequence[int] | None = None, device=None):
    """Records -> ``SlotBatch`` (slot-major keys + lod, dense [B, sum(dims)]): the layout every sparse op of this
    framework consumes.  ``sparse_slots`` / ``dense_slots``: slot indices in output order."""
    from .ops.pull_box_sparse import SlotBatch
    B, S = len(records), len(sparse_slots)
    pos = {s: i for i, s in enumerate(sparse_slots)}
    per = [[[] for _ in range(B)] for _ in range(S)]
    for b, r in enumerate(records):
        for sg, sl in zip(r.u64_signs.tolist(), r.u64_slots.tolist()):
            i = pos.get(sl)
            if i is not None:
                per[i][b].append(sg)
    lens = np.array([len(per[i][b]) for i in range(S) for b in range(B)], np.int64)
    lod = np.zeros(S * B + 1, np.int32)
    lod[1:] = np.cumsum(lens)
    keys = np.array([k for i in range(S) for b in range(B) for k in per[i][b]], np.uint64).view(np.int64)
    dims = list(dense_dims) if dense_dims is not None else [1] * len(dense_slots)
    dense = np.zeros((B, sum(dims)), np.float32)
    col0 = {s: int(sum(dims[:i])) for i, s in enumerate(dense_slots)}
    fill = {}
    for b, r in enumerate(records):
        fill.clear()
        for v, sl in zip(r.f32_vals.tolist(), r.f32_slots.tolist()):
            if sl in col0:
                j = fill.get(sl, 0)
                if j < dims[dense_slots.index(sl)]:
                    dense[b, col0[sl] + j] = v
                    fill[sl] = j + 1
    t = lambda a: torch.from_numpy(a).to(device) if device is not None else torch.from_numpy(a)   # noqa: E731
    return SlotBatch(t(keys), t(lod), B, S, t(dense))


# ------------------------------------------------------------------------------------------- feeds / dataset
class MultiSlotDataFeed:
    """Streaming queue feed (data_feed.h:1585-1640): files are read in order, nothing is kept in memory, every
    ``batch_size`` parsed lines become one batch."""

    def __init__(self, slot_types: Sequence[str], batch_size: int = 1, used: Sequence[bool] | None = None,
                 pipe_command: str = "cat", parse_ins_id: bool = False, parse_content: bool = False):
        self.slot_types, self.batch_size, self.used, self.pipe_command = list(slot_types), batch_size, used, pipe_command
        self.parse_ins_id, self.parse_content = parse_ins_id, parse_content

    def _lines(self, path: str) -> Iterator[str]:
        if self.pipe_command and self.pipe_command != "cat":
            import subprocess
            with subprocess.Popen(f"cat {path} | {self.pipe_command}", shell=True, stdout=subprocess.PIPE, text=True) as p:
                yield from p.stdout
        else:
            with open(path) as f:
                yield from f

    def records(self, files: Sequence[str], **parse_kw) -> Iterator[Record]:
        parse_kw.setdefault("parse_ins_id", self.parse_ins_id)
        parse_kw.setdefault("parse_content", self.parse_content)
        for path in files:
            for line in self._lines(path):
                r = parse_multislot_line(line, self.slot_types, self.used, **parse_kw)
                if r is not None:
                    yield r

    def batches(self, files: Sequence[str]) -> Iterator[list]:
        cur = []
        for r in self.records(files):
            cur.append(r)
            if len(cur) == self.batch_size:
                yield cur
                cur = []
        if cur:
            yield cur


class MultiSlotDataset:
    """``MultiSlotDataset`` + ``MultiSlotInMemoryDataFeed`` (data_set.cc:50-1903): Records in memory, local and
    global shuffle, merge by instance id, PV grouping (``PaddleBoxDataFeed``)."""

    def __init__(self, slot_types: Sequence[str], batch_size: int = 1, used: Sequence[bool] | None = None):
        self.feed = MultiSlotDataFeed(slot_types, batch_size, used)
        self.batch_size = batch_size
        self.filelist: list = []
        self.records: list = []
        self.parse_ins_id = self.parse_content = False
        self.merge_size = 0
        self.pvs: list | None = None
        self.rank, self.world, self.group = 0, 1, None

    # -- configuration (names of the reference) --
    def set_filelist(self, files): self.filelist = list(files)
    def set_batch_size(self, n): self.batch_size = self.feed.batch_size = n
    def set_parse_ins_id(self, v): self.parse_ins_id = bool(v)
    def set_parse_content(self, v): self.parse_content = bool(v)
    def set_pipe_command(self, cmd): self.feed.pipe_command = cmd
    def set_merge_by_lineid(self, merge_size: int = 2): self.merge_size = merge_size; self.parse_ins_id = True

    def set_trainer(self, rank: int, world: int, group=None):
        self.rank, self.world, self.group = rank, world, group

    # -- pass data --
    def load_into_memory(self) -> int:
        mine = self.filelist[self.rank::self.world] if self.world > 1 else self.filelist
        self.records = list(self.feed.records(mine, parse_ins_id=self.parse_ins_id, parse_content=self.parse_content))
        return len(self.records)

    def release_memory(self): self.records, self.pvs = [], None
    def get_memory_data_size(self) -> int: return len(self.records)

    def local_shuffle(self, seed: int = 0) -> None:
        rs = np.random.RandomState(seed)
        self.records = [self.records[i] for i in rs.permutation(len(self.records))]

    def global_shuffle(self, seed: int = 0) -> int:
        """Every record goes to trainer ``hash(ins_id) % world`` (random when it carries no id), serialized in the
        Record wire format and exchanged in one all-to-all (the reference sends the same bytes through the PS
        RPC client, data_set.cc:771-877); then a local shuffle.  Returns the local record count afterwards."""
        if self.world > 1:
            rs = np.random.RandomState(seed + self.rank)
            dest = [[] for _ in range(self.world)]
            for r in self.records:
                if r.ins_id:
                    d = int.from_bytes(hashlib.md5(r.ins_id.encode()).digest()[:8], "little") % self.world
                else:
                    d = int(rs.randint(self.world))
                dest[d].append(r)
            from .parallel.cluster import Cluster
            parts = [np.frombuffer(serialize_records(d), np.uint8) for d in dest]
            got = Cluster(self.group).all_to_all_arrays(parts)
            self.records = [r for g in got for r in deserialize_records(np.asarray(g, np.uint8).tobytes())]
        self.local_shuffle(seed)
        return len(self.records)

    def merge_by_lineid(self) -> int:
        """Records sharing an instance id become one (MultiSlotDataset::MergeByInsId, data_set.cc:1387-1581):
        groups whose size differs from ``merge_size`` (> 0) are dropped; a sparse slot may carry values in at
        most one record of a group (else the group is dropped as conflicting); float slots: first record that
        has the slot wins.  Returns the number of dropped instances."""
        groups: dict = {}
        for r in self.records:
            groups.setdefault(r.ins_id, []).append(r)
        out, dropped = [], 0
        for ins, rs_ in groups.items():
            if self.merge_size > 0 and len(rs_) != self.merge_size:
                dropped += len(rs_)
                continue
            seen_u, conflict = set(), False
            us, usl, fs, fsl, seen_f = [], [], [], [], set()
            for r in rs_:
                slots_here = set(r.u64_slots.tolist())
                if slots_here & seen_u:
                    conflict = True
                    break
                seen_u |= slots_here
                us.append(r.u64_signs); usl.append(r.u64_slots)
                new_f = [i for i, s in enumerate(r.f32_slots.tolist()) if s not in seen_f]
                fs.append(r.f32_vals[new_f]); fsl.append(r.f32_slots[new_f])
                seen_f |= set(r.f32_slots.tolist())
            if conflict:
                dropped += len(rs_)
                continue
            out.append(Record(np.concatenate(us), np.concatenate(usl), np.concatenate(fs), np.concatenate(fsl), ins,
                              rs_[0].content, rs_[0].search_id, rs_[0].rank, rs_[0].cmatch, rs_[0].uid))
        self.records = out
        return dropped

    # -- PaddleBoxDataFeed: page views --
    def preprocess_instance(self) -> int:
        """Group the records of one ``search_id`` into a page view, ads ordered by rank
        (PaddleBoxDataFeed / PvInstance, data_set.cc:1100-1160); returns the number of PVs."""
        by: dict = {}
        for r in self.records:
            by.setdefault(r.search_id, []).append(r)
        self.pvs = [sorted(v, key=lambda r: r.rank) for _, v in sorted(by.items())]
        return len(self.pvs)

    def postprocess_instance(self) -> None:
        if self.pvs is not None:
            self.records = [r for pv in self.pvs for r in pv]
            self.pvs = None

    def pv_batches(self, pv_batch_size: int) -> Iterator[tuple]:
        """(records of the batch, rank_offset int32 [n, 2*max_rank+1]) per batch of whole page views."""
        from .ops.ctr_ops import build_rank_offset
        assert self.pvs is not None, "call preprocess_instance first"
        for i in range(0, len(self.pvs), pv_batch_size):
            recs = [r for pv in self.pvs[i:i + pv_batch_size] for r in pv]
            ro = build_rank_offset(torch.tensor([r.search_id for r in recs]), torch.tensor([r.rank for r in recs]), 3,
                                   torch.tensor([r.cmatch for r in recs]))
            yield recs, ro

    # -- training iteration --
    def batches(self) -> Iterator[list]:
        for i in range(0, len(self.records), self.batch_size):
            yield self.records[i:i + self.batch_size]

    def slot_batches(self, sparse_slots, dense_slots, dense_dims=None, device=None):
        for recs in self.batches():
            yield records_to_slot_batch(recs, sparse_slots, dense_slots, dense_dims, device)
