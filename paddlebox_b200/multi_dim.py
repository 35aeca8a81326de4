"""Per-slot embedding widths: one sparse table per distinct width.

Reference role: the variable-dimension feature layout — ``PullCopyVariable`` /
``PushCopyVariable`` (fleet/box_wrapper.cu:270-322, :714-875) and HeterPS' dynamic-mf mode
(``dy_mf_search_kernel`` / ``dy_mf_update_kernel``, heter_ps/hashtable_kernel.cu:88-140, rows
addressed through a per-key ``mf_dim``) — where slots of one model use different embedx widths.

A row of a different width is a different record size, so here every width gets its own table
(own HBM shard, host tier, optimizer state layout, pull/push kernels specialised by row stride)
instead of one table with a per-key width field: lookups stay one 128-byte-aligned row read and
no row carries padding up to the widest slot.  ``MultiDimBox`` fans the pass lifecycle out to the
tables and splits a batch's slots by width.
"""
from __future__ import annotations

import os
from typing import Sequence

import numpy as np
import torch

from .box_wrapper import BoxWrapper
from .ops.pull_box_sparse import SlotBatch, sparse_embedding_pool
from .ops.seqpool import SeqpoolAttrs


class MultiDimBox:
    def __init__(self, slot_dims: Sequence[int], **box_kw):
        """``slot_dims[s]`` = embedx width of slot ``s``; ``box_kw`` goes to every ``BoxWrapper``."""
        self.slot_dims = [int(d) for d in slot_dims]
        self.groups = {d: [s for s, sd in enumerate(self.slot_dims) if sd == d] for d in sorted(set(self.slot_dims))}
        self.boxes = {d: BoxWrapper(embedx_dim=d, **box_kw) for d in self.groups}
        self._col_index = {}

    # ---- lifecycle, fanned out -----------------------------------------------------------------
    def initialize(self, conf: dict | None = None, **kw) -> None:
        for b in self.boxes.values():
            b.initialize_gpu_and_load_model(dict(conf or {}), **kw)

    def begin_feed_pass(self, date: str | None = None) -> None:
        for b in self.boxes.values():
            b.begin_feed_pass(date)

    def feed_pass(self, keys, slots) -> None:
        """``slots[i]`` = slot index of ``keys[i]``: every key is registered with its width's table."""
        k = np.asarray(keys.cpu() if torch.is_tensor(keys) else keys)
        s = np.asarray(slots.cpu() if torch.is_tensor(slots) else slots)
        dims = np.asarray(self.slot_dims)[s]
        for d, b in self.boxes.items():
            b.feed_pass(k[dims == d])

    def feed_pass_batch(self, batch: SlotBatch) -> None:
        for d, sub in self.split(batch).items():
            self.boxes[d].feed_pass(sub.keys[:sub.N])

    def end_feed_pass(self) -> None:
        for b in self.boxes.values():
            b.end_feed_pass()

    def begin_pass(self) -> None:
        for b in self.boxes.values():
            b.begin_pass()

    def end_pass(self, need_save_delta: bool = False, delta_path: str = "") -> None:
        for d, b in self.boxes.items():
            b.end_pass(need_save_delta, os.path.join(delta_path, f"dim_{d}") if delta_path else "")

    def save_base(self, batch_model_path: str, xbox_model_path: str = "", date: str = "") -> str:
        return " ".join(f"dim_{d}: " + b.save_base(os.path.join(batch_model_path, f"dim_{d}"),
                                                   os.path.join(xbox_model_path, f"dim_{d}") if xbox_model_path else "",
                                                   date) for d, b in self.boxes.items())

    def save_delta(self, xbox_model_path: str) -> str:
        return " ".join(f"dim_{d}: " + b.save_delta(os.path.join(xbox_model_path, f"dim_{d}"))
                        for d, b in self.boxes.items())

    def load_model(self, path: str, serving: bool = False) -> int:
        return sum(b.load_model(os.path.join(path, f"dim_{d}"), serving) for d, b in self.boxes.items())

    def set_test_mode(self, on: bool) -> None:
        for b in self.boxes.values():
            b.set_test_mode(on)

    def finalize(self) -> None:
        for b in self.boxes.values():
            b.finalize()

    # ---- batch split -------------------------------------------------------------------------------
    def split(self, batch: SlotBatch) -> dict:
        """{width: SlotBatch of that width's slots} (slot order kept inside a group)."""
        B, S = batch.batch_size, batch.num_slots
        assert S == len(self.slot_dims), f"batch has {S} slots, MultiDimBox was built for {len(self.slot_dims)}"
        if len(self.groups) == 1:
            return {self.slot_dims[0]: batch}
        lod = batch.lod.long()
        starts, lens = lod[:-1].view(S, B), (lod[1:] - lod[:-1]).view(S, B)
        out = {}
        for d, slots in self.groups.items():
            sl = torch.as_tensor(slots, device=lod.device)
            l = lens[sl].reshape(-1)
            new_lod = torch.zeros(l.numel() + 1, dtype=torch.int64, device=lod.device)
            new_lod[1:] = torch.cumsum(l, 0)
            n = int(new_lod[-1])                                   # the one host sync of a split
            src0 = starts[sl].reshape(-1) - new_lod[:-1]           # per bag: source offset - destination offset
            idx = torch.repeat_interleave(src0, l, output_size=n) + torch.arange(n, device=lod.device)
            out[d] = SlotBatch(batch.keys[idx], new_lod.to(torch.int32), B, len(slots), None, n_keys=n)
        return out

    # ---- forward ----------------------------------------------------------------------------------
    def out_widths(self, attrs: SeqpoolAttrs) -> list:
        return [attrs.out_width(3 + d) for d in self.slot_dims]

    def embedding_pool(self, batch: SlotBatch, cvm: torch.Tensor, attrs: SeqpoolAttrs | None = None) -> torch.Tensor:
        """pull + fused_seqpool_cvm of every slot at its own width -> [B, sum_s Eo_s], columns in the
        batch's slot order (the reference's per-slot output list, concatenated)."""
        attrs = attrs or SeqpoolAttrs()
        subs = self.split(batch)
        parts = {d: sparse_embedding_pool(subs[d], cvm, self.boxes[d], attrs) for d in self.groups}
        if len(parts) == 1:
            return next(iter(parts.values()))
        key = (id(attrs), str(cvm.device))
        col = self._col_index.get(key)
        if col is None:
            # column c of the result = column col[c] of cat(parts in width order)
            base, where = 0, {}
            for d, slots in self.groups.items():
                w = attrs.out_width(3 + d)
                for j, s in enumerate(slots):
                    where[s] = (base + j * w, w)
                base += len(slots) * w
            col = torch.cat([torch.arange(b, b + w) for b, w in (where[s] for s in range(len(self.slot_dims)))])
            col = self._col_index[key] = col.to(cvm.device)
        return torch.cat([parts[d] for d in self.groups], 1).index_select(1, col)
