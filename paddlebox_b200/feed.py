"""Pipelined host -> device feed of the training step.

The reference hides input latency with a reader thread per device worker that packs the next
mini-batch and copies it to the GPU while the current one trains
(``SlotPaddleBoxDataFeed`` + ``MiniBatchGpuPack``: data_feed.cc:3800-4100, the trainer loop
``BoxPSWorker::TrainFiles`` boxps_worker.cc:1190-1260).  Here the same overlap is expressed
with two CUDA streams and no extra thread on the device side:

    copy stream : H2D of batch i+1 (pinned record-major keys / offsets / dense) into one of two
                  device staging sets
    main stream : batch_build (record-major -> slot-major keys + LoD, written straight into the
                  buffers the captured step graph reads), graph replay (or the eager step),
                  D2H of the loss into a pinned ring

The host runs one step ahead of the device: after enqueueing step i it waits for the
"done" event of step i-1 and reads that loss from pinned memory, so kernel-launch latency is
hidden but every step's loss still reaches the host.
"""
from __future__ import annotations

import ctypes as C
import time
from dataclasses import dataclass
from typing import Callable, Iterable, Optional

import torch

from . import _native as N
from .ops.pull_box_sparse import SlotBatch


@dataclass
class HostBatch:
    """One packed mini-batch in pinned host memory, record-major (instance after instance).

    ``keys``     int64 [n]            feasigns, instance-major then slot-major inside an instance
    ``offsets``  int32 [B*(S+1)]      per instance the S+1 key offsets of its slots
    ``dense``    float32 [B, Dd]|None label + dense features
    """
    keys: torch.Tensor
    offsets: torch.Tensor
    dense: Optional[torch.Tensor]
    batch_size: int
    num_slots: int
    n_keys: int
    ins_ids: Optional[list] = None
    release: Optional[Callable[[], None]] = None     # called once the H2D copies have completed
    meta: Optional[dict] = None                      # search_id / rank / cmatch / rank_offset tensors

    def h2d_bytes(self) -> int:
        b = self.n_keys * 8 + self.offsets.numel() * 4
        return b + (self.dense.numel() * 4 if self.dense is not None else 0)


class _DeviceStage:
    def __init__(self, key_cap: int, n_off: int, dense_shape, dev):
        self.keys = torch.empty(key_cap, dtype=torch.int64, device=dev)
        self.offsets = torch.empty(n_off, dtype=torch.int32, device=dev)
        self.dense = torch.empty(dense_shape, dtype=torch.float32, device=dev) if dense_shape else None
        self.built = torch.cuda.Event()      # batch_build that read this stage has been enqueued+run
        self.copied = torch.cuda.Event()


class PipelinedFeed:
    """Feeds :class:`HostBatch` es into a :class:`~paddlebox_b200.trainer.CTRTrainer`."""

    RING = 4
    LAG = 2          # steps the host runs ahead of the loss read-back (< RING)

    def __init__(self, trainer, max_keys: int | None = None):
        self.trainer = trainer
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self.max_keys = int(max_keys or trainer.box.max_keys_per_batch)
        self.stages: list[_DeviceStage] = []
        self.scratch = torch.zeros(1024, dtype=torch.int32, device=self.dev)
        self.loss_pin = torch.zeros(self.RING, dtype=torch.float32).pin_memory()
        self._tgt: SlotBatch | None = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.debug_skip_h2d = False   # diagnosis switches (benchmarks only)
        self.debug_skip_d2h = False
        self.debug_skip_build = False
        self.host_enqueue_s = 0.0     # host time spent enqueueing work
        self.host_wait_s = 0.0        # host time spent waiting for the device (0 => the host is the bottleneck)

    # -- buffers ---------------------------------------------------------------------------
    def _target(self, hb: HostBatch) -> SlotBatch:
        st = getattr(self.trainer, "_static", None) if getattr(self.trainer, "_graph", None) is not None else None
        if st is not None and st.batch_size == hb.batch_size and st.num_slots == hb.num_slots:
            return st
        t = self._tgt
        if t is None or t.batch_size != hb.batch_size or t.num_slots != hb.num_slots:
            B, S = hb.batch_size, hb.num_slots
            keys = torch.zeros(self.max_keys, dtype=torch.int64, device=self.dev)
            lod = torch.zeros(S * B + 1, dtype=torch.int32, device=self.dev)
            dense = torch.zeros(tuple(hb.dense.shape), device=self.dev) if hb.dense is not None else None
            t = self._tgt = SlotBatch(keys, lod, B, S, dense, n_keys=self.max_keys, n_dev=lod[S * B:])
        return t

    def _stage(self, i: int, hb: HostBatch) -> _DeviceStage:
        j = i & 1
        n_off = hb.offsets.numel()
        shape = tuple(hb.dense.shape) if hb.dense is not None else None
        if j >= len(self.stages):
            self.stages.append(_DeviceStage(self.max_keys, n_off, shape, self.dev))
        s = self.stages[j]
        if s.offsets.numel() < n_off or (shape and (s.dense is None or tuple(s.dense.shape) != shape)):
            torch.cuda.current_stream(self.dev).synchronize()
            self.copy_stream.synchronize()
            s = self.stages[j] = _DeviceStage(self.max_keys, n_off, shape, self.dev)
        return s

    # -- pipeline stages -------------------------------------------------------------------
    def _enqueue_copy(self, i: int, hb: HostBatch):
        if hb.n_keys > self.max_keys:
            raise ValueError(f"batch has {hb.n_keys} keys; the feed was sized for {self.max_keys}")
        s = self._stage(i, hb)
        if self.debug_skip_h2d and i >= 2:          # diagnosis only: reuse what the stage already holds
            return s
        cp = self.copy_stream
        cp.wait_event(s.built)                      # the stage's previous reader is done
        with torch.cuda.stream(cp):
            s.keys[:hb.n_keys].copy_(hb.keys[:hb.n_keys], non_blocking=True)
            s.offsets[:hb.offsets.numel()].copy_(hb.offsets, non_blocking=True)
            if hb.dense is not None:
                s.dense.copy_(hb.dense, non_blocking=True)
            s.copied.record(cp)
        self.h2d_bytes += hb.h2d_bytes()
        return s

    def _enqueue_step(self, i: int, hb: HostBatch, s: _DeviceStage):
        tr = self.trainer
        main = torch.cuda.current_stream(self.dev)
        main.wait_event(s.copied)
        tgt = self._target(hb)
        B, S = hb.batch_size, hb.num_slots
        if not (self.debug_skip_build and i >= 2):
            fold = hb.dense is not None and s.dense.is_contiguous() and tgt.dense.is_contiguous()
            N.check(N.cuda().pbx_batch_build_dense(N.ptr(s.keys), N.ptr(s.offsets), B, S, N.ptr(tgt.keys), N.ptr(tgt.lod),
                                                   N.ptr(self.scratch), N.ptr(s.dense) if fold else None,
                                                   N.ptr(tgt.dense) if fold else None,
                                                   C.c_int64(s.dense.numel() if fold else 0), N.stream_ptr(main)),
                    "pbx_batch_build")
            if hb.dense is not None and not fold:
                tgt.dense.copy_(s.dense, non_blocking=True)
        s.built.record(main)
        tgt.ins_ids = hb.ins_ids
        tgt.meta = hb.meta
        if tgt is getattr(tr, "_static", None) and getattr(tr, "_graph", None) is not None and not tr.cfg.is_test:
            tr._graph.replay()
            tr.steps += 1
            loss = tr._graph_loss
        else:
            batch = SlotBatch(tgt.keys[:max(hb.n_keys, 1)], tgt.lod, B, S, tgt.dense, n_keys=hb.n_keys,
                              ins_ids=hb.ins_ids, meta=hb.meta)
            loss = tr._eager_step(batch)
        slot = i % self.RING
        if not self.debug_skip_d2h:
            # the device stores the loss into the pinned ring itself (a copy-engine D2H between two
            # graph launches costs ~10 us of stream latency)
            N.check(N.cuda().pbx_store_host(N.ptr(loss.detach()), C.c_void_p(self.loss_pin.data_ptr() + 4 * slot), 1,
                                            N.stream_ptr(main)), "pbx_store_host")
            self.d2h_bytes += 4
        done = torch.cuda.Event()
        done.record(main)
        return done

    def _retire(self, pending, losses, on_loss):
        i, hb, done = pending
        t0 = time.perf_counter()
        done.synchronize()
        self.host_wait_s += time.perf_counter() - t0
        v = float(self.loss_pin[i % self.RING])
        if hb.release is not None:
            hb.release()
        if losses is not None:
            losses.append(v)
        if on_loss is not None:
            on_loss(i, v)

    def run(self, host_batches: Iterable[HostBatch], on_loss: Callable[[int, float], None] | None = None,
            collect: bool = True, after_enqueue: Callable[[int], None] | None = None):
        """Train on every host batch; returns (steps, instances, [loss per step]).
        ``after_enqueue(i)`` runs on the host right after step i was enqueued (stream-ordered work
        put there executes between step i and step i+1)."""
        it = iter(host_batches)
        losses = [] if collect else None
        first = next(it, None)
        if first is None:
            return 0, 0, losses or []
        i, n_ins, pending = 0, 0, []
        cur, cur_stage = first, self._enqueue_copy(0, first)
        while cur is not None:
            t0 = time.perf_counter()
            nxt = next(it, None)
            nxt_stage = self._enqueue_copy(i + 1, nxt) if nxt is not None else None   # overlaps step i
            done = self._enqueue_step(i, cur, cur_stage)
            self.host_enqueue_s += time.perf_counter() - t0
            n_ins += cur.batch_size
            if after_enqueue is not None:
                after_enqueue(i)
            pending.append((i, cur, done))
            # the host stays LAG steps ahead of the loss it reads: enqueue jitter of one rank does
            # not stall the (collective) step of all ranks
            while len(pending) > self.LAG:
                self._retire(pending.pop(0), losses, on_loss)
            cur, cur_stage = nxt, nxt_stage
            i += 1
        while pending:
            self._retire(pending.pop(0), losses, on_loss)
        return i, n_ins, losses if losses is not None else []
