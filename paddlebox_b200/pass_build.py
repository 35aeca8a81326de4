"""Pass build / write-back pipeline: host tier <-> HBM shard, streamed.

What the reference does with three channels and a host pool (``PSGPUWrapper``:
``data_ready -> buildcpu_ready -> gpu_free``, fleet/ps_gpu_wrapper.cc:885-955; ``BuildPull`` :337-682;
``BuildGPUTask`` :684-883; the pass-scoped ``MemoryPool`` / ``HBMMemoryPool``, heter_ps/mem_pool.h:28-102;
chunked ``build_ps`` on alternating streams, heter_ps/heter_comm_inl.h:671-731; ``EndPass`` dumping HBM back
to the CPU table, ps_gpu_wrapper.cc:983-1079) is one small pipeline here:

    fetch workers (CPU threads: host table -> pinned staging slot)
        --> ready queue --> device stage (H2D of slot k on a copy stream, insert kernel on the compute
                            stream once the copy's event fired) --> free queue --> fetch workers

* ``StagingPool`` is the MemoryPool: ``depth`` slots of pinned host memory (keys + value rows of one chunk)
  with a device twin each, allocated once and reused by every pass — no per-chunk allocation, no pageable
  copy.  While the device copies / inserts slot k, the fetch workers fill slot k+1 .. k+depth-1.
* ``PassBuilder.build`` streams a key list through it (BeginPass); ``PassBuilder.write_back`` streams the
  touched rows the other way in bounded chunks (EndPass).
* Both report rows/s and bytes/s (``last_build`` / ``last_write_back``) so that a pass at bench scale can be
  compared with the PCIe rate (``bench.py --real-pass``).

Without a GPU (tests, CPU plumbing config) the device stage is a callable, so the pipeline logic — chunking,
slot recycling, ordering, completeness, error propagation — runs anywhere.
"""
from __future__ import annotations

import queue
import threading
import time
from typing import Callable

import numpy as np
import torch


class StagingSlot:
    __slots__ = ("keys", "rows", "d_keys", "d_rows", "n", "event", "index")

    def __init__(self, chunk: int, stride: int, device: torch.device | None, index: int):
        pin = device is not None and device.type == "cuda"
        self.keys = torch.empty(chunk, dtype=torch.int64, pin_memory=pin)
        self.rows = torch.empty(chunk, stride, dtype=torch.float32, pin_memory=pin)
        self.d_keys = torch.empty(chunk, dtype=torch.int64, device=device) if pin else None
        self.d_rows = torch.empty(chunk, stride, dtype=torch.float32, device=device) if pin else None
        self.n = 0
        self.event = torch.cuda.Event() if pin else None
        self.index = index


class StagingPool:
    """``depth`` pinned staging slots of ``chunk`` rows (+ device twins); reused across passes
    (reference MemoryPool / HBMMemoryPool, heter_ps/mem_pool.h:28-102)."""

    def __init__(self, chunk: int, stride: int, device: torch.device | None = None, depth: int = 3):
        assert chunk > 0 and depth >= 2
        self.chunk, self.stride, self.device, self.depth = int(chunk), int(stride), device, depth
        self.slots = [StagingSlot(self.chunk, self.stride, device, i) for i in range(depth)]

    def bytes_per_row(self) -> int:
        return 8 + 4 * self.stride

    def host_bytes(self) -> int:
        return self.depth * self.chunk * self.bytes_per_row()


class PassBuilder:
    """Streams rows between a host table and a device table through a ``StagingPool``.

    ``host_table``  : object with ``fetch_rows_into(keys_u64, out_rows, create)`` or ``fetch_rows`` and
                      ``store_rows(keys, rows, mark_dirty)`` (``paddlebox_b200.host.HostTable``)
    ``insert``      : ``f(d_keys int64 [n], d_rows fp32 [n, stride])`` — ``SparseTable.insert``; on a CPU-only
                      box any callable (it then receives the pinned-less host views)
    """

    def __init__(self, host_table, insert: Callable, stride: int, device: torch.device | None = None,
                 chunk: int = 1 << 20, depth: int = 3, fetch_threads: int = 2):
        self.host_table, self.insert = host_table, insert
        self.device = device if (device is not None and device.type == "cuda") else None
        self.pool = StagingPool(chunk, stride, self.device, depth)
        self.fetch_threads = max(1, fetch_threads)
        self.copy_stream = torch.cuda.Stream(device=self.device) if self.device is not None else None
        self.last_build: dict = {}
        self.last_write_back: dict = {}

    # ------------------------------------------------------------------ BeginPass
    def _fetch_into(self, keys: np.ndarray, slot: StagingSlot) -> None:
        n = keys.size
        slot.n = n
        slot.keys[:n].numpy()[:] = keys.view(np.int64)
        out = slot.rows[:n].numpy()
        fn = getattr(self.host_table, "fetch_rows_into", None)
        if fn is not None:
            fn(keys, out, True)
        else:
            out[:] = self.host_table.fetch_rows(keys, create=True)

    def build(self, keys: np.ndarray) -> dict:
        """Insert the rows of ``keys`` (uint64, unique) into the device table; returns timing / rate."""
        t0 = time.time()
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        chunk = self.pool.chunk
        n_chunks = (keys.size + chunk - 1) // chunk
        free_q: "queue.Queue[StagingSlot]" = queue.Queue()
        ready_q: "queue.Queue" = queue.Queue()
        for s in self.pool.slots:
            free_q.put(s)
        next_chunk = [0]
        lock = threading.Lock()
        errors: list = []
        fetch_s = [0.0]

        def fetch_worker():
            try:
                while True:
                    with lock:
                        c = next_chunk[0]
                        if c >= n_chunks:
                            return
                        next_chunk[0] = c + 1
                    slot = free_q.get()
                    if slot is None:            # shutdown after an error elsewhere
                        free_q.put(None)
                        return
                    t1 = time.time()
                    self._fetch_into(keys[c * chunk:(c + 1) * chunk], slot)
                    with lock:
                        fetch_s[0] += time.time() - t1
                    ready_q.put((c, slot))
            except Exception as e:  # noqa: BLE001
                errors.append(e)
                ready_q.put((-1, None))

        workers = [threading.Thread(target=fetch_worker, daemon=True) for _ in range(min(self.fetch_threads, max(n_chunks, 1)))]
        for w in workers:
            w.start()
        done, in_flight = 0, []
        compute = torch.cuda.current_stream(self.device) if self.device is not None else None
        try:
            while done < n_chunks:
                c, slot = ready_q.get()
                if slot is None:
                    raise errors[0]
                n = slot.n
                if self.device is not None:
                    with torch.cuda.stream(self.copy_stream):
                        slot.d_keys[:n].copy_(slot.keys[:n], non_blocking=True)
                        slot.d_rows[:n].copy_(slot.rows[:n], non_blocking=True)
                        slot.event.record(self.copy_stream)
                    compute.wait_event(slot.event)
                    self.insert(slot.d_keys[:n], slot.d_rows[:n])      # synchronises on its status word
                else:
                    self.insert(slot.keys[:n], slot.rows[:n])
                free_q.put(slot)
                done += 1
        finally:
            free_q.put(None)
            for w in workers:
                w.join(timeout=30)
        dt = max(time.time() - t0, 1e-9)
        nbytes = keys.size * self.pool.bytes_per_row()
        self.last_build = {"rows": int(keys.size), "chunks": int(n_chunks), "seconds": dt, "rows_per_s": keys.size / dt,
                           "GB_per_s": nbytes / dt / 1e9, "host_fetch_s": fetch_s[0],
                           "staging_host_bytes": self.pool.host_bytes()}
        return self.last_build

    # ------------------------------------------------------------------ EndPass
    def write_back(self, d_keys: torch.Tensor, d_rows: torch.Tensor, mark_dirty: bool = True) -> dict:
        """Store device rows (the touched rows of the pass) into the host table in bounded chunks:
        D2H of chunk k+1 (copy stream, pinned slot) overlaps the host-table store of chunk k."""
        t0 = time.time()
        n_total = int(d_keys.numel())
        chunk = self.pool.chunk
        slots = self.pool.slots
        pending = []                                   # (slot, n) whose copy has been enqueued
        store_s = 0.0

        def drain(item):
            nonlocal store_s
            slot, n = item
            if slot.event is not None:
                slot.event.synchronize()
            t1 = time.time()
            self.host_table.store_rows(slot.keys[:n].numpy().view(np.uint64), slot.rows[:n].numpy(), mark_dirty=mark_dirty)
            store_s += time.time() - t1

        if self.device is not None and n_total:
            self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
        for i, b in enumerate(range(0, n_total, chunk)):
            slot = slots[i % len(slots)]
            if len(pending) == len(slots):             # the slot we are about to reuse must have been stored
                drain(pending.pop(0))
            n = min(chunk, n_total - b)
            if self.device is not None:
                with torch.cuda.stream(self.copy_stream):
                    slot.keys[:n].copy_(d_keys[b:b + n], non_blocking=True)
                    slot.rows[:n].copy_(d_rows[b:b + n], non_blocking=True)
                    slot.event.record(self.copy_stream)
            else:
                slot.keys[:n].copy_(d_keys[b:b + n])
                slot.rows[:n].copy_(d_rows[b:b + n])
            pending.append((slot, n))
        while pending:
            drain(pending.pop(0))
        dt = max(time.time() - t0, 1e-9)
        self.last_write_back = {"rows": n_total, "seconds": dt, "rows_per_s": n_total / dt,
                                "GB_per_s": n_total * self.pool.bytes_per_row() / dt / 1e9, "host_store_s": store_s}
        return self.last_write_back
