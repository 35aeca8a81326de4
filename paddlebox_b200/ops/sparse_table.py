"""HBM shard of the sparse parameter table (one per GPU).

Python face of ``csrc/cuda/table.cu``.  Role in the reference: ``HashTable`` +
``HBMMemoryPool`` + the optimizer functors (heter_ps/hashtable.h:1-219,
hashtable_kernel.cu:34-337, optimizer.cuh.h, mem_pool.h) and, for BoxPS proper, the
HBM tier of the closed ``libbox_ps.so``.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N

OPTIMIZERS = {"adagrad": 0, "adam": 1, "shared_adam": 2}


class SparseTable:
    """Open-addressing key -> value-row table in device memory.

    capacity      number of slots (keep the load factor <= 0.75 like the reference,
                  heter_comm.h:338)
    embedx_dim    width of the embedx vector (reference default 8, box_wrapper.h:652)
    optimizer     "adagrad" | "adam" | "shared_adam"
    """

    def __init__(self, capacity: int, embedx_dim: int = 8, optimizer: str = "adagrad",
                 config: N.SparseOptConfig | None = None, seed: int = 0, device: int | None = None):
        if not torch.cuda.is_available():
            raise RuntimeError("SparseTable needs a CUDA device (use paddlebox_b200.host.HostTable on CPU)")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.capacity = int(capacity)
        self.embedx_dim = int(embedx_dim)
        self.optimizer = optimizer
        self.config = config or N.SparseOptConfig()
        self.layout = N.make_layout(self.embedx_dim, OPTIMIZERS[optimizer])
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            N.check(N.cuda().pbx_table_create(C.byref(self._h), C.c_uint64(self.capacity), self.embedx_dim,
                                              OPTIMIZERS[optimizer], C.byref(self.config), C.c_uint64(seed)),
                    "pbx_table_create")

    # -- geometry ---------------------------------------------------------
    @property
    def hidden(self) -> int:
        return 3 + self.embedx_dim

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def __del__(self):
        try:
            if self._h:
                N.cuda().pbx_table_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # -- configuration ------------------------------------------------------
    def set_config(self, config: N.SparseOptConfig) -> None:
        self.config = config
        N.check(N.cuda().pbx_table_set_config(self._h, C.byref(config)))

    def set_test_mode(self, on: bool) -> None:
        """Test mode: pushes do not update the table (reference ``SetTestMode``)."""
        N.check(N.cuda().pbx_table_set_test_mode(self._h, int(bool(on))))

    # -- pass lifecycle -----------------------------------------------------
    def clear(self) -> None:
        N.check(N.cuda().pbx_table_clear(self._h, N.stream_ptr()))

    def insert(self, keys: torch.Tensor, rows: torch.Tensor | None = None) -> None:
        """Bulk insert (BeginPass staging).  ``keys`` int64/uint64 on device; ``rows``
        optional [n, layout.stride] fp32 full value rows (from a checkpoint or the host
        tier); new keys without rows get the optimizer's initial state."""
        keys = keys.contiguous()
        assert keys.is_cuda and keys.dtype in (torch.int64, torch.uint64)
        if rows is not None:
            rows = rows.contiguous()
            assert rows.shape == (keys.numel(), self.layout.stride) and rows.dtype == torch.float32
        N.check(N.cuda().pbx_table_insert(self._h, N.ptr(keys), C.c_int64(keys.numel()), N.ptr(rows),
                                          N.stream_ptr()), "pbx_table_insert")

    def fill_synthetic(self, n_total: int, key_seed: int = 1, rank: int = 0, nranks: int = 1,
                       create_mf: bool = True) -> None:
        """Populate with the keys ``synthetic_key(i)`` for i < n_total that hash to this
        rank (benchmarks: billions of keys without a host pass)."""
        N.check(N.cuda().pbx_table_fill_synthetic(self._h, C.c_uint64(n_total), C.c_uint64(key_seed), rank, nranks,
                                                  int(create_mf), N.stream_ptr()), "pbx_table_fill_synthetic")

    def size(self) -> int:
        out = C.c_int64()
        N.check(N.cuda().pbx_table_count(self._h, C.byref(out), N.stream_ptr()))
        return out.value

    def dump(self, only_dirty: bool = False, clear_dirty: bool = False):
        """(keys int64 [n], rows fp32 [n, stride]) of every live (or dirty) entry."""
        n = C.c_int64()
        N.check(N.cuda().pbx_table_dump(self._h, None, None, C.c_int64(0), int(only_dirty), 0, C.byref(n),
                                        N.stream_ptr()))
        cnt = n.value
        keys = torch.empty(max(cnt, 1), dtype=torch.int64, device=self.device)
        rows = torch.empty(max(cnt, 1), self.layout.stride, dtype=torch.float32, device=self.device)
        N.check(N.cuda().pbx_table_dump(self._h, N.ptr(keys), N.ptr(rows), C.c_int64(cnt), int(only_dirty),
                                        int(clear_dirty), C.byref(n), N.stream_ptr()))
        return keys[:cnt], rows[:cnt]

    def lookup(self, keys: torch.Tensor) -> torch.Tensor:
        """Local pull records [n, hidden] (no dedup, no cross-GPU routing)."""
        keys = keys.contiguous()
        out = torch.empty(keys.numel(), self.hidden, dtype=torch.float32, device=self.device)
        N.check(N.cuda().pbx_table_lookup(self._h, N.ptr(keys), C.c_int64(keys.numel()), N.ptr(out), self.hidden,
                                          None, N.stream_ptr()))
        return out

    def fetch_rows(self, keys: torch.Tensor, out: torch.Tensor | None = None, with_found: bool = False):
        """Full value rows [n, layout.stride] of ``keys`` (zero rows for absent keys)."""
        keys = keys.contiguous()
        n = keys.numel()
        if out is None:
            out = torch.empty(n, self.layout.stride, dtype=torch.float32, device=self.device)
        found = torch.empty(max(n, 1), dtype=torch.uint8, device=self.device) if with_found else None
        N.check(N.cuda().pbx_table_fetch_rows(self._h, N.ptr(keys), C.c_int64(n), N.ptr(out), N.ptr(found),
                                              N.stream_ptr()), "pbx_table_fetch_rows")
        return (out, found[:n].bool()) if with_found else out

    def update(self, keys: torch.Tensor, push: torch.Tensor) -> None:
        """Apply merged push records [n, >= 4 + D] for *unique* keys."""
        keys = keys.contiguous()
        push = push.contiguous()
        N.check(N.cuda().pbx_table_update(self._h, N.ptr(keys), N.ptr(push), push.shape[1],
                                          C.c_int64(keys.numel()), N.stream_ptr()))

    def shrink(self, decay: float = 0.98, delete_threshold: float = 0.0,
               delete_after_unseen_days: float = 30.0) -> int:
        """Decay show/click, age entries, evict stale ones; returns survivors
        (reference ``ShrinkTable``; rule from ctr_dymf_accessor.cc:68-85)."""
        kept = C.c_int64()
        N.check(N.cuda().pbx_table_shrink(self._h, C.c_float(decay), C.c_float(delete_threshold),
                                          C.c_float(delete_after_unseen_days), C.byref(kept), N.stream_ptr()))
        return kept.value


def synthetic_key(i: torch.Tensor, key_seed: int = 1) -> torch.Tensor:
    """Key generator matching ``k_fill_synth``: mix64(i * golden + seed) >> 1 (int64 math
    with wraparound)."""
    def mix64(k):
        k = k ^ ((k >> 33) & 0x7FFFFFFF)
        k = k * torch.tensor(-49064778989728563, dtype=torch.int64, device=k.device)  # 0xff51afd7ed558ccd
        k = k ^ ((k >> 33) & 0x7FFFFFFF)
        k = k * torch.tensor(-4265267296055464877, dtype=torch.int64, device=k.device)  # 0xc4ceb9fe1a85ec53
        k = k ^ ((k >> 33) & 0x7FFFFFFF)
        return k
    i = i.to(torch.int64)
    x = i * torch.tensor(-7046029254386353131, dtype=torch.int64, device=i.device) + key_seed  # 0x9E3779B97F4A7C15
    return (mix64(x) >> 1) & 0x7FFFFFFFFFFFFFFF
