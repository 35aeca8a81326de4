"""Python faces of the C++ host runtime (libpbx_host.so): host-memory table tier, AUC
calculators, the native dataset, the async dense table.  Everything releases the GIL while
in native code (ctypes does), like the reference's pybind methods (box_helper_py.cc)."""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import numpy as np
import torch

from . import _native as N

OPTIMIZERS = {"adagrad": 0, "adam": 1, "shared_adam": 2}


def _u64p(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_uint64))


def _f32p(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _as_np_u64(keys) -> np.ndarray:
    if isinstance(keys, torch.Tensor):
        keys = keys.detach().cpu().numpy()
    return np.ascontiguousarray(keys).view(np.uint64) if keys.dtype != np.uint64 else np.ascontiguousarray(keys)


class HostTable:
    """Host-memory tier of the sparse table; also the CPU execution path of pull/push
    (csrc/host/host_table.cc)."""

    def __init__(self, embedx_dim: int = 8, optimizer: str = "adagrad", config: N.SparseOptConfig | None = None,
                 seed: int = 0):
        self.embedx_dim = embedx_dim
        self.optimizer = optimizer
        self.config = config or N.SparseOptConfig()
        self.layout = N.make_layout(embedx_dim, OPTIMIZERS[optimizer])
        self._h = C.c_void_p()
        N.hcheck(N.host().pbh_table_create(C.byref(self._h), embedx_dim, OPTIMIZERS[optimizer],
                                           C.byref(self.config), C.c_uint64(seed)))
        N.host().pbh_table_size.restype = C.c_int64
        N.host().pbh_table_shrink.restype = C.c_int64
        N.host().pbh_table_keys.restype = C.c_int64

    def __del__(self):
        try:
            if self._h:
                N.host().pbh_table_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    @property
    def hidden(self) -> int:
        return 3 + self.embedx_dim

    def size(self) -> int:
        return N.host().pbh_table_size(self._h)

    def set_config(self, cfg: N.SparseOptConfig) -> None:
        self.config = cfg
        N.hcheck(N.host().pbh_table_set_config(self._h, C.byref(cfg)))

    def set_test_mode(self, on: bool) -> None:
        N.hcheck(N.host().pbh_table_set_test_mode(self._h, int(on)))

    def set_serving_quant(self, scale: float) -> None:
        """``scale`` > 0: serving-format saves hold embedx as int16 multiples of ``scale``."""
        N.hcheck(N.host().pbh_table_set_serving_quant(self._h, C.c_float(scale)))

    def clear(self) -> None:
        N.hcheck(N.host().pbh_table_clear(self._h))

    def fetch_rows(self, keys, create: bool = True) -> np.ndarray:
        k = _as_np_u64(keys)
        out = np.empty((k.size, self.layout.stride), dtype=np.float32)
        N.hcheck(N.host().pbh_table_fetch_rows(self._h, _u64p(k), C.c_int64(k.size), _f32p(out), int(create)))
        return out

    def fetch_rows_into(self, keys, out: np.ndarray, create: bool = True) -> None:
        """``fetch_rows`` straight into a caller-owned buffer (a pinned staging slot of the pass builder):
        no intermediate array, and the native call runs without the GIL, so several threads fetch at once."""
        k = _as_np_u64(keys)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.shape == (k.size, self.layout.stride)
        N.hcheck(N.host().pbh_table_fetch_rows(self._h, _u64p(k), C.c_int64(k.size), _f32p(out), int(create)))

    def store_rows(self, keys, rows, mark_dirty: bool = True) -> None:
        k = _as_np_u64(keys)
        r = np.ascontiguousarray(rows.detach().cpu().numpy() if isinstance(rows, torch.Tensor) else rows,
                                 dtype=np.float32)
        assert r.shape == (k.size, self.layout.stride)
        N.hcheck(N.host().pbh_table_store_rows(self._h, _u64p(k), C.c_int64(k.size), _f32p(r), int(mark_dirty)))

    def pull(self, keys) -> torch.Tensor:
        k = _as_np_u64(keys)
        out = np.empty((k.size, self.hidden), dtype=np.float32)
        N.hcheck(N.host().pbh_table_pull(self._h, _u64p(k), C.c_int64(k.size), _f32p(out), self.hidden))
        return torch.from_numpy(out)

    def push(self, keys, push_records) -> None:
        k = _as_np_u64(keys)
        p = np.ascontiguousarray(push_records.detach().cpu().numpy() if isinstance(push_records, torch.Tensor)
                                 else push_records, dtype=np.float32)
        N.hcheck(N.host().pbh_table_push(self._h, _u64p(k), C.c_int64(k.size), _f32p(p), p.shape[1]))

    def shrink(self, decay=0.98, delete_threshold=0.0, delete_after_unseen_days=30.0) -> int:
        return N.host().pbh_table_shrink(self._h, C.c_float(decay), C.c_float(delete_threshold),
                                         C.c_float(delete_after_unseen_days))

    def save(self, directory: str, mode: int = 0, shard_base: int = 0, base_threshold: float = 0.0,
             delta_threshold: float = 0.0, clear_dirty: "bool | int" = True) -> int:
        """``clear_dirty``: True / 1 = a delta save (mode 1) clears the dirty flags of the rows it wrote;
        2 = a full save (mode 0) clears them too (the delta baseline is reset, as after ``save_base``)."""
        os.makedirs(directory, exist_ok=True)
        n = C.c_int64()
        N.hcheck(N.host().pbh_table_save(self._h, directory.encode(), mode, shard_base, C.c_float(base_threshold),
                                         C.c_float(delta_threshold), int(clear_dirty), C.byref(n)))
        return n.value

    def load(self, directory: str, serving: bool = False, rank: int = 0, world: int = 1) -> int:
        """Load every ``part-*`` file of ``directory``.  With ``world`` > 1 only the keys that rank ``rank``
        owns are kept, whatever number of ranks wrote the files (resharding on load)."""
        files = [os.path.join(directory, n) for n in sorted(os.listdir(directory)) if n.startswith("part-")]

        def load_one(path: str) -> int:
            n = C.c_int64()
            N.hcheck(N.host().pbh_table_load_file_owned(self._h, path.encode(), int(serving), rank, world, C.byref(n)))
            return n.value

        if len(files) <= 1:
            return sum(load_one(f) for f in files)
        # the native call parses without the GIL and fills rows under shard locks: files load in parallel
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(16, len(files), (os.cpu_count() or 4))) as ex:
            return sum(ex.map(load_one, files))

    # ---- SSD tier ----------------------------------------------------------------------------
    def ssd_open(self, directory: str) -> None:
        """Enable the SSD tier: one append-only file per shard under ``directory``."""
        os.makedirs(directory, exist_ok=True)
        N.hcheck(N.host().pbh_table_ssd_open(self._h, directory.encode()))

    def ssd_close(self, unlink: bool = False) -> None:
        N.hcheck(N.host().pbh_table_ssd_close(self._h, int(unlink)))

    def _i64(self, name: str, *args) -> int:
        fn = getattr(N.host(), name)
        fn.restype = C.c_int64
        return fn(self._h, *args)

    def spill(self, min_unseen_days: float = -1.0, max_mem_rows: int = 0) -> int:
        """Move cold rows to the SSD tier; returns how many went (see include/pbx_host.h)."""
        return self._i64("pbh_table_spill", C.c_float(min_unseen_days), C.c_int64(max_mem_rows))

    def load_ssd2mem(self) -> int:
        return self._i64("pbh_table_load_ssd2mem")

    def mem_rows(self) -> int:
        return self._i64("pbh_table_mem_rows")

    def ssd_rows(self) -> int:
        return self._i64("pbh_table_ssd_rows")

    def keys(self) -> np.ndarray:
        n = N.host().pbh_table_keys(self._h, None, C.c_int64(0))
        out = np.empty(n, dtype=np.uint64)
        if n:
            N.host().pbh_table_keys(self._h, _u64p(out), C.c_int64(n))
        return out


class HostAuc:
    """C++ bucketed AUC calculator (csrc/host/metrics.cc)."""

    def __init__(self, table_size: int = 1000000):
        self.table_size = table_size
        self._h = C.c_void_p()
        N.hcheck(N.host().pbh_auc_create(C.byref(self._h), table_size))

    def __del__(self):
        try:
            if self._h:
                N.host().pbh_auc_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def reset(self):
        N.hcheck(N.host().pbh_auc_reset(self._h))

    def add(self, pred, label, mask=None, float_label: bool = False):
        p = np.ascontiguousarray(pred, dtype=np.float32).reshape(-1)
        y = np.ascontiguousarray(label, dtype=np.float32).reshape(-1)
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.float32).reshape(-1)
        N.hcheck(N.host().pbh_auc_add(self._h, _f32p(p), _f32p(y), _f32p(m) if m is not None else None,
                                      C.c_int64(p.size), int(float_label)))

    def add_uid(self, pred, label, uid):
        p = np.ascontiguousarray(pred, dtype=np.float32).reshape(-1)
        y = np.ascontiguousarray(label, dtype=np.float32).reshape(-1)
        u = np.ascontiguousarray(uid).astype(np.uint64).reshape(-1)
        N.hcheck(N.host().pbh_auc_add_uid(self._h, _f32p(p), _f32p(y), _u64p(u), C.c_int64(p.size)))

    def merge_tables(self, tables: np.ndarray, sums: np.ndarray):
        t = np.ascontiguousarray(tables, dtype=np.float64)
        s = np.ascontiguousarray(sums, dtype=np.float64)
        N.hcheck(N.host().pbh_auc_merge_tables(self._h, t.ctypes.data_as(C.POINTER(C.c_double)),
                                               s.ctypes.data_as(C.POINTER(C.c_double))))

    def export_state(self):
        t = np.empty((2, self.table_size), dtype=np.float64)
        s = np.empty(5, dtype=np.float64)
        N.hcheck(N.host().pbh_auc_export(self._h, t.ctypes.data_as(C.POINTER(C.c_double)),
                                         s.ctypes.data_as(C.POINTER(C.c_double))))
        return t, s

    def import_state(self, t, s):
        t = np.ascontiguousarray(t, dtype=np.float64)
        s = np.ascontiguousarray(s, dtype=np.float64)
        N.hcheck(N.host().pbh_auc_import(self._h, t.ctypes.data_as(C.POINTER(C.c_double)),
                                         s.ctypes.data_as(C.POINTER(C.c_double))))

    def compute(self) -> list:
        out = (C.c_double * 8)()
        N.hcheck(N.host().pbh_auc_compute(self._h, out))
        return list(out)

    def compute_wuauc(self) -> list:
        out = (C.c_double * 4)()
        N.hcheck(N.host().pbh_auc_compute_wuauc(self._h, out))
        return list(out)

    def nan_inf(self):
        out = (C.c_double * 2)()
        N.hcheck(N.host().pbh_auc_nan_inf(self._h, out))
        return out[0], out[1]


class AsyncDenseTable:
    """Host-resident dense parameters + Adam applied by a background thread
    (reference BoxPSAsynDenseTable, boxps_worker.cc:61-370)."""

    def __init__(self, numel: int, lr: float = 1e-3, threads: int = 8, merge_limit: int = 4):
        self.numel = numel
        self._h = C.c_void_p()
        N.hcheck(N.host().pbh_async_dense_create(C.byref(self._h), C.c_int64(numel), C.c_float(lr), threads,
                                                 merge_limit))
        N.host().pbh_async_dense_updates.restype = C.c_int64

    def __del__(self):
        try:
            if self._h:
                N.host().pbh_async_dense_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def set_params(self, flat: torch.Tensor):
        a = np.ascontiguousarray(flat.detach().cpu().numpy(), dtype=np.float32)
        N.hcheck(N.host().pbh_async_dense_set_params(self._h, _f32p(a)))

    def pull(self, out: torch.Tensor | None = None) -> torch.Tensor:
        """Copy of the current parameter vector (``PullDense``)."""
        host = torch.empty(self.numel, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        N.hcheck(N.host().pbh_async_dense_pull(self._h, C.cast(host.data_ptr(), C.POINTER(C.c_float))))
        if out is not None:
            out.copy_(host, non_blocking=True)
            return out
        return host

    def push(self, grad: torch.Tensor):
        """Queue one gradient vector (``PushDense``)."""
        g = np.ascontiguousarray(grad.detach().cpu().numpy(), dtype=np.float32)
        N.hcheck(N.host().pbh_async_dense_push(self._h, _f32p(g)))

    def set_lr_range(self, begin: int, end: int, lr: float):
        """Learning rate of the flat range ``[begin, end)`` (one parameter tensor; ``lr_map``)."""
        N.hcheck(N.host().pbh_async_dense_set_lr_range(self._h, C.c_int64(begin), C.c_int64(end), C.c_float(lr)))

    def wait(self):
        N.hcheck(N.host().pbh_async_dense_wait(self._h))

    def updates(self) -> int:
        return N.host().pbh_async_dense_updates(self._h)


class NativeDataset:
    """Thin handle of the C++ ``pbh::Dataset`` (csrc/host/dataset.cc)."""

    def __init__(self):
        lib = N.host()
        self._h = C.c_void_p()
        N.hcheck(lib.pbh_dataset_create(C.byref(self._h)))
        for fn in ("pbh_dataset_num_records", "pbh_dataset_num_batches", "pbh_dataset_total_feasigns",
                   "pbh_dataset_unique_keys", "pbh_dataset_batch_ins_ids", "pbh_dataset_stats",
                   "pbh_compute_batch_ranges"):
            getattr(lib, fn).restype = C.c_int64

    def __del__(self):
        try:
            if self._h:
                N.host().pbh_dataset_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def set_slots(self, slots: Sequence[dict]):
        n = len(slots)
        names = (C.c_char_p * n)(*[s["name"].encode() for s in slots])
        arr = lambda key: (C.c_int * n)(*[int(s[key]) for s in slots])
        N.hcheck(N.host().pbh_dataset_set_slots(self._h, n, names, arr("type"), arr("is_dense"), arr("is_used"),
                                                arr("dim")))

    def set_filelist(self, files: Sequence[str]):
        n = len(files)
        arr = (C.c_char_p * n)(*[f.encode() for f in files])
        N.hcheck(N.host().pbh_dataset_set_filelist(self._h, n, arr))

    def set(self, key: str, value):
        lib = N.host()
        if isinstance(value, str):
            N.hcheck(lib.pbh_dataset_set_str(self._h, key.encode(), value.encode()))
        elif isinstance(value, float):
            N.hcheck(lib.pbh_dataset_set_float(self._h, key.encode(), C.c_double(value)))
        else:
            N.hcheck(lib.pbh_dataset_set_int(self._h, key.encode(), C.c_int64(int(value))))

    def set_reader_cores(self, cores: Sequence[int]):
        arr = (C.c_int * len(cores))(*cores)
        N.hcheck(N.host().pbh_dataset_set_reader_cores(self._h, arr, len(cores)))

    def set_omit_slots(self, slot_indices: Sequence[int]):
        arr = (C.c_int * len(slot_indices))(*slot_indices)
        N.hcheck(N.host().pbh_dataset_set_omit_slots(self._h, arr, len(slot_indices)))

    def copy_config_from(self, other: "NativeDataset") -> None:
        """Everything but the file list and the rank / world of the file dealing."""
        N.hcheck(N.host().pbh_dataset_copy_config(self._h, other._h))

    def set_input_index(self, index_handle):
        """Attach an ``InputTable`` key index (parse-time ``GetOffsetFunc`` for parser plugins)."""
        N.hcheck(N.host().pbh_dataset_set_input_index(self._h, index_handle))

    def load_into_memory(self):
        N.hcheck(N.host().pbh_dataset_load_into_memory(self._h), "load_into_memory")

    def preload_into_memory(self):
        N.hcheck(N.host().pbh_dataset_preload_into_memory(self._h))

    def wait_preload_done(self):
        N.hcheck(N.host().pbh_dataset_wait_preload_done(self._h), "wait_preload_done")

    def release_memory(self):
        N.hcheck(N.host().pbh_dataset_release_memory(self._h))

    def trim_record_pool(self) -> int:
        return int(N.host().pbh_dataset_trim_record_pool(self._h))

    def idle_records(self) -> int:
        return int(N.host().pbh_dataset_idle_records(self._h))

    def set_record_pool_cap(self, cap: int) -> None:
        N.hcheck(N.host().pbh_dataset_set_record_pool_cap(self._h, C.c_int64(int(cap))))

    def local_shuffle(self, seed: int = 0):
        N.hcheck(N.host().pbh_dataset_local_shuffle(self._h, C.c_uint64(seed)))

    def slots_shuffle(self, slot_indices: Sequence[int], seed: int = 0):
        n = len(slot_indices)
        N.hcheck(N.host().pbh_dataset_slots_shuffle(self._h, (C.c_int * n)(*slot_indices), n, C.c_uint64(seed)))

    def merge_by_ins_id(self, merge_size: int = 0) -> int:
        fn = N.host().pbh_dataset_merge_by_ins_id
        fn.restype = C.c_int64
        return int(fn(self._h, int(merge_size)))

    def slots_shuffle_back(self, slot_indices: Sequence[int], seed: int = 0):
        n = len(slot_indices)
        N.hcheck(N.host().pbh_dataset_slots_shuffle_back(self._h, (C.c_int * n)(*slot_indices), n, C.c_uint64(seed)))

    def slot_first_values(self, slot_index: int):
        """``(values uint64 [n_records], has bool [n_records])``: the first feasign of one sparse slot per record."""
        n = self.num_records()
        out, has = np.zeros(n, dtype=np.uint64), np.zeros(n, dtype=np.uint8)
        N.hcheck(N.host().pbh_dataset_slot_first_values(self._h, int(slot_index), _u64p(out),
                                                        has.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out, has.astype(bool)

    def expand_records(self, slot_index: int, label_slot_index: int, counts, ids, neg) -> None:
        """Record i -> ``counts[i]`` copies with the slot's first feasign replaced by the next ``ids``; copies
        flagged in ``neg`` get the label slot's first value zeroed (TDM layer-wise sampling)."""
        counts = np.ascontiguousarray(counts, dtype=np.int64)
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        neg = np.ascontiguousarray(neg, dtype=np.uint8)
        assert counts.size == self.num_records() and ids.size == int(counts.sum()) == neg.size
        N.hcheck(N.host().pbh_dataset_expand_records(self._h, int(slot_index), int(label_slot_index),
                                                     counts.ctypes.data_as(C.POINTER(C.c_int64)), _u64p(ids),
                                                     neg.ctypes.data_as(C.POINTER(C.c_uint8))))

    def prepare_train(self, nworkers: int, shuffle: bool = True, seed: int = 0, min_batches: int = 0):
        N.hcheck(N.host().pbh_dataset_prepare_train(self._h, nworkers, int(shuffle), C.c_uint64(seed),
                                                    C.c_int64(min_batches)))

    def num_records(self) -> int:
        return N.host().pbh_dataset_num_records(self._h)

    def num_batches(self, worker: int = 0) -> int:
        return N.host().pbh_dataset_num_batches(self._h, worker)

    def n_sparse(self) -> int:
        return N.host().pbh_dataset_n_sparse(self._h)

    def dense_dim(self) -> int:
        return N.host().pbh_dataset_dense_dim(self._h)

    def total_feasigns(self) -> int:
        return N.host().pbh_dataset_total_feasigns(self._h)

    def batch_shape(self, worker: int, b: int):
        B, n = C.c_int(), C.c_int64()
        N.hcheck(N.host().pbh_dataset_batch_shape(self._h, worker, C.c_int64(b), C.byref(B), C.byref(n)))
        return B.value, n.value

    def pack_batch(self, worker: int, b: int, keys_ptr: int, off_ptr: int, dense_ptr: int, slot_major: bool):
        N.hcheck(N.host().pbh_dataset_pack_batch(self._h, worker, C.c_int64(b), C.c_void_p(keys_ptr),
                                                 C.c_void_p(off_ptr), C.c_void_p(dense_ptr), int(slot_major)))

    def batch_meta(self, worker: int, b: int, B: int):
        sid = np.empty(B, dtype=np.uint64)
        rk = np.empty(B, dtype=np.uint32)
        cm = np.empty(B, dtype=np.uint32)
        N.hcheck(N.host().pbh_dataset_batch_meta(self._h, worker, C.c_int64(b), _u64p(sid),
                                                 rk.ctypes.data_as(C.POINTER(C.c_uint32)),
                                                 cm.ctypes.data_as(C.POINTER(C.c_uint32))))
        return sid, rk, cm

    def batch_ins_ids(self, worker: int, b: int) -> list:
        need = N.host().pbh_dataset_batch_ins_ids(self._h, worker, C.c_int64(b), None, C.c_int64(0))
        buf = C.create_string_buffer(max(int(need), 1))
        N.host().pbh_dataset_batch_ins_ids(self._h, worker, C.c_int64(b), buf, C.c_int64(len(buf)))
        return buf.value.decode().split("\n")

    def unique_keys(self) -> np.ndarray:
        n = N.host().pbh_dataset_unique_keys(self._h, None, C.c_int64(0))
        out = np.empty(n, dtype=np.uint64)
        if n:
            N.host().pbh_dataset_unique_keys(self._h, _u64p(out), C.c_int64(n))
        return out

    def save_arena(self, path: str):
        N.hcheck(N.host().pbh_dataset_save_arena(self._h, path.encode()), "pbh_dataset_save_arena")

    def load_arena(self, path: str):
        N.hcheck(N.host().pbh_dataset_load_arena(self._h, path.encode()), "pbh_dataset_load_arena")

    def save_archive(self, path: str):
        N.hcheck(N.host().pbh_dataset_save_archive(self._h, path.encode()))

    def load_archive(self, path: str):
        N.hcheck(N.host().pbh_dataset_load_archive(self._h, path.encode()))

    def save_partitioned(self, prefix: str, nparts: int, mode: int = 0, seed: int = 0):
        """Write ``<prefix>.<rank>.<dst>`` partition files (mode 0 random, 1 search_id, 2 ins_id)."""
        N.hcheck(N.host().pbh_dataset_save_partitioned(self._h, prefix.encode(), nparts, mode, C.c_uint64(seed)))

    def stats(self) -> str:
        buf = C.create_string_buffer(1024)
        N.host().pbh_dataset_stats(self._h, buf, C.c_int64(1024))
        return buf.value.decode()

    def parse_line(self, line: str):
        nu, nf, sid = C.c_int64(), C.c_int64(), C.c_uint64()
        u = np.zeros(4096, dtype=np.uint64)
        f = np.zeros(4096, dtype=np.float32)
        rc = N.host().pbh_dataset_parse_line(self._h, line.encode(), C.byref(nu), C.byref(nf), _u64p(u),
                                             C.c_int64(4096), _f32p(f), C.c_int64(4096), C.byref(sid))
        if rc != 0:
            return None
        return u[:nu.value].copy(), f[:nf.value].copy(), sid.value


def sort_unique_u64(keys: np.ndarray) -> np.ndarray:
    """Sorted unique keys (parallel sample sort on the host pool; ``np.unique`` for large key sets)."""
    k = np.array(keys, dtype=np.uint64, copy=True).reshape(-1)
    fn = N.host().pbh_sort_unique_u64
    fn.restype = C.c_int64
    n = fn(_u64p(k), C.c_int64(k.size))
    return k[:n]


def partition_by_owner(keys: np.ndarray, nranks: int) -> list:
    """``[keys owned by rank 0, by rank 1, ...]`` (views of one array; owner = ``pbx_owner_of``)."""
    k = np.ascontiguousarray(keys, dtype=np.uint64).reshape(-1)
    out = np.empty_like(k)
    counts = np.zeros(nranks, dtype=np.int64)
    N.hcheck(N.host().pbh_partition_by_owner(_u64p(k), C.c_int64(k.size), nranks, _u64p(out),
                                             counts.ctypes.data_as(C.POINTER(C.c_int64))))
    ends = np.cumsum(counts)
    return [out[e - c:e] for c, e in zip(counts.tolist(), ends.tolist())]


def compute_batch_ranges(ins_num: int, batch_size: int, nworkers: int):
    lib = N.host()
    lib.pbh_compute_batch_ranges.restype = C.c_int64
    cap = ins_num // max(batch_size, 1) + 4 * nworkers + 8
    starts = (C.c_int64 * cap)()
    lens = (C.c_int * cap)()
    n = lib.pbh_compute_batch_ranges(C.c_int64(ins_num), batch_size, nworkers, starts, lens, C.c_int64(cap))
    return [(starts[i], lens[i]) for i in range(n)]
