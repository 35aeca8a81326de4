"""Small dense side tables and the file manager that hang off BoxWrapper.

* ``GpuReplicaCache`` — a per-pass float table built on the host, replicated to every GPU and
  gathered by row index (ops ``pull_cache_value`` / ``push_cache_value``;
  fleet/box_wrapper.h:63-122, kernel box_wrapper.cu:1210-1224).
* ``InputTable`` — host ``string -> row`` table; data lines carry the string key, the feed
  replaces it by the row index and the op ``lookup_input`` gathers the row
  (fleet/box_wrapper.h:124-197, pull_box_sparse_op.h:52-89).
* ``BoxFileMgr`` — file-system wrapper exposed to Python (fleet/box_wrapper.h:1016-1041,
  box_wrapper.cc:1326-1397).  The reference talks to AFS/HDFS through the closed library; here
  local paths are handled natively and ``hdfs:// / afs://`` paths go through a configurable
  shell client (``hadoop fs`` style), which is how the open-source ``io/fs.cc`` does it too.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import threading

import torch

from .ops.functional import row_gather


class GpuReplicaCache:
    def __init__(self, emb_dim: int | None = None):
        if emb_dim is None:
            from .flags import FLAGS
            emb_dim = FLAGS.gpu_replica_cache_dim       # FLAGS_gpu_replica_cache_dim (flags.cc)
        self.emb_dim = emb_dim
        self._host: list = []
        self._lock = threading.Lock()
        self._dev: dict = {}

    def add_items(self, emb) -> int:
        """Append one row (list of ``emb_dim`` floats); returns its index (``AddItems``)."""
        assert len(emb) == self.emb_dim
        with self._lock:
            self._host.append([float(v) for v in emb])
            return len(self._host) - 1

    def to_hbm(self, device=None) -> None:
        """Replicate the table to the GPU (``ToHBM``)."""
        device = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available()
                            else torch.device("cpu"))
        t = torch.tensor(self._host, dtype=torch.float32).reshape(-1, self.emb_dim)
        self._dev[str(device)] = t.to(device)

    def pull_cache_value(self, idx: torch.Tensor) -> torch.Tensor:
        """[n, emb_dim] rows for int64 indices (op ``pull_cache_value``)."""
        tab = self._dev.get(str(idx.device))
        if tab is None:
            self.to_hbm(idx.device)
            tab = self._dev[str(idx.device)]
        return row_gather(tab, idx.reshape(-1))

    def push_cache_value(self, idx: torch.Tensor, grad: torch.Tensor) -> None:
        """The replica cache is read-only during a pass; the grad op is a no-op in the
        reference as well (pull_box_sparse_op.cc ``push_cache_value``)."""

    def size(self) -> int:
        return len(self._host)

    def clear(self) -> None:
        with self._lock:
            self._host.clear()
            self._dev.clear()


class InputTable:
    """Host ``string -> row`` side table (``InputTable``, fleet/box_wrapper.h:124-197).  The key
    index is native (``pbh_input_index_*``) so that parser plugins can resolve keys at parse time
    (``GetOffsetFunc``); the rows are gathered on the device by ``lookup_input``."""

    def __init__(self, dim: int):
        import ctypes as C
        from . import _native as N
        self.dim = dim
        self._h = C.c_void_p()
        lib = N.host()
        N.hcheck(lib.pbh_input_index_create(C.byref(self._h)))
        for fn in ("add", "get", "size", "miss"):
            getattr(lib, f"pbh_input_index_{fn}").restype = C.c_uint64
        self._lib = lib
        self._rows: list = [[0.0] * dim]       # row 0 = miss
        self._lock = threading.Lock()
        self._dev = None

    def __del__(self):
        try:
            if self._h:
                self._lib.pbh_input_index_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    @property
    def miss(self) -> int:
        return int(self._lib.pbh_input_index_miss(self._h))

    def add_index_data(self, key: str, vec) -> None:
        import ctypes as C
        assert len(vec) == self.dim
        kb = key.encode()
        fresh = C.c_int()
        with self._lock:
            self._lib.pbh_input_index_add(self._h, kb, C.c_size_t(len(kb)), C.byref(fresh))
            if fresh.value:
                self._rows.append([float(v) for v in vec])
            self._dev = None

    def load_index_file(self, path: str, sep: str = "\t") -> int:
        """Lines ``key<sep>v1 v2 ...`` (the reference's InputIndexDataFeed, data_feed.cc:4574-4633)."""
        n = 0
        with open(path) as f:
            for line in f:
                k, _, rest = line.rstrip("\n").partition(sep)
                vals = [float(x) for x in rest.split()]
                if len(vals) == self.dim:
                    self.add_index_data(k, vals)
                    n += 1
        return n

    def get_index_offset(self, key: str) -> int:
        import ctypes as C
        kb = key.encode()
        return int(self._lib.pbh_input_index_get(self._h, kb, C.c_size_t(len(kb))))

    def lookup_input(self, idx: torch.Tensor) -> torch.Tensor:
        """Rows for int64 offsets (op ``lookup_input``)."""
        if self._dev is None or self._dev.device != idx.device:
            self._dev = torch.tensor(self._rows, dtype=torch.float32, device=idx.device)
        return row_gather(self._dev, idx.reshape(-1))

    def size(self) -> int:
        return len(self._rows) - 1


class BoxFileMgr:
    def __init__(self):
        self._client = None
        self._conf = {}

    def init(self, fs_name: str = "", fs_ugi: str = "", conf_path: str = "", hadoop_bin: str = "hadoop") -> bool:
        self._conf = {"fs_name": fs_name, "fs_ugi": fs_ugi, "conf": conf_path}
        self._client = [hadoop_bin, "fs"]
        if fs_name:
            self._client += [f"-Dfs.default.name={fs_name}"]
        if fs_ugi:
            self._client += [f"-Dhadoop.job.ugi={fs_ugi}"]
        return True

    def destory(self) -> None:     # (sic) the reference spells it this way
        self._client = None

    destroy = destory

    @staticmethod
    def _remote(path: str) -> bool:
        return path.startswith(("hdfs:", "afs:"))

    def _run(self, *args) -> subprocess.CompletedProcess:
        if self._client is None:
            raise RuntimeError("BoxFileMgr.init() has not been called for remote paths")
        return subprocess.run(self._client + list(args), capture_output=True, text=True)

    def list_dir(self, path: str) -> list:
        if self._remote(path):
            out = self._run("-ls", path).stdout.splitlines()
            return [ln.split()[-1] for ln in out if ln and not ln.startswith("Found")]
        return sorted(os.path.join(path, n) for n in os.listdir(path))

    def list_info(self, path: str) -> list:
        """[(path, size in bytes)] of the entries of ``path`` (``BoxFileMgr::list_info``)."""
        if self._remote(path):
            out = self._run("-ls", path).stdout.splitlines()
            rows = [ln.split() for ln in out if ln and not ln.startswith("Found")]
            return [(r[-1], int(r[4])) for r in rows if len(r) >= 8]
        return [(p, os.path.getsize(p)) for p in sorted(os.path.join(path, n) for n in os.listdir(path))]

    def makedir(self, path: str) -> bool:
        if self._remote(path):
            return self._run("-mkdir", "-p", path).returncode == 0
        os.makedirs(path, exist_ok=True)
        return True

    def exists(self, path: str) -> bool:
        if self._remote(path):
            return self._run("-test", "-e", path).returncode == 0
        return os.path.exists(path)

    def download(self, remote: str, local: str) -> bool:
        if self._remote(remote):
            return self._run("-get", remote, local).returncode == 0
        shutil.copy(remote, local)
        return True

    def upload(self, local: str, remote: str) -> bool:
        if self._remote(remote):
            return self._run("-put", "-f", local, remote).returncode == 0
        shutil.copy(local, remote)
        return True

    def remove(self, path: str) -> bool:
        if self._remote(path):
            return self._run("-rm", "-r", path).returncode == 0
        if os.path.isdir(path):
            shutil.rmtree(path)
        elif os.path.exists(path):
            os.remove(path)
        return True

    def file_size(self, path: str) -> int:
        if self._remote(path):
            out = self._run("-du", "-s", path).stdout.split()
            return int(out[0]) if out else -1
        return os.path.getsize(path)

    def dus(self, path: str) -> list:
        if self._remote(path):
            return [(ln.split()[-1], int(ln.split()[0])) for ln in self._run("-du", path).stdout.splitlines() if ln]
        out = []
        for root, _, files in os.walk(path):
            for n in files:
                p = os.path.join(root, n)
                out.append((p, os.path.getsize(p)))
        return out

    def truncate(self, path: str, length: int) -> bool:
        if self._remote(path):
            return self._run("-truncate", str(length), path).returncode == 0
        with open(path, "r+b") as f:
            f.truncate(length)
        return True

    def touch(self, path: str) -> bool:
        if self._remote(path):
            return self._run("-touchz", path).returncode == 0
        open(path, "a").close()
        return True

    def rename(self, src: str, dst: str) -> bool:
        if self._remote(src):
            return self._run("-mv", src, dst).returncode == 0
        os.replace(src, dst)
        return True

    def count(self, path: str) -> int:
        """Number of entries directly under ``path``."""
        return len(self.list_dir(path))

    def finalize(self) -> None:
        self.destory()
