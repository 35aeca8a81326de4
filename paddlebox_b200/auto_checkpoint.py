"""Numbered checkpoints on any file system and the epoch loop that resumes from the last one.

Reference: ``python/paddle/fluid/incubate/checkpoint/checkpoint_saver.py`` (``SerializableBase`` :18-25,
``PaddleModel`` :27-51, ``CheckpointSaver`` :53-223) and ``auto_checkpoint.py`` (``train_epoch_range`` and the
``PADDLE_EDL_*`` environment, :60-140, :560-685).  A checkpoint is the directory
``<path>/__paddle_checkpoint__.<n>``: written as ``....<n>.tmp`` (through a local cache directory when the file
system is remote) and renamed when complete, so a reader never sees a partial one; ``n`` counts up, old ones are
removed by ``clean_redundant_checkpoints``.  What goes into it is a list of serializable objects:
``ModelCheckpoint`` (dense parameters + optimizer moments in the reference's LoDTensor files, and optionally the
sparse table of the process-wide wrapper) and ``TrainStatus`` (epoch counter).  ``train_epoch_range`` is the
loop a job puts around its epochs: it restores the last checkpoint, yields the epochs that remain and saves at
the end of an epoch when the save interval has passed.  The day / pass recovery of online jobs (donefiles,
``FleetUtil``) is separate: fleet_util.py.
"""
from __future__ import annotations

import json
import os
import time

from .utils.fs import FS, LocalFS

_PREFIX = "__paddle_checkpoint__"


class SerializableBase:
    def serialize(self, path: str) -> None:
        raise NotImplementedError

    def deserialize(self, path: str) -> None:
        raise NotImplementedError


class ModelCheckpoint(SerializableBase):
    """Dense state of ``model`` (and of ``trainer``'s optimizer); with ``box`` also this rank's sparse table."""

    def __init__(self, model=None, trainer=None, box=None, name: str = "model"):
        self.model = model if model is not None else getattr(trainer, "model", None)
        self.trainer, self.box, self.name = trainer, box, name

    def _store(self):
        return getattr(self.trainer, "store", None)

    def serialize(self, path: str) -> None:
        from . import checkpoint
        checkpoint.save_persistables(self.model, os.path.join(path, self.name), self._store())
        if self.box is not None:
            self.box._sync_to_host()
            self.box.host_table.save(os.path.join(path, self.name + "_sparse"), mode=0, shard_base=self.box.rank, clear_dirty=False)

    def deserialize(self, path: str) -> None:
        from . import checkpoint
        checkpoint.load_persistables(self.model, os.path.join(path, self.name), self._store())
        if self._store() is not None and hasattr(self._store(), "mark_dirty"):
            self._store().mark_dirty()
        sparse = os.path.join(path, self.name + "_sparse")
        if self.box is not None and os.path.isdir(sparse):
            self.box.host_table.clear()
            self.box.host_table.load(sparse, rank=self.box.rank, world=self.box.world_size)


class TrainStatus(SerializableBase):
    """Epoch counter (auto_checkpoint.py ``TrainEpochRange`` state): ``epoch_no`` is the last FINISHED epoch."""

    def __init__(self, epoch_no: int = -1, **extra):
        self.epoch_no = int(epoch_no)
        self.extra = dict(extra)

    def serialize(self, path: str) -> None:
        with open(os.path.join(path, "train_status.json"), "w") as f:
            json.dump({"epoch_no": self.epoch_no, **self.extra}, f)

    def deserialize(self, path: str) -> None:
        with open(os.path.join(path, "train_status.json")) as f:
            d = json.load(f)
        self.epoch_no = int(d.pop("epoch_no"))
        self.extra = d

    def next(self) -> int:
        return self.epoch_no + 1

    def __eq__(self, other):
        return isinstance(other, TrainStatus) and (self.epoch_no, self.extra) == (other.epoch_no, other.extra)


class CheckpointSaver:
    def __init__(self, fs: FS | None = None):
        self._fs = fs or LocalFS()
        self._checkpoint_prefix = _PREFIX

    def _name(self, root: str, n: int) -> str:
        return f"{root}/{self._checkpoint_prefix}.{n}"

    def save_checkpoint(self, path: str, slists, trainer_id: int | None = None, local_cache_path: str = ".cache"):
        """Serialize every object of ``slists`` into the next numbered checkpoint; ``(real_path, checkpoint_no)``."""
        fs = self._fs
        if not fs.is_exist(path):
            fs.mkdirs(path)
        elif not fs.is_dir(path):
            raise ValueError(f"path:{path} must be a directory")
        no = max(self._get_last_checkpoint_no(path), -1) + 1
        real, tmp = self._name(path, no), self._name(path, no) + ".tmp"
        local = LocalFS()
        remote = fs.need_upload_download()
        if remote:
            saved = f"{local_cache_path}/{self._checkpoint_prefix}.{no}.saved_cache" + (f".{trainer_id}" if trainer_id is not None else "")
        else:
            saved = tmp
        if local.is_exist(saved):
            local.delete(saved)
        local.mkdirs(saved)
        for s in slists:
            s.serialize(saved)
        if remote:
            fs.delete(tmp)
            fs.upload(saved, tmp)
            local.delete(saved)
        fs.mv(tmp, real)
        return real, no

    def load_checkpoint(self, path: str, slists, trainer_id: int | None = None, local_cache_path: str = ".cache",
                        checkpoint_no: int | None = None, ignore_empty: bool = True):
        """Deserialize ``slists`` from checkpoint ``checkpoint_no`` (default: the newest); the path read, or None
        when there is none and ``ignore_empty``."""
        fs = self._fs
        if checkpoint_no is None:
            checkpoint_no = self._get_last_checkpoint_no(path) if fs.is_exist(path) else -1
            if checkpoint_no < 0:
                if not ignore_empty:
                    raise FileNotFoundError(f"Can't find checkpoint under {path}")
                return None
        elif not isinstance(checkpoint_no, int) or checkpoint_no < 0:
            raise ValueError("checkpoint_no must be a non-negative int")
        real = self._name(path, checkpoint_no)
        load = real
        local = LocalFS()
        if fs.need_upload_download():
            load = f"{local_cache_path}/{self._checkpoint_prefix}.{checkpoint_no}.load_cache" + \
                (f".{trainer_id}" if trainer_id is not None else "")
            if not local.is_exist(local_cache_path):
                local.mkdirs(local_cache_path)
            if local.is_exist(load):
                local.delete(load)
            fs.download(real, load)
        for s in slists:
            s.deserialize(load)
        if load != real:
            local.delete(load)
        return real

    def get_checkpoint_no(self, root_path: str) -> list:
        out = []
        for d in self._fs.list_dirs(root_path):
            g = os.path.basename(d.rstrip("/")).split(".")
            if len(g) == 2 and g[0] == self._checkpoint_prefix and g[1].isdigit():
                out.append(int(g[1]))
        return sorted(out)

    def _get_last_checkpoint_no(self, root_path: str) -> int:
        a = self.get_checkpoint_no(root_path)
        return a[-1] if a else -1

    def clean_redundant_checkpoints(self, root_path: str, reserved=()) -> None:
        """Delete every checkpoint but ``reserved`` (default: but the newest)."""
        last = self._get_last_checkpoint_no(root_path)
        if last < 0:
            return
        keep = set(reserved) or {last}
        for n in self.get_checkpoint_no(root_path):
            if n not in keep:
                self._fs.delete(self._name(root_path, n))


def train_epoch_range(max_epoch_num: int, save_checkpoint_inter: float | None = None, path: str | None = None,
                      objects=(), fs: FS | None = None, trainer_id: int | None = None, keep: int = 1):
    """``for epoch in train_epoch_range(10, objects=[ModelCheckpoint(trainer=tr, box=box)], path=...)``: the epochs
    still to run.  Before the first one the newest checkpoint under ``path`` is restored into ``objects``; after
    an epoch a checkpoint is written if ``save_checkpoint_inter`` seconds have passed since the last one (0 =
    every epoch) — always after the last epoch.  Rank 0 writes unless the objects are rank-local (pass
    ``trainer_id``: then every rank writes under ``<path>/rank_<id>``).  Defaults come from the reference's
    environment: ``PADDLE_EDL_HDFS_CHECKPOINT_PATH`` + ``PADDLE_JOB_ID``, ``PADDLE_EDL_SAVE_CHECKPOINT_INTER`` (900 s),
    ``PADDLE_EDL_FS_CACHE``.  Without a path the loop is a plain ``range``."""
    env = os.environ
    if path is None and env.get("PADDLE_EDL_HDFS_CHECKPOINT_PATH"):
        path = os.path.join(env["PADDLE_EDL_HDFS_CHECKPOINT_PATH"], env.get("PADDLE_JOB_ID", "job"))
    if save_checkpoint_inter is None:
        save_checkpoint_inter = float(env.get("PADDLE_EDL_SAVE_CHECKPOINT_INTER", "900"))
    if path is None:
        yield from range(max_epoch_num)
        return
    cache = env.get("PADDLE_EDL_FS_CACHE", ".cache")
    if trainer_id is not None:
        path = f"{path}/rank_{trainer_id}"
    saver = CheckpointSaver(fs)
    status = TrainStatus()
    objs = list(objects) + [status]
    saver.load_checkpoint(path, objs, trainer_id=trainer_id, local_cache_path=cache)
    writer = trainer_id is not None or int(env.get("RANK", env.get("PADDLE_TRAINER_ID", "0"))) == 0
    last_save = time.time()
    for epoch in range(status.next(), max_epoch_num):
        yield epoch
        status.epoch_no = epoch
        if writer and (epoch == max_epoch_num - 1 or time.time() - last_save >= save_checkpoint_inter):
            _, no = saver.save_checkpoint(path, objs, trainer_id=trainer_id, local_cache_path=cache)
            saver.clean_redundant_checkpoints(path, reserved=range(no - keep + 1, no + 1))
            last_save = time.time()
