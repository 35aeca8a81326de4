"""File-system clients with the reference's method names (``paddle.distributed.fleet.utils.fs``:
``LocalFS`` fs.py:120-388, ``HDFSClient`` :428-1154, ``AFSClient`` :1156-1580) — what job scripts use next to
``FleetUtil`` to list data splits, wait for ``_SUCCESS`` files and move models around.

``LocalFS`` works on the local tree.  ``HDFSClient`` drives ``$hadoop_home/bin/hadoop fs -D<k>=<v> ...`` as a
subprocess (retries with a back-off, one overall time-out per operation); ``AFSClient`` is the same shell client
configured by ``init(fs_name, fs_user, fs_passwd, fs_conf)`` — the reference binds AFS through a closed client
library, for which the ``hadoop fs`` front end of an AFS installation is the open equivalent.  Shared semantics:
``ls_dir`` -> ``([dir names], [file names])``; ``mkdirs`` / ``delete`` are idempotent; ``mv`` and ``upload`` refuse
to overwrite unless told to; ``touch(exist_ok=False)`` and ``mv`` raise the FS* exceptions below.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import time
from typing import Sequence


class ExecuteError(Exception):
    pass


class FSFileExistsError(Exception):
    pass


class FSFileNotExistsError(Exception):
    pass


class FSTimeOut(Exception):
    pass


class FSShellCmdAborted(ExecuteError):
    pass


class FS:
    def need_upload_download(self) -> bool:
        raise NotImplementedError

    def list_files_info(self, path_list: Sequence[str]) -> list:
        """``[{"path": p, "size": bytes}]`` for every existing file of ``path_list``."""
        return [{"path": p, "size": self._size(p)} for p in path_list if self.is_file(p)]

    @staticmethod
    def _split_files(files: Sequence[str], trainer_id: int, trainers: int) -> list:
        """This trainer's contiguous share of ``files`` (the first ``len % trainers`` get one more)."""
        per, rem = divmod(len(files), trainers)
        begin = trainer_id * per + min(trainer_id, rem)
        return list(files[begin:begin + per + (1 if trainer_id < rem else 0)])


class LocalFS(FS):
    def ls_dir(self, fs_path: str):
        if not self.is_exist(fs_path):
            return [], []
        dirs, files = [], []
        for name in os.listdir(fs_path):
            (dirs if os.path.isdir(os.path.join(fs_path, name)) else files).append(name)
        return dirs, files

    def list_dirs(self, fs_path: str) -> list:
        return self.ls_dir(fs_path)[0]

    def mkdirs(self, fs_path: str) -> None:
        assert not os.path.isfile(fs_path), f"{fs_path} is already a file"
        os.makedirs(fs_path, exist_ok=True)

    def rename(self, fs_src_path: str, fs_dst_path: str) -> None:
        os.rename(fs_src_path, fs_dst_path)

    def delete(self, fs_path: str) -> None:
        if not self.is_exist(fs_path):
            return
        if os.path.isfile(fs_path) or os.path.islink(fs_path):
            os.remove(fs_path)
        else:
            shutil.rmtree(fs_path)

    def need_upload_download(self) -> bool:
        return False

    def is_file(self, fs_path: str) -> bool:
        return os.path.isfile(fs_path)

    def is_dir(self, fs_path: str) -> bool:
        return os.path.isdir(fs_path)

    def is_exist(self, fs_path: str) -> bool:
        return os.path.exists(fs_path)

    def touch(self, fs_path: str, exist_ok: bool = True) -> None:
        if self.is_exist(fs_path):
            if exist_ok:
                return
            raise FSFileExistsError(fs_path)
        with open(fs_path, "a"):
            pass

    def mv(self, src_path: str, dst_path: str, overwrite: bool = False, test_exists: bool = False) -> None:
        if not self.is_exist(src_path):
            raise FSFileNotExistsError(src_path)
        if overwrite and self.is_exist(dst_path):
            self.delete(dst_path)
        if self.is_exist(dst_path):
            raise FSFileExistsError(dst_path)
        self.rename(src_path, dst_path)

    def upload(self, local_path: str, fs_path: str, multi_processes: int = 1, overwrite: bool = False) -> None:
        self._copy(local_path, fs_path, overwrite)

    def download(self, fs_path: str, local_path: str, multi_processes: int = 1, overwrite: bool = False) -> None:
        self._copy(fs_path, local_path, overwrite)

    def upload_dir(self, local_dir: str, dest_dir: str, overwrite: bool = False) -> None:
        self._copy(local_dir, dest_dir, overwrite)

    def _copy(self, src: str, dst: str, overwrite: bool) -> None:
        if not self.is_exist(src):
            raise FSFileNotExistsError(src)
        if self.is_exist(dst):
            if not overwrite:
                raise FSFileExistsError(dst)
            self.delete(dst)
        if os.path.isdir(src):
            shutil.copytree(src, dst)
        else:
            os.makedirs(os.path.dirname(os.path.abspath(dst)), exist_ok=True)
            shutil.copy(src, dst)

    def cat(self, fs_path: str | None = None) -> str:
        if fs_path and self.is_file(fs_path):
            with open(fs_path) as f:
                return f.read().rstrip("\n")
        return ""

    def _size(self, path: str) -> int:
        return os.path.getsize(path)


class _ShellFS(FS):
    """``hadoop fs`` as a subprocess.  ``time_out`` / ``sleep_inter`` in milliseconds, as in the reference."""

    def __init__(self, hadoop_home: str = "", configs: dict | None = None, time_out: int = 5 * 60 * 1000,
                 sleep_inter: int = 1000):
        home = os.path.expandvars(hadoop_home or "")
        self._cmd = [os.path.join(home, "bin", "hadoop") if home else "hadoop", "fs"]
        for k, v in (configs or {}).items():
            self._cmd.append(f"-D{k}={v}")
        self._time_out, self._sleep_inter = time_out / 1000.0, sleep_inter / 1000.0

    def _run(self, *args, retry_times: int = 5, ok_codes=(0,)):
        """(return code, stdout lines); a failing command is retried until the operation's time-out."""
        t0 = time.time()
        ret, out = 1, ""
        for attempt in range(retry_times + 1):
            try:
                p = subprocess.run(self._cmd + list(args), capture_output=True, text=True,
                                   timeout=max(1.0, self._time_out - (time.time() - t0)))
            except subprocess.TimeoutExpired:
                raise FSTimeOut(f"hadoop fs {' '.join(args)}: no answer within {self._time_out:.0f} s") from None
            except FileNotFoundError:
                raise ExecuteError(f"{self._cmd[0]} not found") from None
            ret, out = p.returncode, p.stdout
            if ret in ok_codes:
                break
            if ret == 134:
                raise FSShellCmdAborted(" ".join(args))
            if time.time() - t0 + self._sleep_inter > self._time_out:
                raise FSTimeOut(f"hadoop fs {' '.join(args)}: still failing after {time.time() - t0:.0f} s")
            if attempt < retry_times:
                time.sleep(self._sleep_inter)
        return ret, out.splitlines()

    def need_upload_download(self) -> bool:
        return True

    # ---- queries (a "test" that answers 1 is an answer, not a failure: no retries) ------------------------
    def _test(self, flag: str, fs_path: str) -> bool:
        ret, _ = self._run("-test", flag, fs_path, retry_times=0, ok_codes=(0, 1))
        return ret == 0

    def is_exist(self, fs_path: str) -> bool:
        return self._test("-e", fs_path)

    def is_dir(self, fs_path: str) -> bool:
        return self._test("-d", fs_path)

    def is_file(self, fs_path: str) -> bool:
        return self.is_exist(fs_path) and not self.is_dir(fs_path)

    def ls_dir(self, fs_path: str):
        if not self.is_exist(fs_path):
            return [], []
        ret, lines = self._run("-ls", fs_path)
        if ret != 0:
            raise ExecuteError(f"ls {fs_path}")
        dirs, files = [], []
        for ln in lines:
            f = ln.split()
            if len(f) != 8:                      # "Found n items" and anything that is not an entry
                continue
            name = os.path.basename(f[7].rstrip("/"))
            (dirs if f[0].startswith("d") else files).append(name)
        return dirs, files

    def list_dirs(self, fs_path: str) -> list:
        return self.ls_dir(fs_path)[0]

    def _size(self, path: str) -> int:
        _, lines = self._run("-ls", path)
        for ln in lines:
            f = ln.split()
            if len(f) == 8:
                return int(f[4])
        return -1

    def cat(self, fs_path: str | None = None) -> str:
        if not fs_path or not self.is_file(fs_path):
            return ""
        ret, lines = self._run("-cat", fs_path)
        if ret != 0:
            raise ExecuteError(f"cat {fs_path}")
        return "\n".join(lines)

    # ---- changes ----------------------------------------------------------------------------------------------
    def mkdirs(self, fs_path: str) -> None:
        if self.is_exist(fs_path):
            return
        ret, _ = self._run("-mkdir", "-p", fs_path)
        if ret != 0:
            raise ExecuteError(f"mkdir {fs_path}")

    def delete(self, fs_path: str) -> None:
        if not self.is_exist(fs_path):
            return
        ret, _ = self._run("-rmr" if self.is_dir(fs_path) else "-rm", fs_path)
        if ret != 0:
            raise ExecuteError(f"rm {fs_path}")

    def touch(self, fs_path: str, exist_ok: bool = True) -> None:
        if self.is_exist(fs_path):
            if exist_ok:
                return
            raise FSFileExistsError(fs_path)
        ret, _ = self._run("-touchz", fs_path)
        if ret != 0:
            raise ExecuteError(f"touchz {fs_path}")

    def mv(self, fs_src_path: str, fs_dst_path: str, overwrite: bool = False, test_exists: bool = True) -> None:
        if overwrite and self.is_exist(fs_dst_path):
            self.delete(fs_dst_path)
        if test_exists:
            if not self.is_exist(fs_src_path):
                raise FSFileNotExistsError(fs_src_path)
            if self.is_exist(fs_dst_path):
                raise FSFileExistsError(fs_dst_path)
        ret, _ = self._run("-mv", fs_src_path, fs_dst_path, retry_times=1)
        if ret != 0:
            raise ExecuteError(f"mv {fs_src_path} {fs_dst_path}")

    rename = mv

    def upload(self, local_path: str, fs_path: str, multi_processes: int = 1, overwrite: bool = False) -> None:
        if not os.path.exists(local_path):
            raise FSFileNotExistsError(local_path)
        if self.is_exist(fs_path):
            if not overwrite:
                raise FSFileExistsError(fs_path)
            self.delete(fs_path)
        ret, _ = self._run("-put", local_path, fs_path)
        if ret != 0:
            raise ExecuteError(f"put {local_path} {fs_path}")

    def upload_dir(self, local_dir: str, dest_dir: str, overwrite: bool = False) -> None:
        self.upload(local_dir, dest_dir, overwrite=overwrite)

    def download(self, fs_path: str, local_path: str, multi_processes: int = 1, overwrite: bool = False) -> None:
        if not self.is_exist(fs_path):
            raise FSFileNotExistsError(fs_path)
        if os.path.exists(local_path):
            if not overwrite:
                raise FSFileExistsError(local_path)
            LocalFS().delete(local_path)
        ret, _ = self._run("-get", fs_path, local_path)
        if ret != 0:
            raise ExecuteError(f"get {fs_path} {local_path}")


class HDFSClient(_ShellFS):
    """``HDFSClient(hadoop_home, {"fs.default.name": ..., "hadoop.job.ugi": ...})``."""


class AFSClient(_ShellFS):
    """``AFSClient(); client.init(fs_name, fs_user, fs_passwd, fs_conf)``."""

    def __init__(self, time_out: int = 5 * 60 * 1000, sleep_inter: int = 1000, hadoop_home: str = ""):
        super().__init__(hadoop_home or os.environ.get("HADOOP_HOME", ""), None, time_out, sleep_inter)

    def init(self, fs_name: str, fs_user: str, fs_passwd: str, fs_conf: str = "") -> None:
        self._cmd = self._cmd[:2]
        if fs_conf:
            self._cmd += ["-conf", fs_conf]
        self._cmd += [f"-Dfs.default.name={fs_name}", f"-Dhadoop.job.ugi={fs_user},{fs_passwd}"]
