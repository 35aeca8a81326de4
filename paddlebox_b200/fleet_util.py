"""``FleetUtil`` / ``GPUPSUtil`` — the job-script helpers PaddleBox users drive a training day with.

Reference: ``python/paddle/fluid/incubate/fleet/utils/fleet_util.py`` (``FleetUtil`` :42-1735, ``GPUPSUtil``
:1736-2200).  The production loop around ``BoxWrapper`` / ``PSGPUWrapper`` is written against it:

    for day in days:
        for pass_id, splits in enumerate(util.get_online_pass_interval(...), 1):
            ... load, begin_pass, train, end_pass ...
            util.print_global_metrics(...)
            util.save_delta_model(out, day, pass_id);  util.write_xbox_donefile(out, day, pass_id, key, data, ...)
        util.save_xbox_base_model(out, day);  util.write_xbox_donefile(out, day, -1, key, data, ...)
        util.save_batch_model(out, day);      util.write_model_donefile(out, day, -1, key, ...)
    # restart:  day, pass_id, path, key = util.get_last_save_model(out, ...);  util.load_model(out, day, pass_id)

What is kept byte-compatible is what OTHER systems read: the directory layout under ``output_path``
(``<day>/<pass>/``, ``<day>/0/``, ``<day>/delta-<pass>/``, ``<day>/base/``), the tab-separated ``donefile.txt``
and the JSON lines of ``xbox_base_done.txt`` / ``xbox_patch_done.txt`` (field names, the ``/000`` suffix of
``input``, the "only append what is newer" rule), and the layout of ``cache_donefile.txt``.  What differs is
where the numbers come from: metrics are read from the wrapper's metric registry (``init_metric``) instead of
named scope variables, models are saved through ``BoxWrapper.save_base / save_delta`` and the dense
persistables through ``checkpoint.py``, and files go through ``BoxFileMgr`` (local paths, or ``hadoop fs`` for
``hdfs:`` / ``afs:``).  The program-drawing / proto helpers of the reference (``draw_from_program``,
``parse_program_proto``, ``check_two_programs`` ...) have no counterpart: there is no program proto here.
"""
from __future__ import annotations

import json
import math
import os
import re
import sys
import tempfile
import time
from collections import OrderedDict
from typing import Sequence

import numpy as np

from . import fleet
from .side_tables import BoxFileMgr

METRIC_NAMES = ("auc", "bucket_error", "mae", "rmse", "actual_ctr", "predicted_ctr", "copc", "mean_predict_qvalue",
                "total_ins_num")


def expand_braces(pattern: str) -> list:
    """``"{20190720..20190729}"`` -> the ten days; ``"{0..23}"`` -> 0..23; ``"a,b"`` / ``"a b"`` -> [a, b]; zero
    padding of the left bound is kept (bash semantics; the reference shells out to ``echo`` for this)."""
    pattern = str(pattern).strip()
    m = re.fullmatch(r"\{(-?\d+)\.\.(-?\d+)\}", pattern)
    if m:
        a, b = m.group(1), m.group(2)
        width = max(len(a), len(b)) if (a.startswith("0") and len(a) > 1) or (b.startswith("0") and len(b) > 1) else 0
        lo, hi = int(a), int(b)
        step = 1 if hi >= lo else -1
        return [str(v).zfill(width) for v in range(lo, hi + step, step)]
    m = re.fullmatch(r"\{(.*)\}", pattern)
    if m:
        return [p for p in m.group(1).split(",") if p != ""]
    return [p for p in re.split(r"[,\s]+", pattern) if p]


def bucket_error(pos: np.ndarray, neg: np.ndarray, max_span: float = 0.01, rel_bound: float = 0.05) -> float:
    """Calibration error over the prediction buckets: neighbouring buckets are pooled until the pool is large
    enough for its click rate to be known to ``rel_bound`` relative error (or spans ``max_span`` of the
    prediction axis), then |actual / predicted - 1| of the pool is averaged, weighted by impressions
    (the reference's ``bucket_error``, fleet_util.py:1403-1444, fleet/metrics.cc:352-397)."""
    pos = np.asarray(pos, dtype=np.float64).reshape(-1)
    neg = np.asarray(neg, dtype=np.float64).reshape(-1)
    n = pos.size
    last, imp, ctr_sum, clk = -1.0, 0.0, 0.0, 0.0
    err_sum = err_cnt = 0.0
    for i in range(n):          # (every bucket, empty ones too: they take part in the span rule)
        show = pos[i] + neg[i]
        ctr = float(i) / n
        if abs(ctr - last) > max_span:
            last, imp, ctr_sum, clk = ctr, 0.0, 0.0, 0.0
        imp += show
        ctr_sum += ctr * show
        clk += pos[i]
        if imp == 0:
            continue
        adjust = ctr_sum / imp
        if adjust == 0:
            continue
        if math.sqrt((1 - adjust) / (adjust * imp)) < rel_bound:
            err_sum += abs(clk / imp / adjust - 1) * imp
            err_cnt += imp
            last = -1.0
    return err_sum / err_cnt if err_cnt > 0 else 0.0


def auc_from_buckets(pos: np.ndarray, neg: np.ndarray) -> float:
    """Area under the ROC curve from per-bucket positive / negative counts, walking the buckets from the
    highest prediction down (trapezoids); 0.5 when one class is absent (fleet_util.py:223-258)."""
    pos = np.asarray(pos, dtype=np.float64).reshape(-1)[::-1]
    neg = np.asarray(neg, dtype=np.float64).reshape(-1)[::-1]
    cp, cn = np.cumsum(pos), np.cumsum(neg)
    area = float(np.sum((cn - np.concatenate([[0.0], cn[:-1]])) * (cp + np.concatenate([[0.0], cp[:-1]])) / 2.0))
    tot_p, tot_n = float(cp[-1]) if cp.size else 0.0, float(cn[-1]) if cn.size else 0.0
    if tot_p * tot_n == 0:
        return 0.5
    return area / (tot_p * tot_n)


class _FS:
    """The few file operations the donefiles need, on local paths or through ``hadoop fs``."""

    def __init__(self, mgr: BoxFileMgr | None = None):
        self.mgr = mgr or BoxFileMgr()

    def configure(self, fs_name: str = "", fs_ugi: str = "", hadoop_home: str = "") -> None:
        if fs_name or fs_ugi:
            hb = os.path.join(os.path.expandvars(hadoop_home), "bin", "hadoop") if hadoop_home else "hadoop"
            self.mgr.init(fs_name, fs_ugi, hadoop_bin=hb if os.path.exists(hb) else "hadoop")

    def is_file(self, path: str) -> bool:
        if self.mgr._remote(path):
            return self.mgr._run("-test", "-f", path).returncode == 0
        return os.path.isfile(path)

    def cat(self, path: str) -> str:
        if self.mgr._remote(path):
            return self.mgr._run("-cat", path).stdout
        with open(path) as f:
            return f.read()

    def write(self, path: str, content: str) -> None:
        """Replace ``path`` by ``content`` (readers never see a partial file on a local path)."""
        if self.mgr._remote(path):
            with tempfile.NamedTemporaryFile("w", suffix=".done", delete=False) as f:
                f.write(content)
                tmp = f.name
            try:
                self.mgr.makedir(os.path.dirname(path))
                if not self.mgr.upload(tmp, path):
                    raise OSError(f"upload to {path} failed")
            finally:
                os.remove(tmp)
            return
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        tmp = f"{path}.tmp.{os.getpid()}"
        with open(tmp, "w") as f:
            f.write(content)
            f.flush()
            os.fsync(f.fileno())
        os.replace(tmp, path)


class FleetUtil:
    """See the module docstring.  ``box``: the ``BoxWrapper`` / ``PSGPUWrapper`` of the job (default: the
    process-wide instance); ``trainer``: a ``CTRTrainer`` whose dense parameters travel with the models."""

    def __init__(self, mode: str = "pslib", box=None, trainer=None, fs: BoxFileMgr | None = None):
        self.mode = mode
        self._box, self.trainer = box, trainer
        self._fs = _FS(fs)

    # ---- plumbing ---------------------------------------------------------------------------------------
    @property
    def box(self):
        if self._box is None:
            from .box_wrapper import BoxWrapper
            self._box = BoxWrapper.get_instance()
        return self._box

    def set_box(self, box) -> None:
        self._box = box

    def set_trainer(self, trainer) -> None:
        self.trainer = trainer

    def rank0_print(self, s) -> None:
        if fleet.worker_index() == 0:
            print(s)
            sys.stdout.flush()

    def rank0_info(self, s) -> None:
        if fleet.worker_index() == 0:
            import logging
            logging.getLogger("paddlebox_b200").info(s)

    def rank0_error(self, s) -> None:
        if fleet.worker_index() == 0:
            import logging
            logging.getLogger("paddlebox_b200").error(s)

    @staticmethod
    def _paths(output_path: str, day, pass_id, kind: str) -> str:
        out, day, pass_id = output_path.rstrip("/"), str(day), str(pass_id)
        if kind == "model":
            return f"{out}/{day}/{pass_id if pass_id != '-1' else '0'}/"
        if kind == "xbox":
            return f"{out}/{day}/delta-{pass_id}/" if pass_id != "-1" else f"{out}/{day}/base/"
        raise ValueError(kind)

    # ---- metrics ------------------------------------------------------------------------------------------
    def set_zero(self, metric_name: str | None = None, **_) -> None:
        """Reset one metric (or all) of the registry — the reference zeroes the named stat variables."""
        names = [metric_name] if metric_name else self.box.get_metric_name_list()
        for n in names:
            m = self.box.get_metric(n)
            m.calc.reset()
            if getattr(m, "_dev_table", None) is not None:
                m._dev_table.zero_()
                m._dev_sums.zero_()

    def get_global_metrics(self, metric_name: str = "auc", **_) -> list:
        """``[auc, bucket_error, mae, rmse, actual_ctr, predicted_ctr, copc, mean_predict_qvalue,
        total_ins_num]`` over all workers; reads out and resets the metric.  (``mean_predict_qvalue`` is the mean
        prediction: the q-value stat of the reference's example net is not part of the registry metrics.)"""
        if metric_name not in self.box.get_metric_name_list():
            self.rank0_print(f"not found metric {metric_name}")
            return [None] * 9
        msg = self.box.get_metric_msg(metric_name)
        auc, berr, mae, rmse, actual, pred, copc, size = msg[:8]
        return [auc, berr, mae, rmse, actual, pred, copc, pred, int(size)]

    def get_global_auc(self, metric_name: str = "auc", **_):
        m = self.get_global_metrics(metric_name)
        return m[0]

    def print_global_auc(self, metric_name: str = "auc", print_prefix: str = "", **_) -> None:
        self.rank0_print(f"{print_prefix} global auc = {self.get_global_auc(metric_name)}")

    def print_global_metrics(self, metric_name: str = "auc", print_prefix: str = "", **_) -> list:
        m = self.get_global_metrics(metric_name)
        if m[0] is None:
            return m
        self.rank0_print("%s global AUC=%.6f BUCKET_ERROR=%.6f MAE=%.6f RMSE=%.6f Actural_CTR=%.6f Predicted_CTR=%.6f "
                         "COPC=%.6f MEAN Q_VALUE=%.6f Ins number=%s" % ((print_prefix,) + tuple(m)))
        return m

    @staticmethod
    def metrics_from_stats(stat_pos, stat_neg, sqrerr: float, abserr: float, prob: float, q: float, pos_ins: float,
                           total_ins: float) -> list:
        """The same nine numbers from raw (already all-reduced) statistics, for nets that keep their own bucket
        tensors the way the reference's example does (fleet_util.py:1257-1444)."""
        total = float(total_ins)
        if total <= 0:
            return [0.5, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0]
        actual, pred = float(pos_ins) / total, float(prob) / total
        return [auc_from_buckets(stat_pos, stat_neg), bucket_error(stat_pos, stat_neg), float(abserr) / total,
                math.sqrt(float(sqrerr) / total), actual, pred, (actual / pred if abs(pred) > 1e-6 else 0.0),
                float(q) / total, int(total)]

    # ---- models --------------------------------------------------------------------------------------------
    def _save_dense(self, path: str) -> None:
        """Dense persistables travel with every full model (the reference's pslib keeps them in the same
        model directory as dense tables); written by rank 0 into ``<model>/dense``."""
        if self.trainer is not None and fleet.worker_index() == 0:
            from . import checkpoint
            checkpoint.save_persistables(self.trainer.model, os.path.join(path, "dense"), self.trainer.store)

    def _load_dense(self, path: str) -> None:
        d = os.path.join(path, "dense")
        if self.trainer is not None and os.path.isdir(d):
            from . import checkpoint
            checkpoint.load_persistables(self.trainer.model, d, self.trainer.store, strict=False)
            self.trainer.store.mark_dirty()

    def save_fleet_model(self, path: str, mode: int = 0) -> int:
        """``mode`` 0 = full training state, 1 = delta (serving format), 2 = serving base, 3 = batch model."""
        box = self.box
        box._sync_to_host()
        # only a delta save moves the delta baseline (UpdateStatAfterSave, ctr_dymf_accessor.cc:151-175)
        pol = fleet.table_policy          # thresholds of the job's strategy (fleet.distributed_optimizer), else everything
        n = box.host_table.save(path, mode=0 if mode == 3 else mode, shard_base=box.rank, clear_dirty=(mode == 1),
                                base_threshold=float(pol.get("base_threshold", 0.0)) if mode == 2 else 0.0,
                                delta_threshold=float(pol.get("delta_threshold", 0.0)) if mode == 1 else 0.0)
        if mode in (0, 3):
            self._save_dense(path)
        fleet.barrier_worker()
        return n

    def load_fleet_model(self, path: str, mode: int = 0) -> int:
        n = self.box.load_model(path, serving=mode in (1, 2))
        self._load_dense(path)
        fleet.barrier_worker()
        return n

    def load_fleet_model_one_table(self, table_id, path: str) -> int:
        return self.load_fleet_model(path)          # one sparse table per wrapper

    def load_model(self, output_path: str, day, pass_id) -> int:
        path = self._paths(output_path, day, pass_id, "model")
        self.rank0_error(f"going to load_model {path}")
        n = self.load_fleet_model(path)
        self.rank0_error("load_model done")
        return n

    def save_model(self, output_path: str, day, pass_id) -> int:
        path = self._paths(output_path, day, pass_id, "model")
        self.rank0_print(f"going to save_model {path}")
        n = self.save_fleet_model(path, 0)
        self.rank0_print("save_model done")
        return n

    def save_batch_model(self, output_path: str, day) -> int:
        path = self._paths(output_path, day, -1, "model")           # <day>/0/
        self.rank0_print(f"going to save_model {path}")
        n = self.save_fleet_model(path, 3)
        self.rank0_print("save_batch_model done")
        return n

    def save_delta_model(self, output_path: str, day, pass_id) -> int:
        path = self._paths(output_path, day, pass_id, "xbox")       # <day>/delta-<pass>/
        self.rank0_print(f"going to save_delta_model {path}")
        n = self.save_fleet_model(path, 1)
        self.rank0_print("save_delta_model done")
        return n

    def save_xbox_base_model(self, output_path: str, day) -> int:
        path = self._paths(output_path, day, -1, "xbox")            # <day>/base/
        self.rank0_print(f"going to save_xbox_base_model {path}")
        n = self.save_fleet_model(path, 2)
        self.rank0_print("save_xbox_base_model done")
        return n

    def save_cache_model(self, output_path: str, day, pass_id, mode: int = 1, **kwargs) -> int:
        """The serving cache: the hot rows (score >= ``cache_threshold``, default the table's base threshold)
        in serving format under ``<day>/delta-<pass>/<table>_cache``; returns the number of keys over all
        workers (fleet_util.py:752-787)."""
        table_id = int(kwargs.get("table_id", 0))
        base = self._paths(output_path, day, pass_id, "xbox").rstrip("/")
        return self._save_cache(f"{base}/{table_id:03d}_cache", float(kwargs.get("cache_threshold", 1.0)))

    def save_cache_base_model(self, output_path: str, day, **kwargs) -> int:
        table_id = int(kwargs.get("table_id", 0))
        base = self._paths(output_path, day, -1, "xbox").rstrip("/")
        return self._save_cache(f"{base}/{table_id:03d}_cache", float(kwargs.get("cache_threshold", 1.0)))

    def _save_cache(self, path: str, threshold: float) -> int:
        box = self.box
        self.rank0_print(f"going to save_cache_model {path}")
        box._sync_to_host()
        n = box.host_table.save(path, mode=2, shard_base=box.rank, base_threshold=threshold, clear_dirty=False)
        total = int(fleet.util.all_reduce(int(n), "sum"))
        self.rank0_print("save_cache_model done")
        return total

    def pull_all_dense_params(self, scope=None, program=None) -> dict:
        """Name -> tensor of every dense parameter (the reference copies them from the PS into the scope;
        here every rank already holds the synchronised values)."""
        if self.trainer is None:
            return {}
        model = self.trainer.model
        named = model.named_dense_params() if hasattr(model, "named_dense_params") else model.named_parameters()
        return {n: p.detach().cpu() for n, p in named}

    def save_paddle_params(self, output_path: str = "", day="", pass_id="", **_) -> str:
        """Dense parameters alone, in the reference's persistables format, under ``<day>/<pass>/dnn_plugin``
        (``save_paddle_params`` / ``save_paddle_inference_model``, fleet_util.py:866-1054)."""
        path = os.path.join(self._paths(output_path, day, pass_id, "model"), "dnn_plugin")
        if self.trainer is not None and fleet.worker_index() == 0:
            from . import checkpoint
            checkpoint.save_persistables(self.trainer.model, path)
        fleet.barrier_worker()
        return path

    save_paddle_inference_model = save_paddle_params

    # ---- donefiles ---------------------------------------------------------------------------------------------
    def _xbox_str(self, output_path: str, day: str, model_path: str, xbox_base_key: int, data_path: str,
                  hadoop_fs_name: str, mode: str) -> str:
        d = OrderedDict()
        d["id"] = str(xbox_base_key) if mode == "base" else str(int(time.time()))
        d["key"] = str(xbox_base_key)
        if model_path.startswith(("hdfs:", "afs:")):
            model_path = model_path[model_path.find(":") + 1:]
        d["input"] = (hadoop_fs_name or "") + model_path.rstrip("/") + "/000"
        d["record_count"] = "111111"
        d["partition_type"] = "2"
        d["job_name"] = "default_job_name"
        d["ins_tag"] = "feasign"
        d["ins_path"] = data_path
        job, inst = os.environ.get("JOB_ID", ""), os.environ.get("INSTANCE_ID", "")
        a, b = inst.find(job), inst.find("--")
        d["job_id"] = inst[a:b] if (job and a != -1 and b != -1) else job
        d["monitor_data"] = ""
        d["monitor_path"] = output_path.rstrip("/") + "/monitor/" + day + ".txt"
        d["mpi_size"] = str(fleet.worker_num())
        return json.dumps(d)

    def write_model_donefile(self, output_path: str, day, pass_id, xbox_base_key, hadoop_fs_name: str = "",
                             hadoop_fs_ugi: str = "", hadoop_home: str = "$HADOOP_HOME",
                             donefile_name: str = "donefile.txt") -> bool:
        """One line ``day \\t key \\t model_path \\t pass_id \\t 0`` per saved model; a (day, pass) that is
        already listed is not written twice.  Returns whether a line was added (rank 0)."""
        day, pass_id, key = str(day), str(pass_id), int(xbox_base_key)
        model_path = self._paths(output_path, day, pass_id, "model")
        wrote = False
        if fleet.worker_index() == 0:
            self._fs.configure(hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
            path = output_path.rstrip("/") + "/" + donefile_name
            line = "%s\t%d\t%s\t%s\t%d" % (day, key, model_path, pass_id, 0)
            pre = self._fs.cat(path).strip("\n") if self._fs.is_file(path) else ""
            rows = [r.split("\t") for r in pre.split("\n") if r.strip()]
            if any(len(r) > 3 and int(r[0]) == int(day) and int(r[3]) == int(pass_id) for r in rows):
                self.rank0_error(f"not write {donefile_name} because {day}/{pass_id} already exists")
            else:
                self._fs.write(path, (pre + "\n" if pre else "") + line + "\n")
                self.rank0_error(f"write {day}/{pass_id} {donefile_name} succeed")
                wrote = True
        fleet.barrier_worker()
        return wrote

    def write_xbox_donefile(self, output_path: str, day, pass_id, xbox_base_key, data_path, hadoop_fs_name: str = "",
                            hadoop_fs_ugi: str = "", monitor_data=None, hadoop_home: str = "$HADOOP_HOME",
                            donefile_name: str | None = None) -> bool:
        """The serving side's manifest: one JSON line per delta (``xbox_patch_done.txt``) or per base model
        (``pass_id = -1``: ``xbox_base_done.txt``); only a model newer than the last listed one is appended."""
        day, pass_id, key = str(day), str(pass_id), int(xbox_base_key)
        mode = "patch" if pass_id != "-1" else "base"
        model_path = self._paths(output_path, day, pass_id, "xbox")
        donefile_name = donefile_name or ("xbox_patch_done.txt" if mode == "patch" else "xbox_base_done.txt")
        if isinstance(data_path, (list, tuple)):
            data_path = ",".join(data_path)
        wrote = False
        if fleet.worker_index() == 0:
            self._fs.configure(hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
            path = output_path.rstrip("/") + "/" + donefile_name
            entry = self._xbox_str(output_path, day, model_path, key, data_path, hadoop_fs_name, mode)
            pre = self._fs.cat(path).strip("\n") if self._fs.is_file(path) else ""
            stale = False
            if pre:
                last = json.loads(pre.split("\n")[-1])["input"].split("/")
                last_day, last_pass = int(last[-3]), last[-2].split("-")[-1]
                last_pass = int(last_pass) if last_pass.lstrip("-").isdigit() else -1      # ".../base/000"
                stale = int(day) < last_day or (int(day) == last_day and int(pass_id) <= last_pass)
            if stale:
                self.rank0_error(f"not write {donefile_name} because {day}/{pass_id} already exists")
            else:
                self._fs.write(path, (pre + "\n" if pre else "") + entry + "\n")
                self.rank0_error(f"write {day}/{pass_id} {donefile_name} succeed")
                wrote = True
        fleet.barrier_worker()
        return wrote

    def write_cache_donefile(self, output_path: str, day, pass_id, key_num, hadoop_fs_name: str = "",
                             hadoop_fs_ugi: str = "", hadoop_home: str = "$HADOOP_HOME",
                             donefile_name: str = "sparse_cache.meta", **kwargs) -> bool:
        day, pass_id = str(day), str(pass_id)
        table_id, file_num = int(kwargs.get("table_id", 0)), kwargs.get("file_num", 16)
        model_path = self._paths(output_path, day, pass_id, "xbox").rstrip("/") + f"/{table_id:03d}_cache"
        wrote = False
        if fleet.worker_index() == 0:
            self._fs.configure(hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
            path = model_path + "/" + donefile_name
            if self._fs.is_file(path):
                self.rank0_error(f"not write because {path} already exists")
            else:
                self._fs.write(path, "file_prefix:part\npart_num:%s\nkey_num:%d\n" % (file_num, int(key_num)))
                self.rank0_error(f"write {path} succeed")
                wrote = True
        fleet.barrier_worker()
        return wrote

    def _last_json(self, output_path: str, name: str, fs_name: str, fs_ugi: str, hadoop_home: str):
        self._fs.configure(fs_name, fs_ugi, hadoop_home)
        path = output_path.rstrip("/") + "/" + name
        if not self._fs.is_file(path):
            return None
        rows = [r for r in self._fs.cat(path).split("\n") if r.strip()]
        return json.loads(rows[-1]) if rows else None

    def get_last_save_xbox_base(self, output_path: str, hadoop_fs_name: str = "", hadoop_fs_ugi: str = "",
                                hadoop_home: str = "$HADOOP_HOME") -> list:
        """``[day, path, xbox_base_key]`` of the last base model, ``[-1, -1, now]`` when there is none."""
        last = self._last_json(output_path, "xbox_base_done.txt", hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
        if last is None:
            return [-1, -1, int(time.time())]
        parts = last["input"].split("/")
        return [int(parts[-3]), "/".join(parts[:-1]), int(last["key"])]

    def get_last_save_xbox(self, output_path: str, hadoop_fs_name: str = "", hadoop_fs_ugi: str = "",
                           hadoop_home: str = "$HADOOP_HOME") -> list:
        """``[day, pass, path, xbox_base_key]`` of the last delta, ``[-1, -1, "", now]`` when there is none."""
        last = self._last_json(output_path, "xbox_patch_done.txt", hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
        if last is None:
            return [-1, -1, "", int(time.time())]
        parts = last["input"].split("/")
        return [int(parts[-3]), int(parts[-2].split("-")[-1]), "/".join(parts[:-1]), int(last["key"])]

    def get_last_save_model(self, output_path: str, hadoop_fs_name: str = "", hadoop_fs_ugi: str = "",
                            hadoop_home: str = "$HADOOP_HOME") -> list:
        """``[day, pass, path, xbox_base_key]`` of the last line of ``donefile.txt``."""
        self._fs.configure(hadoop_fs_name, hadoop_fs_ugi, hadoop_home)
        path = output_path.rstrip("/") + "/donefile.txt"
        if not self._fs.is_file(path):
            return [-1, -1, "", int(time.time())]
        rows = [r for r in self._fs.cat(path).split("\n") if r.strip()]
        if not rows:
            return [-1, -1, "", int(time.time())]
        c = rows[-1].split("\t")
        return [int(c[0]), int(c[3]), c[2], int(c[1])]

    # ---- the online schedule -----------------------------------------------------------------------------------
    def get_online_pass_interval(self, days, hours, split_interval, split_per_pass, is_data_hourly_placed) -> list:
        """The day cut into passes: ``split_interval`` minutes per data split, ``split_per_pass`` splits per
        pass, only the hours in ``hours`` (``"{0..23}"``); a split is named ``HHMM`` (or ``HH`` when the data is
        placed hourly).  Returns ``[[split, ...] per pass]`` (fleet_util.py:1196-1255)."""
        hours_l = expand_braces(hours)
        split_interval, split_per_pass = int(split_interval), int(split_per_pass)
        splits_per_day = 24 * 60 // split_interval
        left, right = int(hours_l[0]), int(hours_l[-1])
        names = []
        for i in range(splits_per_day):
            h, m = divmod(i * split_interval, 60)
            if left <= h <= right:
                names.append("%02d" % h if is_data_hourly_placed else "%02d%02d" % (h, m))
        passes = splits_per_day // split_per_pass
        out = []
        for p in range(passes):
            chunk = names[p * split_per_pass:(p + 1) * split_per_pass]
            if len(chunk) < split_per_pass:
                break                          # (the reference indexes past the end here when hours are restricted)
            out.append(chunk)
        return out


class GPUPSUtil(FleetUtil):
    """``FleetUtil`` for the GPU parameter server (fleet_util.py:1736-2200): the same helpers with the file-system
    client handed in (``set_fsclient`` / ``init``) instead of per-call hadoop arguments."""

    def __init__(self, fs_client=None, box=None, trainer=None):
        super().__init__("pslib", box=box, trainer=trainer, fs=fs_client if isinstance(fs_client, BoxFileMgr) else None)
        self._afs = fs_client

    def init(self, fs_name: str, fs_user: str, fs_passwd: str, fs_conf: str = "") -> None:
        self._fs.mgr.init(fs_name, f"{fs_user},{fs_passwd}", conf_path=fs_conf)

    def set_fsclient(self, fs_client) -> None:
        self._afs = fs_client
        if isinstance(fs_client, BoxFileMgr):
            self._fs = _FS(fs_client)

    # the GPUPS flavour takes no hadoop arguments: positional compatibility with the reference's signatures
    def get_last_save_xbox_base(self, output_path: str, *_, **__) -> list:
        return super().get_last_save_xbox_base(output_path)

    def get_last_save_xbox(self, output_path: str, *_, **__) -> list:
        return super().get_last_save_xbox(output_path)

    def get_last_save_model(self, output_path: str, *_, **__) -> list:
        return super().get_last_save_model(output_path)

    def write_model_donefile(self, output_path, day, pass_id, xbox_base_key, donefile_name="donefile.txt", **_):
        return super().write_model_donefile(output_path, day, pass_id, xbox_base_key, donefile_name=donefile_name)

    def write_xbox_donefile(self, output_path, day, pass_id, xbox_base_key, data_path, hadoop_fs_name="",
                            monitor_data=None, donefile_name=None, **_):
        return super().write_xbox_donefile(output_path, day, pass_id, xbox_base_key, data_path,
                                           hadoop_fs_name=hadoop_fs_name, donefile_name=donefile_name)

    def write_cache_donefile(self, output_path, day, pass_id, key_num, donefile_name="sparse_cache.meta", **kwargs):
        return super().write_cache_donefile(output_path, day, pass_id, key_num, donefile_name=donefile_name, **kwargs)
