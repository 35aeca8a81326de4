"""AUC-family metrics (reference: ``MetricMsg`` and its subclasses,
fleet/box_wrapper.h:265-376, box_wrapper.cc:372-886; ``BasicAucCalculator``,
fleet/metrics.cc:38-409, :501-587; registry ``InitMetric`` box_wrapper.cc:916-1025).

Each metric owns a C++ ``HostAuc`` calculator (csrc/host/metrics.cc).  CUDA predictions are
bucketed *on the device* by ``pbx_auc_bucket`` (csrc/cuda/dense_ops.cu) into a
``[2, buckets]`` fp64 table that is merged into the calculator only when the metric is
read — the reference copies pred/label to the host and buckets under a mutex every batch
(metrics.cc:115-136).  ``get_metric_msg`` reduces the tables over the data-parallel group
(reference: MPI / Gloo allreduce, metrics.cc:284-332) and returns
``[auc, bucket_error, mae, rmse, actual_ctr, predicted_ctr, copc, size]``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as N
from .host import HostAuc


def parse_cmatch_rank(x: torch.Tensor):
    """High 32 bits = cmatch, low 8 bits = rank (box_wrapper.h:364-368)."""
    x = x.to(torch.int64)
    return (x >> 32) & 0xFFFFFFFF, x & 0xFF


def accumulate_float_labels(table: torch.Tensor, sums: torch.Tensor, pred: torch.Tensor, label: torch.Tensor,
                            mask: torch.Tensor | None, mode: int) -> None:
    """Device-side twin of ``AucCalculator::add_data`` for float labels (csrc/host/metrics.cc; reference
    fleet/metrics.cc:75-136): ``mode`` 1 = a label in [0, 1] is split fractionally between the negative and the
    positive table; ``mode`` 2 = continuous values, only the five sums.  ``table`` fp64 [2, n], ``sums`` fp64 [5] =
    (abs err, sq err, pred sum, label sum, count).  Runs on whatever device the tensors live on."""
    p = pred.reshape(-1).double()
    y = label.reshape(-1).double()
    w = torch.ones_like(p) if mask is None else (mask.reshape(-1) != 0).double()
    if mode != 2:
        p = torch.nan_to_num(p, nan=0.0, posinf=1.0, neginf=0.0).clamp_(0.0, 1.0)
        n = table.shape[1]
        pos = (p * n).long().clamp_(max=n - 1)
        table[0].index_add_(0, pos, (1.0 - y) * w)
        table[1].index_add_(0, pos, y * w)
    d = p - y
    sums += torch.stack([(d.abs() * w).sum(), (d * d * w).sum(), (p * w).sum(), (y * w).sum(), w.sum()])


class MetricMsg:
    def __init__(self, name="", label_varname="label", pred_varname="pred", metric_phase=1, bucket_size=1000000,
                 device=None, float_label=False, **_):
        self.name = name
        self.label_varname, self.pred_varname = label_varname, pred_varname
        self.metric_phase = metric_phase
        self.bucket_size = bucket_size
        self.float_label = float_label
        self.calc = HostAuc(bucket_size)
        self.device = device
        self._dev_table = None
        self._dev_sums = None

    # ---- data intake -------------------------------------------------------------
    def _add(self, pred: torch.Tensor, label: torch.Tensor, mask: torch.Tensor | None = None):
        pred = pred.detach().reshape(-1).float()
        label = label.detach().reshape(-1).float()
        if mask is not None:
            mask = mask.detach().reshape(-1).float()
        if pred.is_cuda and not self.float_label:
            if self._dev_table is None:
                self._dev_table = torch.zeros(2, self.bucket_size, dtype=torch.float64, device=pred.device)
                self._dev_sums = torch.zeros(5, dtype=torch.float64, device=pred.device)
            N.check(N.cuda().pbx_auc_bucket(N.ptr(pred.contiguous()), N.ptr(label.contiguous()),
                                            N.ptr(mask.contiguous()) if mask is not None else None,
                                            C.c_int64(pred.numel()), self.bucket_size, N.ptr(self._dev_table),
                                            N.ptr(self._dev_sums), N.stream_ptr()), "pbx_auc_bucket")
        elif pred.is_cuda:
            # float-label modes: accumulated on the device with a few PyTorch ops (no host sync per batch, so
            # a captured training step keeps replaying); merged into the host calculator at read-out
            if self._dev_table is None:
                self._dev_table = torch.zeros(2, self.bucket_size, dtype=torch.float64, device=pred.device)
                self._dev_sums = torch.zeros(5, dtype=torch.float64, device=pred.device)
            accumulate_float_labels(self._dev_table, self._dev_sums, pred, label, mask, int(self.float_label))
        else:
            self.calc.add(pred.cpu().numpy(), label.cpu().numpy(), None if mask is None else mask.cpu().numpy(),
                          self.float_label)

    def add_data(self, tensors: dict):
        self._add(tensors[self.pred_varname], tensors[self.label_varname])

    # ---- read-out ------------------------------------------------------------------
    def _flush_device(self):
        if self._dev_table is not None:
            self.calc.merge_tables(self._dev_table.cpu().numpy(), self._dev_sums.cpu().numpy())
            self._dev_table.zero_()
            self._dev_sums.zero_()

    def _reduce(self, group):
        if group is None:
            return
        import torch.distributed as dist
        t, s = self.calc.export_state()
        buf = torch.from_numpy(np.concatenate([t.reshape(-1), s]))
        if dist.get_backend(group) == "nccl":
            buf = buf.cuda()
        dist.all_reduce(buf, group=group)
        buf = buf.cpu().numpy()
        self.calc.import_state(buf[:2 * self.bucket_size].reshape(2, -1), buf[2 * self.bucket_size:])

    def get_metric_msg(self, group=None) -> list:
        self._flush_device()
        self._reduce(group)
        out = self.calc.compute()
        self.calc.reset()
        from .flags import FLAGS
        if FLAGS.enable_debug_print_metrics_info:          # verbose AUC computation (metrics.cc:29-31)
            names = ("auc", "bucket_error", "mae", "rmse", "actual_ctr", "predicted_ctr", "copc", "size")
            print(f"[metric {self.name}] " + " ".join(f"{n}={v:.6g}" for n, v in zip(names, out)), flush=True)
        return out

    def get_nan_inf_msg(self) -> list:
        self._flush_device()
        return list(self.calc.nan_inf())


class MaskMetricMsg(MetricMsg):
    """Only rows whose mask is non-zero count (MaskAucCalculator)."""

    def __init__(self, mask_varname="mask", **kw):
        super().__init__(**kw)
        self.mask_varname = mask_varname

    def add_data(self, tensors):
        self._add(tensors[self.pred_varname], tensors[self.label_varname], tensors[self.mask_varname])


class MultiMaskMetricMsg(MetricMsg):
    """Rows where every listed mask variable equals its listed value (MultiMaskAucCalculator)."""

    def __init__(self, mask_varname="", mask_varvalue_list="", **kw):
        super().__init__(**kw)
        self.mask_names = mask_varname.split()
        self.mask_values = [int(v) for v in str(mask_varvalue_list).split()]
        assert len(self.mask_names) == len(self.mask_values)

    def add_data(self, tensors):
        keep = None
        for n, v in zip(self.mask_names, self.mask_values):
            m = tensors[n].reshape(-1) == v
            keep = m if keep is None else keep & m
        self._add(tensors[self.pred_varname], tensors[self.label_varname], keep.float())


class CmatchRankMetricMsg(MetricMsg):
    """Rows whose (cmatch, rank) is in ``cmatch_rank_group`` ("c_r c_r ..."; with
    ``ignore_rank`` only cmatch is compared) (CmatchRankAucCalculator)."""

    def __init__(self, cmatch_rank_group="", cmatch_rank_varname="cmatch_rank", ignore_rank=False, **kw):
        super().__init__(**kw)
        self.cmatch_rank_varname = cmatch_rank_varname
        self.ignore_rank = ignore_rank
        self.pairs = []
        for tok in cmatch_rank_group.split():
            if ignore_rank:
                self.pairs.append((int(tok), 0))
            else:
                c, r = tok.split("_")
                self.pairs.append((int(c), int(r)))

    def _select(self, tensors):
        cm, rk = parse_cmatch_rank(tensors[self.cmatch_rank_varname].reshape(-1))
        keep = torch.zeros_like(cm, dtype=torch.bool)
        for c, r in self.pairs:
            keep |= (cm == c) if self.ignore_rank else ((cm == c) & (rk == r))
        return keep

    def add_data(self, tensors):
        self._add(tensors[self.pred_varname], tensors[self.label_varname], self._select(tensors).float())


class CmatchRankMaskMetricMsg(CmatchRankMetricMsg):
    def __init__(self, mask_varname="mask", **kw):
        super().__init__(**kw)
        self.mask_varname = mask_varname

    def add_data(self, tensors):
        keep = self._select(tensors) & (tensors[self.mask_varname].reshape(-1) != 0)
        self._add(tensors[self.pred_varname], tensors[self.label_varname], keep.float())


class MultiTaskMetricMsg(MetricMsg):
    """One prediction variable per (cmatch, rank) pair; a row feeds the prediction whose pair
    it matches (MultiTaskAucCalculator, box_wrapper.cc:372-455)."""

    def __init__(self, cmatch_rank_group="", cmatch_rank_varname="cmatch_rank", **kw):
        super().__init__(**kw)
        self.cmatch_rank_varname = cmatch_rank_varname
        self.pairs = [tuple(int(v) for v in tok.split("_")) for tok in cmatch_rank_group.split()]
        self.pred_names = self.pred_varname.split()
        assert len(self.pairs) == len(self.pred_names), "cmatch_rank group and pred list sizes differ"

    def add_data(self, tensors):
        cm, rk = parse_cmatch_rank(tensors[self.cmatch_rank_varname].reshape(-1))
        label = tensors[self.label_varname]
        for (c, r), pn in zip(self.pairs, self.pred_names):
            self._add(tensors[pn], label, ((cm == c) & (rk == r)).float())


class FloatMaskMetricMsg(MaskMetricMsg):
    """Float labels in [0, 1] split fractionally between the two tables (FloatMaskAucCalculator)."""

    def __init__(self, **kw):
        super().__init__(float_label=True, **kw)


class ContinueMaskMetricMsg(MaskMetricMsg):
    """Regression-style metric on masked rows: only mae / rmse / mean prediction / mean label
    are meaningful (ContinueMaskCalculator).  Values are unbounded: nothing is clamped to [0, 1]
    (add_unlock_data_with_continue_label, fleet/metrics.cc:98-105)."""

    def __init__(self, **kw):
        super().__init__(float_label=2, **kw)

    def get_metric_msg(self, group=None):
        self._flush_device()
        self._reduce(group)
        t, s = self.calc.export_state()
        n = max(s[4], 1.0)
        self.calc.reset()
        # [mae, rmse, predicted mean, actual mean, size]
        return [s[0] / n, float(np.sqrt(s[1] / n)), s[2] / n, s[3] / n, s[4]]


class WuAucMetricMsg(MetricMsg):
    """Per-user AUC: uauc (mean over users) and wuauc (instance-weighted)
    (metrics.cc:501-587)."""

    def __init__(self, uid_varname="uid", **kw):
        super().__init__(**kw)
        self.uid_varname = uid_varname

    FLUSH_EVERY = 256      # batches kept on the device before one copy to the host

    def add_data(self, tensors):
        p, y, u = tensors[self.pred_varname].detach(), tensors[self.label_varname].detach(), tensors[self.uid_varname].detach()
        if p.is_cuda:
            # per-user records stay on the device; one host copy per FLUSH_EVERY batches instead of one per batch
            buf = self.__dict__.setdefault("_pending", [])
            buf.append((p.reshape(-1).float().clone(), y.reshape(-1).float().clone(), u.reshape(-1).clone()))
            if len(buf) >= self.FLUSH_EVERY:
                self._flush_uid()
            return
        self.calc.add_uid(p.float().cpu().numpy(), y.float().cpu().numpy(), u.cpu().numpy())

    def _flush_uid(self):
        buf = self.__dict__.get("_pending") or []
        if buf:
            p = torch.cat([b[0] for b in buf]).cpu().numpy()
            y = torch.cat([b[1] for b in buf]).cpu().numpy()
            u = torch.cat([b[2] for b in buf]).cpu().numpy()
            buf.clear()
            self.calc.add_uid(p, y, u)

    def get_metric_msg(self, group=None):
        """``[uauc, wuauc, user_cnt, size]``.  Over several ranks the per-user AUCs are computed where the
        records are (a user seen on two ranks counts as two users there, as in the reference) and the SUMS are
        reduced: uauc = sum(auc) / users, wuauc = sum(auc * ins) / size (``GetWuAucMetricMsg``, metrics.h:774-805)."""
        self._flush_uid()
        uauc, wuauc, users, size = self.calc.compute_wuauc()
        self.calc.reset()
        if group is not None:
            import torch.distributed as dist
            t = torch.tensor([users, size, uauc * users, wuauc * size], dtype=torch.float64)
            if dist.get_backend(group) == "nccl":
                t = t.cuda()
            dist.all_reduce(t, group=group)
            users, size, su, sw = t.tolist()
            uauc, wuauc = su / (users + 1e-10), sw / (size + 1e-10)
        return [uauc, wuauc, users, size]


class NanInfMetricMsg(MetricMsg):
    """Counts NaN / Inf predictions (NanInfCalculator, box_wrapper.cc:853-886)."""

    def __init__(self, **kw):
        super().__init__(**kw)
        self.nan = 0
        self.inf = 0
        self.total = 0

    def add_data(self, tensors):
        p = tensors[self.pred_varname].detach().reshape(-1)
        self.total += p.numel()
        if p.is_cuda:                       # counted on the device, read at get_nan_inf_msg: no sync per batch
            if self.__dict__.get("_dev_cnt") is None:
                self._dev_cnt = torch.zeros(2, dtype=torch.int64, device=p.device)
            self._dev_cnt += torch.stack([torch.isnan(p).sum(), torch.isinf(p).sum()])
            return
        self.nan += int(torch.isnan(p).sum())
        self.inf += int(torch.isinf(p).sum())

    def get_nan_inf_msg(self):
        if self.__dict__.get("_dev_cnt") is not None:
            c = self._dev_cnt.cpu().tolist()
            self.nan += int(c[0])
            self.inf += int(c[1])
            self._dev_cnt.zero_()
        out = [self.nan, self.inf, self.total]
        self.nan = self.inf = self.total = 0
        return out

    def get_metric_msg(self, group=None):
        return self.get_nan_inf_msg()


_REGISTRY = {
    "AucCalculator": MetricMsg,
    "MultiTaskAucCalculator": MultiTaskMetricMsg,
    "CmatchRankAucCalculator": CmatchRankMetricMsg,
    "MaskAucCalculator": MaskMetricMsg,
    "MultiMaskAucCalculator": MultiMaskMetricMsg,
    "CmatchRankMaskAucCalculator": CmatchRankMaskMetricMsg,
    "FloatMaskAucCalculator": FloatMaskMetricMsg,
    "ContinueMaskCalculator": ContinueMaskMetricMsg,
    "WuAucCalculator": WuAucMetricMsg,
    "NanInfCalculator": NanInfMetricMsg,
}


def create_metric(method: str, **kw) -> MetricMsg:
    if method not in _REGISTRY:
        raise ValueError(f"unsupported metric {method!r}; supported: {sorted(_REGISTRY)}")
    cls = _REGISTRY[method]
    kw = {k: v for k, v in kw.items() if v not in (None,)}
    # drop empty optional names so that subclass defaults apply
    for opt in ("mask_varname", "cmatch_rank_varname", "cmatch_rank_group"):
        if kw.get(opt, None) == "":
            kw.pop(opt)
    return cls(**kw)


class Metric:
    """``core.Metric()`` — the metric registry of the PSLIB / PSGPU path (fleet/metrics.h:640-820,
    pybind/metrics_py.cc): the same calculators as ``BoxWrapper.init_metric``, for jobs that run without a
    ``BoxWrapper``.  One instance per process; the trainers feed it next to the wrapper's metrics
    (``add_metric_data``), results are reduced over the ``torch.distributed`` world."""
    _instance = None

    def __new__(cls):
        if cls._instance is None:
            cls._instance = super().__new__(cls)
            cls._instance._metrics = {}
            cls._instance._phase = 1
        return cls._instance

    @classmethod
    def get_instance(cls):
        return cls._instance

    @classmethod
    def reset_instance(cls) -> None:
        cls._instance = None

    def init_metric(self, method: str, name: str, label_varname: str = "label", pred_varname: str = "pred",
                    cmatch_rank_varname: str = "", mask_varname: str = "", uid_varname: str = "", metric_phase: int = 1,
                    cmatch_rank_group: str = "", ignore_rank: bool = False, bucket_size: int = 1000000) -> None:
        kw = dict(name=name, label_varname=label_varname, pred_varname=pred_varname, cmatch_rank_varname=cmatch_rank_varname,
                  mask_varname=mask_varname, metric_phase=metric_phase, cmatch_rank_group=cmatch_rank_group,
                  ignore_rank=ignore_rank, bucket_size=bucket_size)
        if uid_varname:
            kw["uid_varname"] = uid_varname
        self._metrics[name] = create_metric(method, **kw)

    def flip_phase(self) -> None:
        self._phase = (self._phase + 1) % 2

    def phase(self) -> int:
        return self._phase

    def get_metric_name_list(self, metric_phase: int = -1) -> list:
        return [n for n, m in self._metrics.items() if metric_phase < 0 or m.metric_phase == metric_phase]

    def add_metric_data(self, tensors: dict) -> None:
        for m in self._metrics.values():
            if m.metric_phase == self._phase:
                m.add_data(tensors)

    @staticmethod
    def _group():
        import torch.distributed as dist
        return dist.group.WORLD if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None

    def _get(self, name: str):
        if name not in self._metrics:
            raise ValueError("The metric name you provided is not registered.")
        return self._metrics[name]

    def get_metric_msg(self, name: str) -> list:
        """[auc, bucket_error, mae, rmse, actual_ctr, predicted_ctr, copc, size]; resets the metric."""
        return self._get(name).get_metric_msg(self._group())

    def get_wuauc_metric_msg(self, name: str) -> list:
        """[user_cnt, size, sum of user AUCs, sum of (user AUC * instances), uauc, wuauc]; resets the metric."""
        uauc, wuauc, users, size = self._get(name).get_metric_msg(self._group())
        return [users, size, uauc * users, wuauc * size, uauc, wuauc]


class Auc:
    """``fluid.layers.auc(input, label, curve='ROC', num_thresholds=2**12 - 1, topk=1, slide_steps=1)`` as an object
    (layers/metric_op.py:100-240, operators/metrics/auc_op.h): bucket counts of positives / negatives by
    ``floor(p * num_thresholds)``, accumulated over the calls.  ``__call__`` returns what the layer returns —
    ``(auc over everything seen, auc over the last slide_steps batches, [batch_stat_pos, batch_stat_neg, stat_pos,
    stat_neg])`` — and the last two are the statistics ``fleet.metrics.auc(stat_pos, stat_neg)`` reduces over the
    workers.  ``slide_steps`` 0 makes the batch AUC cover everything as well."""

    def __init__(self, curve: str = "ROC", num_thresholds: int = 2 ** 12 - 1, topk: int = 1, slide_steps: int = 1):
        if curve != "ROC":
            raise ValueError("only the ROC curve is computed from bucket statistics")
        self.num_thresholds, self.slide_steps = int(num_thresholds), int(slide_steps)
        n = self.num_thresholds + 1
        self.stat_pos = torch.zeros(1, n, dtype=torch.int64)
        self.stat_neg = torch.zeros(1, n, dtype=torch.int64)
        self._window: list = []

    @staticmethod
    def area(stat_pos: torch.Tensor, stat_neg: torch.Tensor) -> float:
        pos_b, neg_b = stat_pos.reshape(-1).flip(0).double(), stat_neg.reshape(-1).flip(0).double()
        pos_run = pos_b.cumsum(0)
        area = float((neg_b * (2 * pos_run - pos_b) / 2).sum())
        pos, neg = float(pos_run[-1]), float(neg_b.sum())
        return 0.5 if pos * neg == 0 else area / (pos * neg)

    def __call__(self, input: torch.Tensor, label: torch.Tensor):
        p = input.detach()
        p = (p[:, 1] if p.dim() == 2 and p.shape[1] == 2 else p.reshape(-1)).float().cpu()
        y = label.detach().reshape(-1).cpu() > 0
        b = (p * self.num_thresholds).long().clamp_(0, self.num_thresholds)
        n = self.num_thresholds + 1
        bp, bn = torch.bincount(b[y], minlength=n).view(1, n), torch.bincount(b[~y], minlength=n).view(1, n)
        self.stat_pos += bp
        self.stat_neg += bn
        if self.slide_steps > 0:
            self._window.append((bp, bn))
            del self._window[:-self.slide_steps]
            wp, wn = sum(w[0] for w in self._window), sum(w[1] for w in self._window)
        else:
            wp, wn = self.stat_pos, self.stat_neg
        return self.area(self.stat_pos, self.stat_neg), self.area(wp, wn), [wp, wn, self.stat_pos, self.stat_neg]

    def reset(self) -> None:
        self.stat_pos.zero_()
        self.stat_neg.zero_()
        self._window.clear()


_AUC_LAYERS: dict = {}


def auc(input: torch.Tensor, label: torch.Tensor, curve: str = "ROC", num_thresholds: int = 2 ** 12 - 1, topk: int = 1,
        slide_steps: int = 1, name: str = "auc"):
    """Function form of :class:`Auc`; the accumulated statistics live under ``name`` (one per metric of the model)."""
    key = (name, num_thresholds, slide_steps)
    if key not in _AUC_LAYERS:
        _AUC_LAYERS[key] = Auc(curve, num_thresholds, topk, slide_steps)
    return _AUC_LAYERS[key](input, label)
