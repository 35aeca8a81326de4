"""Job-wide metrics from per-worker statistics: ``fleet.metrics.sum / max / min / auc / mae / rmse / mse / acc``.

Reference: ``python/paddle/distributed/fleet/metrics/metric.py`` (:24-420).  A job script keeps local statistics
(the positive / negative bucket counts of an AUC layer, summed absolute / squared errors, instance counts) and
asks for the value over all workers; the statistics are all-reduced over the fleet world (``fleet.util``) and the
metric is computed from the sums.  Inputs are numpy arrays, torch tensors or numbers (the reference's "variable
name + scope" form has no counterpart without a static graph).
"""
from __future__ import annotations

import builtins
import math

import numpy as np

__all__ = ["sum", "max", "min", "auc", "mae", "rmse", "mse", "acc"]


def _np(v) -> np.ndarray:
    if hasattr(v, "detach"):
        v = v.detach().cpu().numpy()
    return np.asarray(v, dtype=np.float64)


def _reduce(v, mode: str, util=None) -> np.ndarray:
    if util is None:
        from . import fleet
        util = fleet.util
    a = _np(v)
    return np.asarray(util.all_reduce(a.reshape(-1), mode), dtype=np.float64).reshape(a.shape)


def sum(input, scope=None, util=None):          # noqa: A001  (the reference's names)
    return _reduce(input, "sum", util)


def max(input, scope=None, util=None):          # noqa: A001
    return _reduce(input, "max", util)


def min(input, scope=None, util=None):          # noqa: A001
    return _reduce(input, "min", util)


def auc(stat_pos, stat_neg, scope=None, util=None) -> float:
    """AUC over all workers from per-threshold-bucket counts of positives / negatives (``[1, buckets]`` or
    ``[buckets]``, bucket = floor(pred * buckets)): trapezoids from the highest bucket down (metric.py:144-224)."""
    pos_b = _reduce(stat_pos, "sum", util).reshape(-1)[::-1]
    neg_b = _reduce(stat_neg, "sum", util).reshape(-1)[::-1]
    pos_run = np.cumsum(pos_b)
    pos_prev = pos_run - pos_b
    area = float(np.sum(neg_b * (pos_prev + pos_run) / 2.0))
    pos, neg = (float(pos_run[-1]), float(neg_b.sum())) if pos_b.size else (0.0, 0.0)
    return 0.5 if pos * neg == 0 else area / (pos * neg)


def _ratio(num, den, util) -> float:
    n = float(_reduce(num, "sum", util).reshape(-1)[0])
    d = float(_reduce(den, "sum", util).reshape(-1)[0])
    return n / d


def mae(abserr, total_ins_num, scope=None, util=None) -> float:
    return _ratio(abserr, total_ins_num, util)


def mse(sqrerr, total_ins_num, scope=None, util=None) -> float:
    return _ratio(sqrerr, total_ins_num, util)


def rmse(sqrerr, total_ins_num, scope=None, util=None) -> float:
    return math.sqrt(_ratio(sqrerr, total_ins_num, util))


def acc(correct, total, scope=None, util=None) -> float:
    return _ratio(correct, total, util)


del builtins
