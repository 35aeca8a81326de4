"""Named integer/float counters (reference ``platform/monitor.h`` STAT registry,
platform/monitor.h:35-147; e.g. STAT_total_feasign_num_in_mem, data_feed.cc:37)."""
from __future__ import annotations

import threading

_LOCK = threading.Lock()
_STATS: dict = {}


def stat_add(name: str, value=1):
    with _LOCK:
        _STATS[name] = _STATS.get(name, 0) + value


def stat_set(name: str, value):
    with _LOCK:
        _STATS[name] = value


def stat_get(name: str, default=0):
    with _LOCK:
        return _STATS.get(name, default)


def stat_reset(name: str | None = None):
    with _LOCK:
        if name is None:
            _STATS.clear()
        else:
            _STATS.pop(name, None)


def stat_all() -> dict:
    with _LOCK:
        return dict(_STATS)
