"""Stage timers and device-side step timing.

The reference brackets every stage with host wall-clock ``platform::Timer`` pairs and a
``cudaStreamSynchronize`` (fleet/box_wrapper.h:394-419, box_wrapper.cc:1085-1138), which
serialises the pipeline.  Here host timers are only used for host stages (feed pass,
begin/end pass); device stages are timed with CUDA events on the launching stream, which
costs no synchronisation until the numbers are read.
"""
from __future__ import annotations

import time
from collections import defaultdict

import torch


class StageTimers:
    def __init__(self):
        self.reset()

    def reset(self):
        self.total = defaultdict(float)
        self.count = defaultdict(int)

    def add(self, name: str, seconds: float):
        self.total[name] += seconds
        self.count[name] += 1

    class _Span:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            self.t0 = time.perf_counter()
            return self

        def __exit__(self, *a):
            self.owner.add(self.name, time.perf_counter() - self.t0)

    def span(self, name: str):
        return self._Span(self, name)

    def report(self) -> str:
        return " ".join(f"{k}={v * 1e3:.2f}ms/{self.count[k]}" for k, v in sorted(self.total.items()))


class CudaStageTimers:
    """Device-timed stages: ``with t.span('pull'):`` records two events on the current
    stream; ``summary()`` synchronises once and returns mean milliseconds per stage."""

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self.events = defaultdict(list)

    class _Span:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            if self.owner.enabled:
                self.a = torch.cuda.Event(enable_timing=True)
                self.b = torch.cuda.Event(enable_timing=True)
                self.a.record()
            return self

        def __exit__(self, *a):
            if self.owner.enabled:
                self.b.record()
                self.owner.events[self.name].append((self.a, self.b))

    def span(self, name: str):
        return self._Span(self, name)

    def summary(self) -> dict:
        if not self.enabled:
            return {}
        torch.cuda.synchronize()
        out = {k: sum(a.elapsed_time(b) for a, b in v) / max(len(v), 1) for k, v in self.events.items()}
        self.events.clear()
        return out
