"""Pipeline-by-section: a model cut into consecutive sections, each on its own device with its
own worker threads, micro-batches flowing through bounded queues.

Reference role: ``BoxPSOptimizer(cut_list=..., place_list=..., concurrency_list=..., queue_size=...)``
splits the program at the cut variables (``_split_program``, python/paddle/fluid/optimizer.py:
7496-7575) for the stock ``PipelineTrainer`` / ``SectionWorker`` (framework/section_worker.cc);
BoxPSWorker itself never runs sections, so this is the API's behaviour rather than a
PaddleBox hot path.

Forward of micro-batch i in section s overlaps forward of micro-batch i+1 in section s-1 (one
thread pool per section, CUDA work of different sections on different devices / streams).  The
backward pass runs after the last micro-batch's forward, through autograd, which already crosses
devices; gradients accumulate over the micro-batches (GPipe schedule).
"""
from __future__ import annotations

import queue
import threading
from typing import Callable, Sequence

import torch
import torch.nn as nn


def split_sequential(model: nn.Sequential, cut_list: Sequence[str]) -> list:
    """Cut ``model`` AFTER each child named in ``cut_list`` -> list of ``nn.Sequential`` sections."""
    names = [n for n, _ in model.named_children()]
    missing = [c for c in cut_list if c not in names]
    if missing:
        raise ValueError(f"cut_list names {missing} are not children of the model ({names})")
    sections, cur = [], []
    for n, m in model.named_children():
        cur.append((n, m))
        if n in cut_list:
            sections.append(cur)
            cur = []
    if cur:
        sections.append(cur)
    from collections import OrderedDict
    return [nn.Sequential(OrderedDict(sec)) for sec in sections]


class SectionPipeline:
    def __init__(self, sections: Sequence[nn.Module], place_list: Sequence | None = None,
                 concurrency_list: Sequence[int] | None = None, queue_size: int = 30):
        self.sections = list(sections)
        n = len(self.sections)
        self.places = [torch.device(p) for p in (place_list or ["cpu"] * n)]
        self.concurrency = list(concurrency_list or [1] * n)
        assert len(self.places) == n and len(self.concurrency) == n
        for sec, dev in zip(self.sections, self.places):
            sec.to(dev)
        self.queue_size = queue_size

    def parameters(self):
        for s in self.sections:
            yield from s.parameters()

    def forward_backward(self, micro_batches: Sequence[torch.Tensor], loss_fn: Callable, targets: Sequence) -> float:
        """Runs every micro-batch through the sections (pipelined), then backward; returns the mean loss.
        Gradients accumulate into ``.grad`` (divide the learning rate or call with a mean-reducing loss)."""
        n_mb, n_sec = len(micro_batches), len(self.sections)
        qs = [queue.Queue(self.queue_size) for _ in range(n_sec + 1)]
        errors: list = []

        def worker(s: int):
            sec, dev = self.sections[s], self.places[s]
            while True:
                item = qs[s].get()
                if item is None:
                    qs[s].put(None)              # let the section's other threads see the end marker
                    return
                i, x = item
                try:
                    qs[s + 1].put((i, sec(x.to(dev, non_blocking=True))))
                except Exception as e:           # surface in the caller, keep the pipeline draining
                    errors.append(e)
                    qs[s + 1].put((i, None))

        threads = []
        for s in range(n_sec):
            for _ in range(max(1, self.concurrency[s])):
                t = threading.Thread(target=worker, args=(s,), daemon=True)
                t.start()
                threads.append((s, t))
        for i, mb in enumerate(micro_batches):
            qs[0].put((i, mb))
        outs: dict = {}
        while len(outs) < n_mb:
            i, y = qs[n_sec].get()
            outs[i] = y
        for s in range(n_sec):                  # stop section s once everything upstream is done
            qs[s].put(None)
        for _, t in threads:
            t.join()
        if errors:
            raise errors[0]
        total = 0.0
        for i in range(n_mb):
            loss = loss_fn(outs[i], targets[i].to(outs[i].device)) / n_mb
            loss.backward()
            total += float(loss.detach())
        return total
