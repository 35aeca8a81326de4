"""AUC runner: feature-importance evaluation by slot shuffling.

Reference: ``BoxWrapper::InitializeAucRunner`` / ``RecordReplace`` / ``RecordReplaceBack`` /
``FlipPhase`` (fleet/box_wrapper.h:908-1011, fleet/box_wrapper.cc:212-368) and the feed-side
replacers (data_feed.h:1772-1962).  The reference evaluates a group of slots by replacing
their feasigns in every record with feasigns drawn from a reservoir of other records, running
an inference pass, and reading the metric: the AUC drop is the importance of the group.

Here the donor of record i is record perm(i) of the same pass (a seeded permutation in the
native dataset, ``csrc/host/dataset.cc``), which is undone exactly afterwards, so a dataset
can be evaluated for many groups in a row without reloading.  The model runs in test mode
(no pushes, no embedx creation), like the reference's auc-runner mode (``Mode() == 1``).
"""
from __future__ import annotations

from typing import Sequence

import torch

from .box_wrapper import BoxWrapper


class AucRunner:
    """``runner = AucRunner(box, [["slot_3"], ["slot_7", "slot_8"]], metric="auc")``
    ``runner.run(trainer, dataset)`` -> {"base": auc, "groups": [{"slots":..., "auc":..., "drop":...}]}"""

    def __init__(self, box: BoxWrapper | None = None, slot_eval: Sequence[Sequence[str]] = (), metric: str = "auc",
                 thread_num: int = 1, pool_size: int = 0, slot_list: Sequence[str] = ()):
        self.box = box or BoxWrapper.get_instance()
        self.slot_eval = [list(g) for g in slot_eval]
        self.metric = metric
        self.thread_num, self.pool_size = thread_num, pool_size
        known = set(slot_list)
        if known:
            for g in self.slot_eval:
                missing = [s for s in g if s not in known]
                if missing:
                    raise ValueError(f"auc runner: slots {missing} are not in slot_list")
        self.mode = 1                   # reference Mode(): 1 = auc runner
        self.round_id = 0

    def eval_slot_set(self) -> set:
        return {s for g in self.slot_eval for s in g}

    # -- one inference pass over the dataset, returns the metric's AUC ---------------------------
    def _infer_pass(self, trainer, dataset) -> float:
        self.box.get_metric_msg(self.metric)          # reading a metric resets it
        if dataset._prepared_workers == 0:
            dataset.prepare_train(trainer.world, trainer.rank)
        with torch.no_grad():
            for batch in dataset.batches(worker=trainer.rank):
                trainer.infer_step(batch)
        if self.box.use_cuda:
            torch.cuda.synchronize()
        return float(self.box.get_metric_msg(self.metric)[0])

    def run(self, trainer, dataset) -> dict:
        """Evaluate every slot group on the (already loaded, pass-begun) dataset."""
        was_test = getattr(self.box, "_test_mode", False)
        old_phase = self.box.phase()
        mphase = self.box.get_metric(self.metric).metric_phase
        self.box.set_test_mode(True)
        self.box.set_phase(mphase)
        try:
            base = self._infer_pass(trainer, dataset)
            out = {"base": base, "groups": []}
            for g in self.slot_eval:
                dataset.slots_shuffle(g)                 # RecordReplace
                try:
                    auc = self._infer_pass(trainer, dataset)
                finally:
                    dataset.slots_shuffle_back(g)        # RecordReplaceBack
                self.round_id += 1
                out["groups"].append({"slots": g, "auc": auc, "drop": base - auc})
            return out
        finally:
            self.box.set_test_mode(was_test)
            self.box.set_phase(old_phase)
