"""Online prediction from the serving ("xbox") models.

The reference trains with BoxPS and exports ``xbox`` base / delta models (``SaveBase`` /
``SaveDelta``, fleet/box_wrapper.cc:1286-1315) that a separate serving system loads: feature values
without optimizer state, a base per day plus a delta per pass.  ``Predictor`` is that consumer for
models written here: it loads a base model and any number of deltas (in order; later rows replace
earlier ones), the dense parameters saved by ``checkpoint.save_persistables``, places the rows in
the device table once, and then only pulls — test mode: no key is created, nothing is updated.
"""
from __future__ import annotations

import os
from typing import Sequence

import numpy as np
import torch

from .box_wrapper import BoxWrapper


class Predictor:
    def __init__(self, model: torch.nn.Module, xbox_base: str, xbox_deltas: Sequence[str] = (),
                 dense_dir: str | None = None, embedx_dim: int = 8, feature_type: int = 0,
                 pull_embedx_scale: float = 1.0, device: torch.device | None = None,
                 table_capacity: int | None = None, max_keys_per_batch: int = 1 << 20):
        if BoxWrapper._instance is not None:
            raise RuntimeError("a BoxWrapper already exists in this process; call BoxWrapper.reset_instance() first")
        self.box = BoxWrapper.set_instance(embedx_dim=embedx_dim, feature_type=feature_type,
                                           pull_embedx_scale=pull_embedx_scale, device=device, rank=0, world_size=1)
        n_rows = self._count_rows([xbox_base, *xbox_deltas])
        cap = table_capacity or max(1 << 16, int(n_rows / 0.6))
        self.box.initialize_gpu_and_load_model({}, table_capacity=cap, max_keys_per_batch=max_keys_per_batch)
        self.rows_loaded = self.box.load_model(xbox_base, serving=True)
        for d in xbox_deltas:
            self.rows_loaded += self.box.load_model(d, serving=True)
        self.model = model.to(self.box.device)
        if dense_dir:
            from .checkpoint import load_persistables
            load_persistables(self.model, dense_dir, strict=False)
            fn = getattr(self.model, "mark_weights_dirty", None)
            if fn:
                fn()
        self.model.eval()
        # the whole serving model is the working set: stage it once, then only pull
        keys = self.box.host_table.keys()
        self.box.begin_feed_pass()
        self.box.feed_pass(keys)
        self.box.end_feed_pass()
        self.box.begin_pass()
        self.box.set_test_mode(True)

    @staticmethod
    def _count_rows(dirs) -> int:
        n = 0
        for d in dirs:
            for name in os.listdir(d):
                if name.startswith("part-"):
                    with open(os.path.join(d, name), "rb") as f:
                        n += sum(1 for _ in f)
        return n

    @property
    def num_features(self) -> int:
        return self.box.host_table.size()

    def predict(self, batch) -> torch.Tensor:
        """Predictions ``[B]`` for one ``SlotBatch`` (label column of ``batch.dense`` is ignored by the
        output; unknown feasigns pull zeros)."""
        with torch.no_grad():
            _, outs = self.model(batch)
        return outs["pred"]

    def predict_dataset(self, dataset) -> np.ndarray:
        if dataset._prepared_workers == 0:
            dataset.prepare_train(1, 0, shuffle=False)
        preds = [self.predict(b).float().cpu().numpy() for b in dataset.batches(worker=0)]
        return np.concatenate(preds) if preds else np.zeros(0, np.float32)

    def close(self) -> None:
        BoxWrapper.reset_instance()

    # ---- following a training job through its donefiles -------------------------------------------------------
    def apply_delta(self, xbox_delta: str) -> int:
        """Bring a newer delta model into the live predictor: rows of the delta replace / extend the loaded ones.
        On a GPU the device table is rebuilt from the host tier (the working set is the whole serving model)."""
        n = self.box.load_model(xbox_delta, serving=True)
        self.rows_loaded += n
        if self.box.use_cuda:
            self.box.set_test_mode(False)
            self.box.end_pass()
            self.box.begin_feed_pass()
            self.box.feed_pass(self.box.host_table.keys())
            self.box.end_feed_pass()
            self.box.begin_pass()
            self.box.set_test_mode(True)
        return n

    @classmethod
    def from_donefiles(cls, model: torch.nn.Module, output_path: str, **kw) -> "Predictor":
        """The newest serving model a training job has published under ``output_path`` in the reference's
        donefile convention (``FleetUtil.write_xbox_donefile`` / ``write_model_donefile``): the last base of
        ``xbox_base_done.txt``, every delta of ``xbox_patch_done.txt`` written after that base, and the dense
        parameters next to the last batch model of ``donefile.txt`` (``<day>/<pass>/dense``)."""
        base, deltas, dense = serving_models_from_donefiles(output_path)
        p = cls(model, base, deltas, dense_dir=dense, **kw)
        p._donefile_root, p._applied = output_path, {base, *deltas}
        return p

    def refresh_from_donefiles(self) -> int:
        """Apply the deltas published since the last look; returns how many.  (A NEW base means a new day: build a
        new predictor from it.)"""
        root = getattr(self, "_donefile_root", None)
        if root is None:
            raise RuntimeError("this predictor was not built with from_donefiles")
        _, deltas, _ = serving_models_from_donefiles(root)
        fresh = [d for d in deltas if d not in self._applied]
        for d in fresh:
            self.apply_delta(d)
            self._applied.add(d)
        return len(fresh)


def _local_path(input_field: str) -> str:
    """``hdfs://ns/out/20240102/delta-2/000`` -> ``/out/20240102/delta-2`` (the ``/000`` part suffix and a
    ``scheme://host`` prefix removed)."""
    import re
    p = re.sub(r"^[a-zA-Z][a-zA-Z0-9+.-]*://[^/]*", "", input_field)
    return p[:-4] if p.endswith("/000") else p


def serving_models_from_donefiles(output_path: str):
    """``(base dir, [delta dirs written after that base, oldest first], dense dir or None)`` from the donefiles
    of ``output_path`` (formats of ``fleet_util.py``).  A base is written at the END of its day, so the deltas
    that belong on top of it are those of later days."""
    import json
    root = output_path.rstrip("/")

    def lines(name):
        p = os.path.join(root, name)
        return [ln for ln in open(p).read().split("\n") if ln.strip()] if os.path.isfile(p) else []
    bases = [json.loads(ln) for ln in lines("xbox_base_done.txt")]
    if not bases:
        raise FileNotFoundError(f"no base model listed in {root}/xbox_base_done.txt")
    base = _local_path(bases[-1]["input"])
    base_day = int(base.rstrip("/").split("/")[-2])
    deltas = []
    for ln in lines("xbox_patch_done.txt"):
        d = _local_path(json.loads(ln)["input"])
        if int(d.rstrip("/").split("/")[-2]) > base_day:
            deltas.append(d)
    dense = None
    rows = [ln.split("\t") for ln in lines("donefile.txt")]
    for r in reversed(rows):
        cand = os.path.join(r[2], "dense") if len(r) > 2 else ""
        if cand and os.path.isdir(cand):
            dense = cand
            break
    return base, deltas, dense
