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
