#!/usr/bin/env python
"""Serve predictions from what examples/train_ctr.py published.

    python examples/train_ctr.py --cpu --out /tmp/job        # writes models + donefiles
    python examples/serve.py --job /tmp/job [--cpu]

Reads the donefiles (newest base model + every delta recorded after it), loads the serving ("xbox") models and
the dense parameters into a ``serving.Predictor`` and scores a few synthetic instances — the consumer side of
``save_base`` / ``save_delta``.
"""
from __future__ import annotations

import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from paddlebox_b200.models.ctr import build_model                    # noqa: E402
from paddlebox_b200.ops.pull_box_sparse import SlotBatch             # noqa: E402
from paddlebox_b200.serving import Predictor                         # noqa: E402
from paddlebox_b200.utils.donefile import ModelManifest              # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--job", required=True, help="the --out directory of examples/train_ctr.py")
    ap.add_argument("--model", default="deepfm", choices=["deepfm", "ctr_dnn", "wide_deep"])
    ap.add_argument("--slots", type=int, default=8)
    ap.add_argument("--dense", type=int, default=4)
    ap.add_argument("--dim", type=int, default=8)
    ap.add_argument("--key-space", type=int, default=200)
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()
    base, deltas = ModelManifest(os.path.join(a.job, "donefiles")).latest()
    if base is None:
        raise SystemExit(f"no base model recorded under {a.job}/donefiles")
    dense_dir = (deltas[-1] if deltas else base).get("dense") or base.get("dense")
    device = torch.device("cuda", 0) if torch.cuda.is_available() and not a.cpu else torch.device("cpu")
    model = build_model(a.model, num_slots=a.slots, embedx_dim=a.dim, dense_dim=a.dense, hidden=(64, 32))
    pred = Predictor(model, base["input"], [d["input"] for d in deltas], dense_dir=dense_dir, embedx_dim=a.dim,
                     device=device, table_capacity=1 << 16, max_keys_per_batch=1 << 12)
    print(f"loaded base {base['input']} + {len(deltas)} delta(s): {pred.num_features} features")
    B, S = 16, a.slots
    gen = torch.Generator().manual_seed(0)
    keys = torch.cat([torch.randint(1, a.key_space + 1, (B,), generator=gen) + s * a.key_space for s in range(S)])
    lod = torch.arange(S * B + 1, dtype=torch.int32)
    dense = torch.cat([torch.zeros(B, 1), torch.rand(B, a.dense, generator=gen)], 1)
    batch = SlotBatch(keys.to(device), lod.to(device), B, S, dense.to(device))
    p = pred.predict(batch).float().cpu()
    print("predictions:", " ".join(f"{v:.4f}" for v in p.tolist()))
    pred.close()


if __name__ == "__main__":
    main()
