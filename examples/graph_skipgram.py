#!/usr/bin/env python
"""Node embeddings from random walks (DeepWalk-style) on the sparse table.

    python examples/graph_skipgram.py [--cpu]

Builds a small community graph, keeps its CSR on the device (``graph.DeviceGraph``), generates
walk pairs + negatives (``GraphDataGenerator``) and trains ``models.graph.SkipGram`` through the
stock pull / push path — the flow of the reference's GPU graph engine + ``GraphDataGenerator``.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from paddlebox_b200.box_wrapper import BoxWrapper             # noqa: E402
from paddlebox_b200.graph import GraphDataGenerator, GraphTable   # noqa: E402
from paddlebox_b200.models.graph import SkipGram               # noqa: E402
from paddlebox_b200.trainer import CTRTrainer, TrainerConfig   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--communities", type=int, default=4)
    ap.add_argument("--size", type=int, default=16)
    ap.add_argument("--epochs", type=int, default=8)
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()
    device = torch.device("cuda", 0) if torch.cuda.is_available() and not a.cpu else torch.device("cpu")
    rs = np.random.RandomState(0)
    src, dst = [], []
    for c in range(a.communities):
        base = 1000 * (c + 1)
        for i in range(a.size):
            for j in rs.choice(a.size, 5, replace=False):
                if i != j:
                    src += [base + i, base + j]
                    dst += [base + j, base + i]
        nxt = 1000 * ((c + 1) % a.communities + 1)              # one bridge to the next community
        src += [base, nxt]
        dst += [nxt, base]
    g = GraphTable()
    g.add_edges(np.array(src, np.uint64), np.array(dst, np.uint64))
    g.build()
    print(f"graph: {g.num_nodes()} nodes, {g.num_edges()} edges")
    box = BoxWrapper.set_instance(embedx_dim=8, device=device, rank=0, world_size=1)
    box.initialize_gpu_and_load_model({"mf_create_thresholds": 0.0, "mf_initial_range": 0.3, "mf_learning_rate": 0.2},
                                      table_capacity=1 << 14, max_keys_per_batch=1 << 14)
    gen = GraphDataGenerator(g, walk_len=8, window=2, batch_size=512, neg=2, walks_per_node=4, device=device)
    box.begin_feed_pass("20240101")
    box.feed_pass(gen.all_keys())
    box.end_feed_pass()
    box.begin_pass()
    trainer = CTRTrainer(SkipGram(8).to(device), box, TrainerConfig(dense_lr=1e-2))
    for ep in range(a.epochs):
        losses = [float(trainer.train_step(b).detach()) for b in gen.batches()]
        print(f"epoch {ep}: {len(losses)} batches, loss {np.mean(losses):.4f}")
    box.end_pass()
    BoxWrapper.reset_instance()


if __name__ == "__main__":
    main()
