"""Graph-embedding model over the sparse table (DeepWalk / metapath2vec skip-gram).

The reference trains node embeddings by feeding ``GraphDataGenerator`` walk pairs through
``pull_gpups_sparse`` and a user-built dot-product head (data_feed.h:892-975); this is that head.
Input batches come from ``paddlebox_b200.graph.GraphDataGenerator``: slot 0 = centre node,
slot 1 = context node, ``dense[:, 0]`` = 1 for a walk pair, 0 for a sampled negative.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..box_wrapper import BoxWrapper
from ..ops.pull_box_sparse import SlotBatch, sparse_embedding_pool
from ..ops.seqpool import SeqpoolAttrs


class SkipGram(nn.Module):
    def __init__(self, embedx_dim: int = 16):
        super().__init__()
        self.embedx_dim = embedx_dim
        self.attrs = SeqpoolAttrs(use_cvm=False, cvm_offset=2)      # columns: embed_w, embedx
        self.bias = nn.Parameter(torch.zeros(1))

    def forward(self, batch: SlotBatch):
        label = batch.dense[:, 0]
        cvm = torch.stack([torch.ones_like(label), label], 1)
        E = 1 + self.embedx_dim
        x = sparse_embedding_pool(batch, cvm, BoxWrapper.get_instance(), self.attrs)      # [B, 2 * E]
        centre, context = x[:, 1:E], x[:, E + 1:2 * E]
        logit = (centre * context).sum(1) + x[:, 0] + x[:, E] + self.bias
        loss = nn.functional.binary_cross_entropy_with_logits(logit, label)
        return loss, {"pred": torch.sigmoid(logit), "label": label}
