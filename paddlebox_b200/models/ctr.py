"""CTR model families on top of the sparse ops and the tcgen05 towers.

The reference ships no model code (the user script builds a static Program from
``_pull_box_sparse`` + ``fused_seqpool_cvm`` + fc towers); BASELINE.json names
the families the framework must run: CTR-DNN, Wide&Deep, DeepFM, MMoE.  All of them share
the sparse front end:

    ids --pull_box_sparse--> [show, clk, embed_w, embedx] --fused_seqpool_cvm--> [B, S*(3+D)]

with per-slot columns ``[log(show+1), log(clk+1)-log(show+1), embed_w, embedx_0..D-1]``.

Batch convention: ``batch.dense`` is ``[B, n_labels + dense_dim]`` — label column(s) first.
``forward`` returns ``(loss, {"pred": ..., "label": ...})``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..box_wrapper import BoxWrapper
from ..ops import functional as F
from ..ops.mlp import FusedMLP, GroupedMLP
from ..ops.pull_box_sparse import SlotBatch, sparse_embedding_pool
from ..ops.seqpool import SeqpoolAttrs


class CTRModelBase(nn.Module):
    n_labels = 1
    # every dense op of these models consumes the pooled embeddings, so the push (= backward of the
    # pull, the first op) runs after all dense gradients are final: the trainer may fold the dense
    # allreduce into the push kernel.  A subclass that adds a dense-only branch must set this False.
    dense_grads_final_before_push = True

    def __init__(self, num_slots: int, embedx_dim: int, dense_dim: int, use_data_norm: bool = False,
                 seqpool_attrs: SeqpoolAttrs | None = None):
        super().__init__()
        self.num_slots, self.embedx_dim, self.dense_dim = num_slots, embedx_dim, dense_dim
        self.slot_width = 3 + embedx_dim
        self.attrs = seqpool_attrs or SeqpoolAttrs(use_cvm=True, cvm_offset=2)
        self.sparse_width = num_slots * self.attrs.out_width(self.slot_width)
        self.input_width = self.sparse_width + dense_dim
        self.data_norm = F.DataNorm(self.input_width) if use_data_norm else None

    # ---- shared front end --------------------------------------------------------------
    def split_dense(self, batch: SlotBatch):
        d = batch.dense
        labels = d[:, :self.n_labels]
        feats = d[:, self.n_labels:self.n_labels + self.dense_dim]
        return labels, feats

    def sparse_features(self, batch: SlotBatch, click: torch.Tensor, dense: torch.Tensor | None = None):
        """[B, S*Eo (+ Dd)] pooled slot embeddings (and dense features appended in the same
        buffer); ``click`` [B] becomes the clk column of the push."""
        cvm = torch.stack([torch.ones_like(click), click], 1)
        return sparse_embedding_pool(batch, cvm, BoxWrapper.get_instance(), self.attrs, dense=dense)

    def features(self, batch: SlotBatch):
        """(x, pooled, labels): x [B, input_width] tower input, pooled = its sparse columns (a view)."""
        labels, feats = self.split_dense(batch)
        x = self.sparse_features(batch, labels[:, 0].contiguous(), feats if self.dense_dim else None)
        pooled = x[:, :self.sparse_width]
        if self.data_norm is not None:
            x = self.data_norm(x)
        return x, pooled, labels

    def dense_towers(self):
        return [m for m in self.modules() if isinstance(m, FusedMLP)]

    def mark_weights_dirty(self):
        for m in self.dense_towers():
            m.mark_weights_dirty()

    def named_dense_params(self):
        """(checkpoint name, parameter) pairs with the reference's naming."""
        seen = set()
        for m in self.dense_towers():
            for n, p in m.named_dense_params():
                seen.add(id(p))
                yield n, p
        for n, p in self.named_parameters():
            if id(p) not in seen:
                yield n.replace(".", "_"), p
        if self.data_norm is not None:
            for suffix in ("batch_size", "batch_sum", "batch_square_sum"):
                yield f"{self.data_norm.name}.{suffix}", getattr(self.data_norm, suffix)


class CtrDnn(CTRModelBase):
    """Plain DNN over pooled slot embeddings + dense features (BASELINE config 1)."""

    def __init__(self, num_slots=26, embedx_dim=8, dense_dim=13, hidden=(512, 256, 128), **kw):
        super().__init__(num_slots, embedx_dim, dense_dim, **kw)
        self.tower = FusedMLP([self.input_width, *hidden, 1], name="fc")

    def forward(self, batch: SlotBatch):
        x, _, labels = self.features(batch)
        loss, pred = F.ctr_head(self.tower(x), labels[:, 0], tower_mod=self.tower)
        return loss, {"pred": pred, "label": labels[:, 0]}


class WideDeep(CTRModelBase):
    """Wide (first-order weights of every slot + linear dense) & Deep (BASELINE config 2)."""

    def __init__(self, num_slots=26, embedx_dim=8, dense_dim=13, hidden=(512, 256, 128), **kw):
        super().__init__(num_slots, embedx_dim, dense_dim, **kw)
        self.tower = FusedMLP([self.input_width, *hidden, 1], name="deep_fc")
        self.wide_dense = nn.Parameter(torch.zeros(max(dense_dim, 1)))
        self.wide_bias = nn.Parameter(torch.zeros(1))

    def forward(self, batch: SlotBatch):
        x, pooled, labels = self.features(batch)
        E = self.attrs.out_width(self.slot_width)
        wide = pooled[:, 2::E].sum(1)                                   # sum of embed_w over slots
        if self.dense_dim:
            wide = wide + x[:, self.sparse_width:] @ self.wide_dense
        loss, pred = F.ctr_head(self.tower(x), labels[:, 0], extra=wide, bias=self.wide_bias, tower_mod=self.tower)
        return loss, {"pred": pred, "label": labels[:, 0]}


class DeepFM(CTRModelBase):
    """FM first + second order over slot embeddings, plus a deep tower (BASELINE config 3)."""

    def __init__(self, num_slots=26, embedx_dim=16, dense_dim=13, hidden=(512, 256, 128), **kw):
        super().__init__(num_slots, embedx_dim, dense_dim, **kw)
        self.tower = FusedMLP([self.input_width, *hidden, 1], name="fc")
        self.bias = nn.Parameter(torch.zeros(1))

    def forward(self, batch: SlotBatch):
        x, pooled, labels = self.features(batch)
        E = self.attrs.out_width(self.slot_width)
        if self.data_norm is None:
            # x's leading columns are the pooled slots: tower and FM term share one autograd node
            deep, fm = self.tower.forward_with_fm(x, self.num_slots, E, 2, 3, self.embedx_dim)
        else:
            fm = F.fm_interaction(pooled, self.num_slots, E, 2, 3, self.embedx_dim)
            deep = self.tower(x)
        loss, pred = F.ctr_head(deep, labels[:, 0], extra=fm, bias=self.bias, tower_mod=self.tower)
        return loss, {"pred": pred, "label": labels[:, 0]}


class MMoE(CTRModelBase):
    """Multi-gate mixture of experts for multi-task CTR/CVR (BASELINE config 4): shared
    expert towers (grouped tcgen05 GEMMs), one softmax gate and one tower per task."""

    def __init__(self, num_slots=300, embedx_dim=64, dense_dim=0, num_experts=8, expert_hidden=(1024, 512),
                 tower_hidden=(256, 128), num_tasks=2, **kw):
        self.n_labels = num_tasks
        super().__init__(num_slots, embedx_dim, dense_dim, **kw)
        self.num_experts, self.num_tasks = num_experts, num_tasks
        # the gate logits of every task ride on the experts' wide first layer: ONE bf16 copy of
        # x, one forward GEMM, one input-gradient GEMM for experts and gates together
        self.experts = GroupedMLP(num_experts, [self.input_width, *expert_hidden], final_relu=True, name="expert",
                                  extra_linear=num_tasks * num_experts)
        self.towers = nn.ModuleList([FusedMLP([expert_hidden[-1], *tower_hidden, 1], name=f"tower{t}")
                                     for t in range(num_tasks)])

    def forward(self, batch: SlotBatch):
        x, _, labels = self.features(batch)
        ex, gate_logits = self.experts(x)                      # [E, B, H], [B, T*E]
        gate_logits = gate_logits.view(-1, self.num_tasks, self.num_experts)
        losses, preds = [], []
        for t in range(self.num_tasks):
            g = torch.softmax(gate_logits[:, t], dim=1)        # [B, E]
            mix = torch.einsum("be,ebh->bh", g, ex)
            l, p = F.ctr_head(self.towers[t](mix), labels[:, t], tower_mod=self.towers[t])
            losses.append(l)
            preds.append(p)
        loss = torch.stack(losses).sum()
        out = {"pred": preds[0], "label": labels[:, 0]}
        for t in range(self.num_tasks):
            out[f"pred_{t}"] = preds[t]
            out[f"label_{t}"] = labels[:, t]
        return loss, out


class MultiDimDnn(nn.Module):
    """DNN over slots of different embedding widths (``paddlebox_b200.multi_dim.MultiDimBox``:
    the reference's variable-dimension pull layout, box_wrapper.cu:270-322)."""

    def __init__(self, mbox, dense_dim: int = 0, hidden=(64, 32), seqpool_attrs: SeqpoolAttrs | None = None):
        super().__init__()
        self.mbox, self.dense_dim = mbox, dense_dim
        self.attrs = seqpool_attrs or SeqpoolAttrs(use_cvm=True, cvm_offset=2)
        self.sparse_width = sum(mbox.out_widths(self.attrs))
        self.tower = FusedMLP([self.sparse_width + dense_dim, *hidden, 1], name="fc")

    def mark_weights_dirty(self):
        self.tower.mark_weights_dirty()

    def forward(self, batch: SlotBatch):
        label = batch.dense[:, 0].contiguous()
        cvm = torch.stack([torch.ones_like(label), label], 1)
        x = self.mbox.embedding_pool(batch, cvm, self.attrs)
        if self.dense_dim:
            x = torch.cat([x, batch.dense[:, 1:1 + self.dense_dim]], 1)
        loss, pred = F.ctr_head(self.tower(x), label, tower_mod=self.tower)
        return loss, {"pred": pred, "label": label}


MODEL_REGISTRY = {"ctr_dnn": CtrDnn, "wide_deep": WideDeep, "deepfm": DeepFM, "mmoe": MMoE}


def build_model(name: str, **kw) -> CTRModelBase:
    return MODEL_REGISTRY[name](**kw)
