"""Tree-based deep model (TDM) sampling over an item tree.

Reference role: ``Dataset.tdm_sample`` (pybind data_set_py.cc; framework/data_set.cc ``TDMSample``
over the ``index_dataset`` tree) and the ``tdm_sampler`` / ``tdm_child`` operators: for every
instance's item id, every tree layer contributes the item's ancestor on that layer as a positive
sample plus ``neg`` nodes of the same layer drawn uniformly without the positive.

``TDMTree`` here is a complete ``branch``-ary tree laid out by heap index (node code c has children
``c * branch + 1 ..``); items sit on the last layer in the order given.  Node id = code + 1 (0 is the
padding id), so node ids are directly usable as feasigns of a sparse slot.  Everything is tensor
ops: it runs on whichever device the ids live on.
"""
from __future__ import annotations

from typing import Sequence

import torch


class TDMTree:
    def __init__(self, item_ids: Sequence[int] | torch.Tensor, branch: int = 2):
        items = torch.as_tensor(item_ids, dtype=torch.int64)
        self.branch = int(branch)
        self.n_items = items.numel()
        height = 1
        while self.branch ** (height - 1) < self.n_items:
            height += 1
        self.height = height                                   # layers 0 (root) .. height-1 (leaves)
        self.layer_start = [(self.branch ** l - 1) // (self.branch - 1) for l in range(height + 1)]
        order = torch.argsort(items)
        self._sorted_items = items[order]
        self._leaf_of_sorted = order                           # position among the leaves
        # nodes of a layer that have at least one item below them
        self.layer_size = [min(self.branch ** l, -(-self.n_items // self.branch ** (height - 1 - l)))
                           for l in range(height)]

    def to(self, device) -> "TDMTree":
        self._sorted_items = self._sorted_items.to(device)
        self._leaf_of_sorted = self._leaf_of_sorted.to(device)
        return self

    def leaf_position(self, item_ids: torch.Tensor):
        idx = torch.searchsorted(self._sorted_items, item_ids).clamp_(max=self.n_items - 1)
        known = self._sorted_items[idx] == item_ids
        return self._leaf_of_sorted[idx], known

    def travel(self, item_ids: torch.Tensor) -> torch.Tensor:
        """[n, height] node ids of every item's path root -> leaf (0 for unknown items)."""
        pos, known = self.leaf_position(item_ids)
        cols = []
        for l in range(self.height):
            in_layer = torch.div(pos, self.branch ** (self.height - 1 - l), rounding_mode="floor")
            cols.append(torch.where(known, in_layer + self.layer_start[l] + 1, torch.zeros_like(pos)))
        return torch.stack(cols, 1)

    def children(self, node_ids: torch.Tensor):
        """(child ids [n, branch], is_leaf-layer mask) — the ``tdm_child`` op."""
        code = node_ids - 1
        kids = code[:, None] * self.branch + 1 + torch.arange(self.branch, device=node_ids.device)[None, :]
        layer_of_kid = torch.bucketize(kids, torch.tensor(self.layer_start[1:], device=node_ids.device), right=True)
        ok = (node_ids[:, None] > 0) & (layer_of_kid < self.height)
        in_layer = kids - torch.tensor(self.layer_start, device=node_ids.device)[layer_of_kid.clamp(max=self.height)]
        size = torch.tensor(self.layer_size + [0], device=node_ids.device)[layer_of_kid.clamp(max=self.height)]
        ok = ok & (in_layer < size)
        return torch.where(ok, kids + 1, torch.zeros_like(kids)), ok & (layer_of_kid == self.height - 1)


def tdm_sample(tree: TDMTree, item_ids: torch.Tensor, neg_per_layer: Sequence[int], start_layer: int = 1,
               generator: torch.Generator | None = None, output_positive: bool = True):
    """``(nodes, labels, mask)``, each ``[n, sum_l (pos + neg_l)]`` for layers ``start_layer ..
    height-1``: per layer the ancestor (label 1) then ``neg_per_layer[l - start_layer]`` other nodes
    of the layer (label 0), distinct from the positive.  ``mask`` is 0 for unknown items and for
    negatives a layer is too small to supply."""
    path = tree.travel(item_ids)
    n, dev = item_ids.numel(), item_ids.device
    nodes, labels, mask = [], [], []
    for li, l in enumerate(range(start_layer, tree.height)):
        neg = int(neg_per_layer[li])
        pos = path[:, l]
        known = pos > 0
        if output_positive:
            nodes.append(pos[:, None])
            labels.append(torch.ones(n, 1, dtype=torch.int64, device=dev))
            mask.append(known[:, None].long())
        if neg == 0:
            continue
        size = tree.layer_size[l]
        avail = max(size - 1, 0)
        # draw from the layer minus the positive: r in [0, size-1), shifted past the positive's slot
        r = (torch.rand(n, neg, device=dev, generator=generator) * max(avail, 1)).long().clamp_(max=max(avail - 1, 0))
        pos_in = (pos - 1 - tree.layer_start[l]).clamp(min=0)
        r = r + (r >= pos_in[:, None]).long()
        ok = known[:, None] & (torch.arange(neg, device=dev)[None, :] < avail).expand(n, neg) & (avail > 0)
        nodes.append(torch.where(ok, r + tree.layer_start[l] + 1, torch.zeros_like(r)))
        labels.append(torch.zeros(n, neg, dtype=torch.int64, device=dev))
        mask.append(ok.long())
    return torch.cat(nodes, 1), torch.cat(labels, 1), torch.cat(mask, 1)


def tdm_child(x: torch.Tensor, tree: TDMTree):
    """``fluid.contrib.layers.tdm_child`` (contrib/layers/nn.py): children of every node id in ``x``
    and the mask of those that are items (leaves).  Shapes ``x.shape + [branch]``."""
    kids, leaf = tree.children(x.reshape(-1))
    return kids.view(*x.shape, tree.branch), leaf.view(*x.shape, tree.branch).long()


def tdm_sampler(x: torch.Tensor, neg_samples_num_list: Sequence[int], tree: TDMTree, output_positive: bool = True,
                output_list: bool = False, start_layer: int = 1, generator: torch.Generator | None = None):
    """``fluid.contrib.layers.tdm_sampler``: per tree layer the positive node of every item in ``x``
    plus ``neg_samples_num_list[layer]`` negatives.  Returns ``(out, labels, mask)``; with
    ``output_list`` each is a list with one tensor per layer."""
    nodes, labels, mask = tdm_sample(tree, x.reshape(-1), neg_samples_num_list, start_layer, generator, output_positive)
    if not output_list:
        return nodes, labels, mask
    widths = [int(n) + (1 if output_positive else 0) for n in neg_samples_num_list[:tree.height - start_layer]]
    return tuple(list(torch.split(t, widths, dim=1)) for t in (nodes, labels, mask))
