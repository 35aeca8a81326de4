"""``pull_box_sparse`` / ``push_box_sparse`` and the fused embedding+pool op.

Reference: ``fluid.layers._pull_box_sparse`` (python/paddle/fluid/layers/nn.py:793) and the
operator pair (operators/pull_box_sparse_op.h:92-211): N id tensors -> N embedding tensors
``[n_i, size]`` with columns ``[show, clk, embed_w, embedx...]``; the grad op *is* the sparse
optimizer step (no dense gradient flows out).  ``_pull_box_extended_sparse``
(contrib/layers/nn.py:1678; operators/pull_box_extended_sparse_op.h:25-193) adds a second
"expand" embedding.

Two entry points:

* ``pull_box_sparse(batch)``                — API parity: materialises the [N, size] embeddings
  (all slots in one tensor, ``split_by_slot`` gives the reference's per-slot list); backward
  merges per-occurrence gradients and pushes.
* ``sparse_embedding_pool(batch, cvm, ...)`` — the fast path every model here uses: pull and
  ``fused_seqpool_cvm`` fused, pooled output read straight from the de-duplicated value
  buffer; backward is ONE kernel doing seqpool/CVM backward + merge-by-key + all-to-all +
  optimizer (csrc/cuda/sparse_a2a.cu ``k_sparse_push`` mode 0).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from .. import _native as N
from ..box_wrapper import BoxWrapper
from . import reference as R
from .seqpool import SeqpoolAttrs, seqpool_cvm_forward, seqpool_cvm_backward


@dataclass
class SlotBatch:
    """One mini-batch of sparse ids in the slot-major layout all sparse ops consume."""
    keys: torch.Tensor            # int64 [N], ordered by (slot, instance)
    lod: torch.Tensor             # int32 [S*B + 1] bag offsets, bag = slot * B + ins
    batch_size: int
    num_slots: int
    dense: torch.Tensor | None = None     # fp32 [B, dense_dim] (label + dense features)
    n_keys: int | None = None             # host-known N (avoids a sync); None -> keys.numel()
    ins_ids: list | None = None
    meta: dict | None = None              # search_id / rank / cmatch tensors
    n_dev: torch.Tensor | None = None     # int32 device scalar with the real key count (static-shape
                                          # batches replayed from a CUDA graph); usually lod[S*B:]

    @property
    def N(self) -> int:
        return self.keys.numel() if self.n_keys is None else self.n_keys

    def key2slot(self) -> torch.Tensor:
        n = self.N
        out = torch.empty(max(n, 1), dtype=torch.int32, device=self.keys.device)
        if self.keys.is_cuda:
            N.check(N.cuda().pbx_key2slot(N.ptr(self.lod), self.batch_size, self.num_slots, C.c_int64(n),
                                          N.ptr(out), N.stream_ptr()))
        else:
            ends = self.lod[self.batch_size::self.batch_size].long()
            out[:n] = torch.bucketize(torch.arange(n), ends, right=True).int()
        return out[:n]

    def slot_lod(self, slot: int) -> torch.Tensor:
        """int64 [B + 1] offsets of one slot's bags inside that slot's rows (``split_by_slot(emb, batch)[slot]``):
        the ``lod`` the ragged ops of ``ops/sequence.py`` take."""
        B = self.batch_size
        seg = self.lod[slot * B:(slot + 1) * B + 1].long()
        return seg - seg[0]

    def slot_ranges(self):
        """[(begin, end)] of every slot's rows in the slot-major key order (host sync)."""
        b = self.lod[::self.batch_size].tolist()
        return list(zip(b[:-1], b[1:]))


_ANCHORS: dict = {}


def _anchor(device) -> torch.Tensor:
    """A requires-grad scalar that ties sparse ops into the autograd graph so that their
    backward (the push) runs even though ids carry no gradient."""
    a = _ANCHORS.get(device)
    if a is None:
        a = _ANCHORS[device] = torch.zeros(1, device=device, requires_grad=True)
    return a


class _PullBoxSparse(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, box, batch, size, offset):
        keys = batch.keys
        vals, restore = box.pull_sparse(keys, batch.n_dev)
        n = batch.N
        hidden = box.hidden
        width = size - 0
        quant = box.feature_type == 3          # int16 pull records: embedx = int16 * scale (box_wrapper.cu:36-43)
        scale = 1.0 if quant else box.pull_embedx_scale
        if keys.is_cuda:
            out = torch.empty(max(n, 1), hidden - offset, dtype=torch.float32, device=keys.device)
            N.check(N.cuda().pbx_pull_scatter(N.ptr(vals), vals.stride(0), N.ptr(restore), C.c_int64(n), hidden,
                                              offset, C.c_float(scale), N.ptr(out), N.stream_ptr()))
            out = out[:n]
        else:
            out = vals[restore.long()][:, offset:hidden].clone()
            if scale != 1.0:
                out[:, max(3 - offset, 0):] *= scale
        if quant:
            ex = out[:, max(3 - offset, 0):]
            q = box.pull_embedx_scale
            ex.copy_(torch.round(ex / q).clamp_(-32768, 32767) * q)
        ctx.box, ctx.batch, ctx.restore, ctx.offset = box, batch, restore, offset
        return out

    @staticmethod
    def backward(ctx, dout):
        box, batch, offset = ctx.box, ctx.batch, ctx.offset
        if offset:
            dout = torch.nn.functional.pad(dout, (offset, 0))
        box.push_sparse_grad(batch.keys, ctx.restore, dout.contiguous(), batch.key2slot(), batch.batch_size)
        return None, None, None, None, None


def pull_box_sparse(batch: SlotBatch, size: int | None = None, offset: int = 0,
                    box: BoxWrapper | None = None) -> torch.Tensor:
    """Embeddings of every id of the batch, [N, size - offset] in slot-major order.

    ``size`` must equal ``cvm_offset + embedx_dim`` of the BoxWrapper (default hidden 11);
    ``offset`` skips leading cvm columns (the op's ``offset`` attr, box_wrapper.cu:92)."""
    box = box or BoxWrapper.get_instance()
    size = size or box.hidden
    if size != box.hidden:
        raise ValueError(f"pull_box_sparse size {size} != cvm_offset + embedx_dim = {box.hidden}")
    return _PullBoxSparse.apply(_anchor(batch.keys.device), box, batch, size, offset)


def split_by_slot(emb: torch.Tensor, batch: SlotBatch, pad_empty: bool = True):
    """The reference's output list: one [n_s, size] tensor per slot.  An empty slot yields one
    zero row when ``FLAGS_enable_pull_box_padding_zero`` (pull_box_sparse_op.h:131-147)."""
    from ..flags import FLAGS
    outs = []
    for b, e in batch.slot_ranges():
        if e == b and pad_empty and FLAGS.enable_pull_box_padding_zero:
            outs.append(torch.zeros(1, emb.shape[1], dtype=emb.dtype, device=emb.device))
        else:
            outs.append(emb[b:e])
    return outs


def pull_box_extended_sparse(batch: SlotBatch, size: int, extend_size: int, box: BoxWrapper | None = None,
                             mask=None, offset: int = 0, expand_only: bool = True):
    """``_pull_box_extended_sparse`` (contrib/layers/nn.py:1678-1750; op pull_box_extended_sparse_op.h:25-193;
    kernels PullCopyNNCross / PullCopyNNCrossWithEmb, box_wrapper.cu:146-268).

    The table row's embedx holds ``embedx_dim = (size - 3) + extend_size`` values: the first part is the base
    embedding, the tail the expand embedding.  Returns ``(out, out_extend)``:

    * ``out``  ``[N, size - offset]`` = ``[cvm columns (minus the first ``offset``), base embedx]``;
    * ``out_extend``: ``expand_only=True`` (layer default) -> ``[N, extend_size]`` the expand embedding alone;
      ``expand_only=False`` -> ``[N, size - offset + extend_size]`` = base row followed by the expand embedding
      (the "WithEmb" kernel).
    * ``mask`` (one int per slot; bit 0 = this slot has an ``Out``, bit 1 = it has an ``OutExtend``): the results
      become LISTS of per-slot tensors holding only the selected slots, in slot order, exactly like the op's
      output lists; a slot without a bit gets no tensor and contributes no gradient for that part.

    One pull serves both outputs, and the backward of both flows into the one push (gradient columns that no
    output selected are zero)."""
    box = box or BoxWrapper.get_instance()
    if size - 3 + extend_size != box.embedx_dim:
        raise ValueError("size - 3 + extend_size must equal the table's embedx_dim")
    if not 0 <= offset <= 3:
        raise ValueError("offset skips leading cvm columns: 0..3")
    full = _PullBoxSparse.apply(_anchor(batch.keys.device), box, batch, box.hidden, 0)
    base = full[:, offset:size]
    ext = full[:, size:] if expand_only else torch.cat([base, full[:, size:]], 1)
    if mask is None or len(mask) == 0:
        return base, ext
    if len(mask) != batch.num_slots:
        raise ValueError(f"mask has {len(mask)} entries for {batch.num_slots} slots")
    outs, outs_ext = [], []
    for flag, (b0, e0) in zip(mask, batch.slot_ranges()):
        if flag & 0x01:
            outs.append(base[b0:e0])
        if flag & 0x02:
            outs_ext.append(ext[b0:e0])
    return outs, outs_ext


class _SparseEmbeddingPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, box, batch, cvm, attrs, dense):
        B, S = batch.batch_size, batch.num_slots
        vals, restore = box.pull_sparse(batch.keys, batch.n_dev)
        sw = S * attrs.out_width(box.hidden)
        if dense is not None and vals.is_cuda:
            # pooled embeddings and dense features land in ONE row-major buffer: no concat pass
            out = torch.empty(B, sw + dense.shape[1], dtype=torch.float32, device=vals.device)
            if dense.stride(1) == 1 and dense.dtype == torch.float32:
                seqpool_cvm_forward(vals, batch.lod, B, S, attrs, box.hidden, restore=restore, out=out, tail=dense)
            else:
                seqpool_cvm_forward(vals, batch.lod, B, S, attrs, box.hidden, restore=restore, out=out)
                out[:, sw:] = dense
        else:
            out = seqpool_cvm_forward(vals, batch.lod, B, S, attrs, box.hidden, restore=restore)
            if dense is not None:
                out = torch.cat([out, dense], 1)
        ctx.box, ctx.batch, ctx.attrs, ctx.restore, ctx.sw = box, batch, attrs, restore, sw
        ctx.save_for_backward(cvm)
        return out

    @staticmethod
    def backward(ctx, dout):
        box, batch, attrs = ctx.box, ctx.batch, ctx.attrs
        (cvm,) = ctx.saved_tensors
        B, S = batch.batch_size, batch.num_slots
        nothing = (None, None, None, None, None, None)    # the anchor only keeps the node alive
        if box._test_mode:
            return nothing
        from ..flags import FLAGS
        fused_ok = (dout.is_cuda and box.router is None and attrs.cvm_offset == 2 and attrs.embedx_concate_size == 1
                    and attrs.variant == 0 and not FLAGS.pbx_deterministic_merge)
        if fused_ok:
            Eo = attrs.out_width(box.hidden)
            if dout.stride(1) != 1:
                dout = dout.contiguous()
            # the pooled gradient is read in place from the (wider) input gradient: row stride = ld
            box.engine.push_pooled(ctx.restore, dout, cvm.contiguous(), batch.lod, B, S, Eo, 2,
                                   attrs.dim_off(), box.push_grad_scale(B), ld_dout=dout.stride(0))
        else:
            dsp = dout[:, :ctx.sw].contiguous()
            dx = seqpool_cvm_backward(dsp, cvm, batch.lod, B, S, batch.N, attrs, box.hidden)
            box.push_sparse_grad(batch.keys, ctx.restore, dx, batch.key2slot(), B)
        return nothing


def sparse_embedding_pool(batch: SlotBatch, cvm: torch.Tensor, box: BoxWrapper | None = None,
                          attrs: SeqpoolAttrs | None = None, dense: torch.Tensor | None = None,
                          **kw) -> torch.Tensor:
    """pull_box_sparse + fused_seqpool_cvm in one differentiable op -> [B, S*Eo].

    ``cvm`` [B, 2] holds (show = 1, clk = label) per instance; in backward it becomes the
    show/click count columns of the push.  With ``dense`` [B, Dd] the result is
    ``[B, S*Eo + Dd]`` — the tower input — written in one buffer (no concat); ``dense`` gets no
    gradient (it is input data)."""
    box = box or BoxWrapper.get_instance()
    attrs = attrs or SeqpoolAttrs(**kw)
    return _SparseEmbeddingPool.apply(_anchor(batch.keys.device), box, batch, cvm, attrs, dense)
