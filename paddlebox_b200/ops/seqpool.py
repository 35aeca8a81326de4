"""fused_seqpool_cvm family: pool variable-length bags of pulled embeddings per
(slot, instance), optional filters / quantisation / embedx-concat, CVM transform —
one kernel forward, one backward (csrc/cuda/seqpool.cu).

Public layers mirror ``fluid.contrib.layers.fused_seqpool_cvm*``
(python/paddle/fluid/contrib/layers/nn.py:1750-2160); kernels replace
operators/fused/fused_seqpool_cvm*_op.cu.

Tensors: ``x`` [N, hidden] value rows in slot-major bag order (or the engine's unique
value buffer + ``restore`` index), ``lod`` int32 [S*B+1]; output [B, S*Eo] — slot s is the
column block [s*Eo, (s+1)*Eo), i.e. the per-slot outputs of the reference already
concatenated.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch

from .. import _native as N
from . import reference as R


@dataclass
class SeqpoolAttrs:
    use_cvm: bool = True
    cvm_offset: int = 2
    pad_value: float = 0.0
    need_filter: bool = False
    show_coeff: float = 0.2
    clk_coeff: float = 1.0
    threshold: float = 0.96
    quant_ratio: int = 0
    clk_filter: bool = False
    embed_threshold_filter: bool = False
    embed_threshold: float = 0.0
    embed_thres_size: int = 0
    embedx_concate_size: int = 1
    embedx_concate_filter: bool = False
    variant: int = 0          # 0 base, 1 with_conv, 2 with_credit
    show_filter: bool = False
    slot_threshold: torch.Tensor | None = None   # with_diff_thres
    trade_w: torch.Tensor | None = None          # tradew

    def __post_init__(self):
        if self.quant_ratio == 0 and self.need_filter:
            self.quant_ratio = 128   # nn.py:1802-1804

    def out_width(self, hidden: int) -> int:
        if self.use_cvm:
            drop = self.clk_filter if self.variant == 0 else self.show_filter
            e = hidden - (1 if drop else 0)
        else:
            e = hidden - self.cvm_offset
        return e * self.embedx_concate_size

    def dim_off(self) -> int:
        if self.use_cvm:
            return 1 if (self.clk_filter if self.variant == 0 else self.show_filter) else 0
        return self.cvm_offset

    def ref_kwargs(self):
        return dict(use_cvm=self.use_cvm, cvm_offset=self.cvm_offset, pad_value=self.pad_value,
                    need_filter=self.need_filter, show_coeff=self.show_coeff, clk_coeff=self.clk_coeff,
                    threshold=self.threshold, quant_ratio=self.quant_ratio, clk_filter=self.clk_filter,
                    embed_threshold_filter=self.embed_threshold_filter, embed_threshold=self.embed_threshold,
                    embed_thres_size=self.embed_thres_size, embedx_concate_size=self.embedx_concate_size,
                    embedx_concate_filter=self.embedx_concate_filter, variant=self.variant,
                    show_filter=self.show_filter, slot_threshold=self.slot_threshold, trade_w=self.trade_w)


def _params(a: SeqpoolAttrs, B, S, hidden, x_stride, out_stride) -> N.SeqpoolParams:
    p = N.SeqpoolParams()
    p.B, p.S, p.hidden, p.x_stride, p.out_stride = B, S, hidden, x_stride, out_stride
    p.Eo = a.out_width(hidden)
    p.use_cvm, p.cvm_offset, p.clk_filter = int(a.use_cvm), a.cvm_offset, int(a.clk_filter)
    p.need_filter, p.quant_ratio = int(a.need_filter), a.quant_ratio
    p.embed_threshold_filter, p.embed_thres_size = int(a.embed_threshold_filter), a.embed_thres_size
    p.concat, p.concat_filter = a.embedx_concate_size, int(a.embedx_concate_filter)
    p.variant, p.show_filter = a.variant, int(a.show_filter)
    p.pad_value, p.show_coeff, p.clk_coeff = a.pad_value, a.show_coeff, a.clk_coeff
    p.threshold, p.embed_threshold = a.threshold, a.embed_threshold
    p.slot_threshold = a.slot_threshold.data_ptr() if a.slot_threshold is not None else None
    p.trade_w = a.trade_w.data_ptr() if a.trade_w is not None else None
    return p


def seqpool_cvm_forward(x: torch.Tensor, lod: torch.Tensor, B: int, S: int, attrs: SeqpoolAttrs, hidden: int,
                        restore: torch.Tensor | None = None, out: torch.Tensor | None = None,
                        tail: torch.Tensor | None = None) -> torch.Tensor:
    """Raw forward.  ``x`` may be wider than ``hidden`` (row stride = x.shape[1]).  ``tail``
    [B, Dd] (GPU, with ``out`` wide enough) is copied behind the pooled columns by the same launch."""
    Eo = attrs.out_width(hidden)
    if not x.is_cuda:
        rows = x[restore.long()] if restore is not None else x
        kw = attrs.ref_kwargs()
        fwd = R.seqpool_cvm_fwd_vec if R.seqpool_plain(kw) else R.seqpool_cvm_fwd
        return fwd(rows[:, :hidden], lod, B, S, **kw)
    if out is None:
        out = torch.empty(B, S * Eo, dtype=torch.float32, device=x.device)
    p = _params(attrs, B, S, hidden, x.stride(0), out.stride(0))
    if tail is not None:
        assert tail.stride(1) == 1 and tail.dtype == torch.float32 and out.shape[1] >= S * Eo + tail.shape[1]
        N.check(N.cuda().pbx_seqpool_cvm_fwd_tail(C.byref(p), N.ptr(x), N.ptr(restore), N.ptr(lod), N.ptr(out),
                                                  N.ptr(tail), tail.stride(0), tail.shape[1], S * Eo,
                                                  N.stream_ptr()), "pbx_seqpool_cvm_fwd_tail")
        return out
    N.check(N.cuda().pbx_seqpool_cvm_fwd(C.byref(p), N.ptr(x), N.ptr(restore), N.ptr(lod), N.ptr(out),
                                         N.stream_ptr()), "pbx_seqpool_cvm_fwd")
    return out


def seqpool_cvm_backward(dout: torch.Tensor, cvm: torch.Tensor, lod: torch.Tensor, B: int, S: int, n_keys: int,
                         attrs: SeqpoolAttrs, hidden: int) -> torch.Tensor:
    if not dout.is_cuda:
        bwd = R.seqpool_cvm_bwd_vec if attrs.embedx_concate_size == 1 else R.seqpool_cvm_bwd
        return bwd(dout, cvm, lod, B, S, hidden, use_cvm=attrs.use_cvm, cvm_offset=attrs.cvm_offset,
                   clk_filter=attrs.clk_filter, embedx_concate_size=attrs.embedx_concate_size,
                   variant=attrs.variant, show_filter=attrs.show_filter)
    dout = dout.contiguous()
    dx = torch.empty(max(n_keys, 1), hidden, dtype=torch.float32, device=dout.device)
    p = _params(attrs, B, S, hidden, hidden, dout.stride(0))
    N.check(N.cuda().pbx_seqpool_cvm_bwd(C.byref(p), N.ptr(dout), N.ptr(cvm.contiguous()), N.ptr(lod), N.ptr(dx),
                                         hidden, N.stream_ptr()), "pbx_seqpool_cvm_bwd")
    return dx[:n_keys]


class _SeqpoolCVM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cvm, lod, B, S, attrs):
        hidden = x.shape[1]
        ctx.save_for_backward(cvm, lod)
        ctx.meta = (B, S, x.shape[0], attrs, hidden)
        return seqpool_cvm_forward(x.contiguous(), lod, B, S, attrs, hidden)

    @staticmethod
    def backward(ctx, dout):
        cvm, lod = ctx.saved_tensors
        B, S, n, attrs, hidden = ctx.meta
        dx = seqpool_cvm_backward(dout, cvm, lod, B, S, n, attrs, hidden)
        return dx, None, None, None, None, None


def fused_seqpool_cvm(x: torch.Tensor, cvm: torch.Tensor, lod: torch.Tensor, batch_size: int, num_slots: int,
                      pool_type: str = "SUM", **attrs) -> torch.Tensor:
    """Differentiable fused_seqpool_cvm over the slot-major embedding tensor ``x`` [N, hidden].

    The gradient w.r.t. ``x`` carries the show/click labels of ``cvm`` in its first
    ``cvm_offset`` columns — that is how counts reach ``push_box_sparse``.
    Returns [B, S*Eo]; use ``split_slots`` for the reference's list-of-tensors view."""
    if pool_type.upper() != "SUM":
        raise ValueError("fused_seqpool_cvm only supports SUM pooling (nn.py:1783-1786)")
    a = attrs.pop("attrs", None) or SeqpoolAttrs(**attrs)
    return _SeqpoolCVM.apply(x, cvm, lod, batch_size, num_slots, a)


def fused_seqpool_cvm_with_conv(x, cvm, lod, batch_size, num_slots, pool_type="SUM", show_filter=False,
                                cvm_offset=3, **kw):
    """cvm columns (show, clk, conv) -> log(show+1), log(clk+1), log(conv+1)-log(clk+1)
    (fused_seqpool_cvm_with_conv_op.cu:35-636)."""
    return fused_seqpool_cvm(x, cvm, lod, batch_size, num_slots, pool_type, variant=1, show_filter=show_filter,
                             cvm_offset=cvm_offset, **kw)


def fused_seqpool_cvm_with_credit(x, cvm, lod, batch_size, num_slots, pool_type="SUM", show_filter=False,
                                  cvm_offset=4, **kw):
    """every cvm column becomes log(x+1) (fused_seqpool_cvm_with_credit_op.cu:34-232)."""
    return fused_seqpool_cvm(x, cvm, lod, batch_size, num_slots, pool_type, variant=2, show_filter=show_filter,
                             cvm_offset=cvm_offset, **kw)


def fused_seqpool_cvm_with_diff_thres(x, cvm, lod, batch_size, num_slots, threshold_vec, pool_type="SUM", **kw):
    """per-slot filter threshold (fused_seqpool_cvm_with_diff_thres_op.cu:35-403)."""
    tv = torch.as_tensor(threshold_vec, dtype=torch.float32, device=x.device)
    return fused_seqpool_cvm(x, cvm, lod, batch_size, num_slots, pool_type, slot_threshold=tv, need_filter=True,
                             **kw)


def fused_seqpool_cvm_tradew(x, cvm, lod, batch_size, num_slots, trade_w, pool_type="SUM", **kw):
    """per-key trade weights scale the embedding columns before pooling
    (fused_seqpool_cvm_tradew_op.cu:34-295)."""
    return fused_seqpool_cvm(x, cvm, lod, batch_size, num_slots, pool_type, trade_w=trade_w.contiguous(), **kw)


def split_slots(out: torch.Tensor, num_slots: int):
    """[B, S*Eo] -> list of S views [B, Eo] (the reference op's output list)."""
    Eo = out.shape[1] // num_slots
    return [out[:, s * Eo:(s + 1) * Eo] for s in range(num_slots)]
