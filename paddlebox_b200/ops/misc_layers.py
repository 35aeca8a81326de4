"""The remaining CTR-adjacent names of ``fluid.contrib.layers`` (python/paddle/fluid/contrib/layers/nn.py):
``fused_embedding_seq_pool`` (:520-583), ``sparse_embedding`` (:969-1097), ``fused_elemwise_activation`` (:57-107)
and the ``pow2_decay_with_linear_warmup`` schedule (:2822-2858).  Plain PyTorch: they are small, bandwidth-light
and not on the sparse-PS hot path (the pooled pull of the parameter server is ``sparse_embedding_pool``).
The vision / NLP ops of that module (``var_conv_2d``, ``tree_conv``, ``multiclass_nms2``, ``bilateral_slice``,
``correlation``, ``search_pyramid_hash``, ``match_matrix_tensor``, ``sequence_topk_avg_pooling``,
``fused_bn_add_act``) are stock-Paddle operators outside the CTR scope of SURVEY §2.5."""
from __future__ import annotations

import torch

_UNARY = {"relu": torch.relu, "tanh": torch.tanh, "sigmoid": torch.sigmoid, "gelu": torch.nn.functional.gelu}
_BINARY = {"elementwise_add": torch.add, "elementwise_mul": torch.mul, "elementwise_sub": torch.sub}


def fused_embedding_seq_pool(ids: torch.Tensor, lod: torch.Tensor, weight: torch.Tensor, combiner: str = "sum",
                             padding_idx: int | None = None) -> torch.Tensor:
    """Embedding lookup + pooling of every sequence in one op: ``ids`` [N] (all sequences back to back), ``lod``
    [B + 1] offsets, ``weight`` [vocab, D] -> [B, D].  Ids equal to ``padding_idx`` contribute zeros (a negative
    ``padding_idx`` counts from the end of the vocabulary, as in the reference)."""
    if combiner != "sum":
        raise ValueError("fused_embedding_seq_pool supports combiner='sum' (as the reference op does)")
    vocab = weight.shape[0]
    ids = ids.reshape(-1).long()
    rows = weight[ids.clamp(0, vocab - 1)]
    if padding_idx is not None:
        pad = padding_idx if padding_idx >= 0 else vocab + padding_idx
        rows = rows * (ids != pad).unsqueeze(1).to(rows.dtype)
    lod = lod.reshape(-1).long()
    B = lod.numel() - 1
    seg = torch.repeat_interleave(torch.arange(B, device=ids.device), lod[1:] - lod[:-1], output_size=ids.numel())
    return torch.zeros(B, weight.shape[1], dtype=rows.dtype, device=rows.device).index_add_(0, seg, rows)


def sparse_embedding(batch, size=None, padding_idx=None, is_test: bool = False, entry=None, table_class: str = "",
                     param_attr=None, dtype: str = "float32", slot=None, box=None) -> torch.Tensor:
    """Embeddings served by the parameter server (the reference's large-scale ``lookup_table`` on the pscore
    tables): per-id rows ``[N, dim]`` in slot-major order — ``pull_box_sparse`` without the cvm columns.  ``size``
    = ``[vocab, dim]`` or ``dim`` (the vocabulary is unbounded here); ``is_test`` pulls without creating rows;
    ids equal to ``padding_idx`` give zeros.  ``entry`` / ``table_class`` / ``param_attr`` configure the CPU tables
    of the reference and are accepted."""
    from ..box_wrapper import BoxWrapper
    from .pull_box_sparse import pull_box_sparse
    box = box or BoxWrapper.get_instance()
    dim = size[-1] if isinstance(size, (list, tuple)) else (size or box.embedx_dim)
    if dim != box.embedx_dim:
        raise ValueError(f"sparse_embedding dim {dim} != embedx_dim {box.embedx_dim} of the wrapper")
    prev = box._test_mode
    if is_test and not prev:
        box.set_test_mode(True)
    try:
        out = pull_box_sparse(batch, offset=box.hidden - box.embedx_dim, box=box)
    finally:
        if is_test and not prev:
            box.set_test_mode(False)
    if padding_idx is not None:
        n = out.shape[0]
        out = out * (batch.keys[:n] != padding_idx).unsqueeze(1).to(out.dtype)
    return out


def fused_elemwise_activation(x: torch.Tensor, y: torch.Tensor, functor_list, axis: int = -1, scale: float = 0.0,
                              save_intermediate_out: bool = True):
    """``functor_list = [binary, unary]`` -> ``binary(x, unary(y))``; ``[unary, binary]`` -> ``unary(binary(x, y))``
    (unary: relu / tanh / sigmoid / gelu / scale; binary: elementwise_add / _mul / _sub).  ``y`` broadcasts onto
    ``x`` starting at dimension ``axis`` (-1: trailing dimensions, Paddle's rule).  Returns the result, or
    ``(result, intermediate)`` with ``save_intermediate_out``."""
    f = [s.strip() for s in (functor_list.split(",") if isinstance(functor_list, str) else functor_list)]
    if len(f) != 2:
        raise ValueError("functor_list needs exactly two functors")

    def unary(name, t):
        if name == "scale":
            return t * scale
        if name not in _UNARY:
            raise ValueError(f"unknown unary functor {name!r}")
        return _UNARY[name](t)

    def bcast(t):
        if t.dim() == x.dim() or axis == -1:
            return t
        return t.reshape(t.shape + (1,) * (x.dim() - axis - t.dim()))

    if f[0] in _BINARY and f[1] not in _BINARY:
        mid = unary(f[1], y)
        out = _BINARY[f[0]](x, bcast(mid))
    elif f[1] in _BINARY and f[0] not in _BINARY:
        mid = _BINARY[f[1]](x, bcast(y))
        out = unary(f[0], mid)
    else:
        raise ValueError(f"functor_list {f} must hold one binary and one unary functor")
    return (out, mid) if save_intermediate_out else out


def pow2_decay_with_linear_warmup(warmup_steps: int, total_steps: int, base_lr: float, end_lr: float, step: int | None
                                  = None):
    """Learning rate at ``step`` (1-based): a linear ramp to ``base_lr`` over ``warmup_steps``, then
    ``(base_lr - end_lr) * (1 - progress)^2 + end_lr`` until ``total_steps``, ``end_lr`` afterwards.  Without
    ``step`` returns the schedule as a function of the step (for ``torch.optim.lr_scheduler.LambdaLR`` divide
    by ``base_lr``)."""
    if total_steps <= warmup_steps:
        raise ValueError("total_steps must be larger than warmup_steps")

    def lr(s: int) -> float:
        if s <= warmup_steps:
            return base_lr * s / max(warmup_steps, 1)
        if s >= total_steps:
            return end_lr
        p = (s - warmup_steps) / (total_steps - warmup_steps)
        return (base_lr - end_lr) * (1.0 - p) ** 2 + end_lr

    return lr if step is None else lr(int(step))
