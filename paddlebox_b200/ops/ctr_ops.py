"""The remaining CTR operator family of the reference, expressed on PyTorch tensors.

These ops are not on the hot path of the headline models; they are written as compositions
of PyTorch ops (differentiable through autograd, CPU and CUDA), with the GEMM-shaped ones
able to route through the tcgen05 kernel.  Each cites the reference op it restates.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import functional as F


# ---------------------------------------------------------------- rank_attention
def rank_attention(x: torch.Tensor, rank_offset: torch.Tensor, rank_param: torch.Tensor, max_rank: int = 3,
                   max_size: int = 0) -> torch.Tensor:
    """PV-aware attention over the ads of one page view.

    ``rank_offset`` [N, 2*max_rank+1] int: col 0 = own rank (1-based, <= 0: invalid), then for
    every k: (rank of the k-th ad of the same PV, row index of that ad).  ``rank_param``
    [max_rank*max_rank*C, out]: one [C, out] block per (own rank, other rank) pair.
    out[i] = sum_k x[idx_k] @ P[own-1, rank_k-1]  over valid k.
    Reference: operators/rank_attention_op.cu:30-392, rank_attention.cu.h:28-195; numpy oracle
    test_rank_attention_op.py:25-108."""
    N_, Cc = x.shape
    out_dim = rank_param.shape[1]
    ro = rank_offset.long()
    own = ro[:, 0] - 1                                   # [N]
    faster = ro[:, 1::2] - 1                             # [N, max_rank]
    index = ro[:, 2::2]                                  # [N, max_rank]
    valid = (own[:, None] >= 0) & (faster >= 0)
    xin = x[index.clamp(min=0)] * valid[..., None].to(x.dtype)        # [N, R, C]
    blocks = rank_param.view(max_rank * max_rank, Cc, out_dim)
    bid = (own.clamp(min=0)[:, None] * max_rank + faster.clamp(min=0))  # [N, R]
    P = blocks[bid]                                                     # [N, R, C, out]
    return torch.einsum("nrc,nrco->no", xin, P)


def rank_attention2(x, rank_offset, rank_param, max_rank: int = 3):
    """Same contract as ``rank_attention`` (the reference's second kernel variant)."""
    return rank_attention(x, rank_offset, rank_param, max_rank)


def build_rank_offset(search_id: torch.Tensor, rank: torch.Tensor, max_rank: int = 3,
                      cmatch: torch.Tensor | None = None, valid_cmatch=(222, 223)) -> torch.Tensor:
    """Rank-offset matrix of a PV-merged batch (reference CopyRankOffsetKernel /
    GetRankOffset, data_feed.cu:1319-1369, data_feed.cc:3531-3591): rows of one ``search_id``
    are one page view; ranks are 1-based, ranks > max_rank are ignored.  With ``cmatch`` given, only ads
    whose cmatch is 222 or 223 take part (data_feed.cc:3557-3560), the others get rank -1."""
    n = search_id.numel()
    out = torch.full((n, 2 * max_rank + 1), -1, dtype=torch.int32)
    if n == 0:
        return out
    sid = search_id.reshape(-1).cpu()
    rk = rank.reshape(-1).cpu().long()
    valid = (rk >= 1) & (rk <= max_rank)
    if cmatch is not None:
        cm = cmatch.reshape(-1).cpu().long()
        ok = torch.zeros_like(valid)
        for c in valid_cmatch:
            ok |= cm == c
        valid &= ok
    out[valid, 0] = rk[valid].int()
    _, inv = torch.unique(sid, return_inverse=True)
    rows = torch.arange(n)
    # table[pv, k] = row of the page view's ad with rank k+1 (the last one if a rank repeats)
    table = torch.full((int(inv.max()) + 1, max_rank), -1, dtype=torch.long)
    order = rows[valid]
    table.index_put_((inv[order], rk[order] - 1), order)
    for k in range(max_rank):
        j = table[inv, k]
        has = (j >= 0) & valid
        out[has, 2 * k + 1] = k + 1
        out[has, 2 * k + 2] = j[has].int()
    return out


# ---------------------------------------------------------------- batch_fc
class BatchFC(nn.Module):
    """``out[s] = x[s] @ W[s] + b[s]`` for every slot pair s (reference batch_fc,
    operators/batch_fc_op.cu:34-567; oracle test_batch_fc_op.py:25)."""

    def __init__(self, num: int, in_dim: int, out_dim: int, act: str | None = None):
        super().__init__()
        self.w = nn.Parameter(torch.randn(num, in_dim, out_dim) / in_dim ** 0.5)
        self.b = nn.Parameter(torch.zeros(num, out_dim))
        self.act = act

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return batch_fc(x, self.w, self.b, self.act)


def batch_fc(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, act: str | None = None, batchcount: int = 0,
             transpose_weight: bool = False) -> torch.Tensor:
    """The three layouts of the reference op (batch_fc_op.cc:43-130, batch_fc_op.cu:200-567):

    * default: ``x`` [S, ins, in], ``w`` [S, in, out], ``b`` [S, out] -> [S, ins, out];
    * ``batchcount`` > 0: ``x`` [ins, bc*in], ``w`` [in, bc*out], ``b`` [1, bc*out] -> [ins, bc*out]; column
      block k of the output is ``x``'s block k times ``w``'s block k;
    * ``transpose_weight`` (needs ``batchcount``): ``x`` [bc, ins, in], ``w`` [in, bc*out], ``b`` [1, bc*out]
      -> [bc, ins, out]."""
    if transpose_weight:
        if batchcount <= 0:
            raise ValueError("with transpose_weight, batchcount should be > 0")
        bc, out = batchcount, w.shape[1] // batchcount
        wb = w.view(w.shape[0], bc, out).permute(1, 0, 2)                       # [bc, in, out]
        y = torch.baddbmm(b.reshape(bc, 1, out), x, wb)
    elif batchcount > 0:
        bc, ins = batchcount, x.shape[0]
        fin, out = x.shape[1] // bc, w.shape[1] // bc
        xb = x.view(ins, bc, fin).permute(1, 0, 2)                              # [bc, ins, in]
        wb = w.view(fin, bc, out).permute(1, 0, 2)                              # [bc, in, out]
        y = torch.baddbmm(b.reshape(bc, 1, out), xb, wb).permute(1, 0, 2).reshape(ins, bc * out)
    else:
        y = torch.baddbmm(b[:, None, :], x, w)
    if act == "relu":
        y = torch.relu(y)
    return y


# ---------------------------------------------------------------- scaled_fc
def scaled_fc(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, input_scale_factor: float = 1.0,
              bias_scale_factor: float = 1.0, grad_scale_factor: float = 256.0) -> torch.Tensor:
    """Half-precision FC with range scaling (reference scaled_fc_op.cu:39-342).  The scales exist only to keep
    the half-precision intermediates in range; net of them the op computes

        out = x @ w + bias * (bias_scale_factor / input_scale_factor)      (inf -> nan on the way out)
        dx  = dout @ w^T,   dw = x^T @ dout,   db = column sums of dout    (unscaled, as the reference)

    with every GEMM operand rounded to 16 bits: the forward product is formed as
    ``input_scale * (x16 @ w16) + bias16 * bias_scale`` and divided by ``input_scale``; the backward operand
    is ``dout * grad_scale / input_scale`` and the results are divided by ``grad_scale``.  The reference rounds
    to fp16; here to bf16, the tcgen05 kernel's operand type (fp32 accumulate)."""
    s_in, s_b, s_g = float(input_scale_factor), float(bias_scale_factor), float(grad_scale_factor)

    def r16(t):
        return t.to(torch.bfloat16).float()

    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, w, bias):
            ctx.save_for_backward(x, w)
            y = (s_in * (r16(x) @ r16(w)) + r16(bias) * s_b) / s_in
            return torch.where(torch.isinf(y), torch.full_like(y, float("nan")), y)

        @staticmethod
        def backward(ctx, g):
            x, w = ctx.saved_tensors
            g16 = r16(g * (s_g / s_in))
            dx = s_in * (g16 @ r16(w).t()) / s_g
            dw = s_in * (r16(x).t() @ g16) / s_g
            return dx, dw, g.sum(0)
    return _Fn.apply(x, w, bias)


def _int8_quant(v: torch.Tensor, expand: float, clip: float, int8_range: float) -> torch.Tensor:
    """``int8(clamp(v * expand, -clip, clip) / interval + 0.5)`` with ``interval = 2 * clip / int8_range`` and
    the C cast's truncation toward zero (scaled_int8fc_op.cu:38-72)."""
    interval = 2.0 * clip / int8_range
    return torch.trunc((v * expand).clamp(-clip, clip) / interval + 0.5).clamp(-128, 127)


def scaled_int8fc(x, w, bias, input_scale_factor=1.0, bias_scale_factor=1.0, grad_scale_factor=1.0,
                  input_expand_factor: float = 1.0, input_clip_factor: float = 2.0, weight_expand_factor: float = 1.0,
                  weight_clip_factor: float = 2.0, int8_range: float = 240.0, dequant_weights: bool = False):
    """int8 FC (reference scaled_int8fc_op.cu:38-440): input and weight are expanded, clipped and quantised
    to int8, multiplied with integer accumulation, and the product is scaled back by
    ``input_interval / (input_expand * weight_expand)`` before the (unscaled) bias is added.  The reference
    leaves the WEIGHT interval out of that rescale (its two-interval line is commented out), so with the
    default clip / range the output is ``1 / weight_interval`` times the real product; ``dequant_weights=True``
    applies both intervals instead.  Backward is the plain fp32 FC gradient (straight-through), as in the
    reference (:405-430).  B200 keeps an int8 tensor-core path, but the towers here run bf16: the integer
    product is formed in fp32, which is exact for these magnitudes."""
    in_int = 2.0 * input_clip_factor / int8_range
    w_int = 2.0 * weight_clip_factor / int8_range

    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, w, bias):
            ctx.save_for_backward(x, w)
            xq = _int8_quant(x, input_expand_factor, input_clip_factor, int8_range)
            wq = _int8_quant(w, weight_expand_factor, weight_clip_factor, int8_range)
            y = (xq @ wq) / (input_expand_factor * weight_expand_factor) * in_int
            if dequant_weights:
                y = y * w_int
            return y + bias

        @staticmethod
        def backward(ctx, g):
            x, w = ctx.saved_tensors
            return g @ w.t(), x.t() @ g, g.sum(0)
    return _Fn.apply(x, w, bias)


# ---------------------------------------------------------------- cross_norm_hadamard
class CrossNormHadamard(nn.Module):
    """For ``fields_num`` field pairs (a_i, b_i) of width ``embed_dim`` laid side by side in the input
    (``[a_0, b_0, a_1, b_1, ...]``), emit per pair the block ``[a_i, b_i, a_i*b_i, <a_i, b_i>]`` —
    ``3*embed_dim + 1`` columns, the last one the dot product — and normalise every output column with
    data_norm style running summaries (reference cross_norm_hadamard.cu.h:44-240: ``nncross_normforward_multi``
    + ``..._multi_sim``; mean = sum/size, scale = sqrt(size/square_sum))."""

    def __init__(self, fields_num: int, embed_dim: int, epsilon: float = 1e-4, summary_decay_rate: float = 0.9999999,
                 sync_stats: bool = False):
        super().__init__()
        self.fields_num, self.embed_dim = fields_num, embed_dim
        self.norm = F.DataNorm(fields_num * (3 * embed_dim + 1), epsilon, summary_decay_rate, sync_stats,
                               name="cross_norm")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B = x.shape[0]
        f, d = self.fields_num, self.embed_dim
        pairs = x.view(B, f, 2, d)
        a, b = pairs[:, :, 0], pairs[:, :, 1]
        had = a * b
        y = torch.cat([a, b, had, had.sum(2, keepdim=True)], dim=2).reshape(B, f * (3 * d + 1))
        return self.norm(y)


# ---------------------------------------------------------------- concat / slice utilities
def fused_concat(inputs, start_index: int = 0, length: int = -1, axis: int = 1) -> torch.Tensor:
    """Column slice ``[start, start+length)`` of every input, concatenated (reference
    fused_concat_op.cu:34-258)."""
    outs = []
    for t in inputs:
        s = max(start_index, 0)                       # the layer clamps a negative start to 0 (nn.py:2822)
        e = t.shape[1] if length < 0 else s + length
        outs.append(t[:, s:e])
    return torch.cat(outs, dim=axis)


def _partial_range(width: int, start_index: int, length: int):
    """Column range of the partial ops: a negative start counts from the end, a negative length means "to
    the end" (np oracle, test_partial_concat_op.py:25-40)."""
    if not -width <= start_index < width:
        raise ValueError(f"start_index {start_index} out of range for {width} columns")
    s = start_index + width if start_index < 0 else start_index
    n = width - s if length < 0 else length
    if s + n > width:
        raise ValueError(f"columns [{s}, {s + n}) exceed the input width {width}")
    return s, s + n


def partial_concat(inputs, start_index: int = 0, length: int = -1) -> torch.Tensor:
    """The same column slice of every input, side by side (operators/partial_concat_op.*)."""
    b, e = _partial_range(inputs[0].shape[1], start_index, length)
    return torch.cat([t[:, b:e] for t in inputs], 1)


def partial_sum(inputs, start_index: int = 0, length: int = -1) -> torch.Tensor:
    """Sum of the same column slice of every input (operators/partial_sum_op.*)."""
    b, e = _partial_range(inputs[0].shape[1], start_index, length)
    return torch.stack([t[:, b:e] for t in inputs], 0).sum(0)


def fused_seqpool_concat(inputs, num_slots: int, offsets=None, dims=None, lod: torch.Tensor | None = None,
                         batch_size: int | None = None) -> torch.Tensor:
    """``fluid.contrib.layers.fused_seqpool_concat`` (contrib/layers/nn.py:2726-2800, fused_concat_op.cu):
    two groups of per-slot sequence embeddings are SUM-pooled by ``lod`` and, slot by slot, the columns
    ``[offsets[g], offsets[g] + dims[g])`` of group g are laid side by side.

    ``inputs``: a pair ``[x1, x2]`` of ``[N, E_g]`` tensors in slot-major bag order sharing ``lod``
    (int32 ``[S*B+1]``) -> ``[B, S * (dims[0] + dims[1])]``; ``dims[g] <= 0`` means "to the end".
    A single already pooled tensor ``[B, S*E]`` is also accepted: every ``(offset, dim)`` pair then selects
    a column range of each slot block (the selections are concatenated group after group)."""
    if torch.is_tensor(inputs):
        pooled = inputs
        E = pooled.shape[1] // num_slots
        blocks = pooled.view(pooled.shape[0], num_slots, E)
        if offsets is None:
            return pooled
        return torch.cat([blocks[:, :, o:o + d].reshape(pooled.shape[0], -1) for o, d in zip(offsets, dims)], 1)
    if len(inputs) != 2:
        raise ValueError("fused_seqpool_concat takes exactly two input groups")
    assert lod is not None and batch_size is not None, "sequence inputs need lod and batch_size"
    B, S = batch_size, num_slots
    lens = (lod[1:] - lod[:-1]).long()
    n = int(lod[-1])
    bag = torch.repeat_interleave(torch.arange(S * B, device=lod.device), lens, output_size=n)
    parts = []
    for g, x in enumerate(inputs):
        start = max(0, offsets[g]) if isinstance(offsets, (list, tuple)) and g < len(offsets) else 0
        dim = dims[g] if isinstance(dims, (list, tuple)) and g < len(dims) else 0
        if dim <= 0:
            dim = x.shape[1] - start
        pooled = torch.zeros(S * B, dim, dtype=x.dtype, device=x.device).index_add_(0, bag, x[:n, start:start + dim])
        parts.append(pooled.view(S, B, dim))
    return torch.cat(parts, 2).permute(1, 0, 2).reshape(B, -1)


def fused_seq_tensor(x: torch.Tensor, ad_input: torch.Tensor, batch_count: int, max_length: int, slot_num: int,
                     fea_emb_dim: int, ad_slot_num: int, ad_slot_offset: int = 0):
    """DIN-style sequence tensors (``fluid.contrib.layers.fused_seq_tensor``; kernels
    fused_seq_tensor_op.cu:14-290).

    ``x`` ``[ins, batch_count * slot_num * max_length * dim]``: per instance and session the behaviour
    embeddings laid out ``[slot][time][dim]``; ``ad_input`` ``[ins, batch_count * ad_slot_num * dim]``: the
    candidate ad's embeddings of the ``ad_slot_num`` slots that start at ``ad_slot_offset``.  Returns, all
    session-major ``[batch_count, ins, max_length, ...]``:

    * ``din``  ``[..., 4 * ad_slot_num * dim]`` = ``[seq, ad, seq - ad, seq * ad]`` of the ad slots,
    * ``mask`` ``[batch_count, ins, max_length]`` = 1 where the time step holds any non-zero value
      (``|sum over slots and dims| > 1e-8``),
    * ``side`` ``[..., (slot_num - ad_slot_num) * dim]`` = the remaining (side-info) slots — those after the
      ad slots when ``ad_slot_offset == 0``, else those from slot 0,
    * ``ad_slot_session`` ``[..., ad_slot_num * dim]`` = the sequence values of the ad slots."""
    ins = x.shape[0]
    bc, L, D, an = batch_count, max_length, fea_emb_dim, ad_slot_num
    seq = x.reshape(ins, bc, slot_num, L, D)
    ad = ad_input.reshape(ins, bc, an, D)
    inp = seq[:, :, ad_slot_offset:ad_slot_offset + an].permute(1, 0, 3, 2, 4)          # [bc, ins, L, an, D]
    ad_b = ad.permute(1, 0, 2, 3)[:, :, None].expand(bc, ins, L, an, D)
    sess = inp.reshape(bc, ins, L, an * D)
    adf = ad_b.reshape(bc, ins, L, an * D)
    din = torch.cat([sess, adf, sess - adf, sess * adf], dim=-1)
    so = an if ad_slot_offset == 0 else 0
    side = seq[:, :, so:so + slot_num - an].permute(1, 0, 3, 2, 4).reshape(bc, ins, L, (slot_num - an) * D)
    mask = (seq.sum(dim=(2, 4)).abs() > 1e-8).to(x.dtype).permute(1, 0, 2).contiguous()
    return din, mask, side, sess.contiguous()


def shuffle_batch(x: torch.Tensor, seed: int | None = None):
    """Row permutation of a batch; returns (shuffled, index) (operators/shuffle_batch_op.*)."""
    g = torch.Generator(device="cpu")
    if seed is not None:
        g.manual_seed(seed)
    idx = torch.randperm(x.shape[0], generator=g).to(x.device)
    return x[idx], idx


def filter_by_instag(ins: torch.Tensor, ins_tag_lod, ins_tags: torch.Tensor, filter_tags, out_val_if_empty: float = 0.0,
                     ins_lod=None):
    """Keep the instances that carry at least one of ``filter_tags`` (reference filter_by_instag_op.h:40-193;
    oracle test_filter_by_instag_op.py).  ``ins_tag_lod``: per-instance tag offsets into ``ins_tags``;
    ``ins_lod``: per-instance ROW offsets into ``ins`` (``is_lod``; default one row per instance).

    Returns ``(out, index_map, loss_weight)``: ``out`` = all rows of the kept instances (differentiable),
    ``index_map`` int64 ``[kept, 3]`` = (first output row, first input row, row count), ``loss_weight``
    ``[kept, 1]`` ones.  With nothing kept: one row of ``out_val_if_empty``, map ``[[0, 1, 1]]``, weight 0.
    ``out.lod`` (attribute) holds the output row offsets."""
    want = set(int(t) for t in filter_tags)
    tags = ins_tags.reshape(-1).tolist()
    n_ins = len(ins_tag_lod) - 1
    rows_lod = list(ins_lod) if ins_lod is not None else list(range(n_ins + 1))
    keep = [i for i in range(n_ins) if any(int(t) in want for t in tags[ins_tag_lod[i]:ins_tag_lod[i + 1]])]
    if not keep:
        out = torch.full((1, ins.shape[1]), out_val_if_empty, dtype=ins.dtype, device=ins.device)
        out.lod = [0, 1]
        return out, torch.tensor([[0, 1, 1]], dtype=torch.int64), torch.zeros(1, 1, dtype=torch.float32, device=ins.device)
    src, imap, out_lod = [], [], [0]
    for i in keep:
        b, e = rows_lod[i], rows_lod[i + 1]
        imap.append([out_lod[-1], b, e - b])
        src.extend(range(b, e))
        out_lod.append(out_lod[-1] + e - b)
    out = ins[torch.tensor(src, dtype=torch.long, device=ins.device)]
    out.lod = out_lod
    return out, torch.tensor(imap, dtype=torch.int64), torch.ones(len(keep), 1, dtype=torch.float32, device=ins.device)


# ---------------------------------------------------------------- fused_seqpool_cvm_with_pcoc
class _SeqpoolCVMWithPCOC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cvm, q_values, lod, B, S, use_cvm, used, mx, pad_value, need_filter, show_coeff, clk_coeff,
                threshold, quant_ratio):
        H = x.shape[1]
        G = B * S
        lens = (lod[1:] - lod[:-1]).long()
        n = int(lod[-1])
        seg = torch.repeat_interleave(torch.arange(G, device=x.device), lens, output_size=n)
        rows = x[:n]
        if need_filter:                                   # whole rows are skipped (kernel QuantFilter :87-120)
            keep = ~((rows[:, 0] - rows[:, 1]) * show_coeff + rows[:, 1] * clk_coeff < threshold)
            rows = rows * keep[:, None].to(x.dtype)
        if quant_ratio > 0:                               # only the columns behind max_cvm_offset (:72-80)
            emb = torch.trunc(rows[:, mx:] * quant_ratio + 0.5) / quant_ratio
            rows = torch.cat([rows[:, :mx], emb], 1)
        pooled = torch.zeros(G, H, dtype=x.dtype, device=x.device).index_add_(0, seg, rows) + pad_value
        if use_cvm:
            pn = used - 4                                 # show, clk, show2, clk2, then pclk_0..pclk_{pn-1}
            lg = torch.log(pooled[:, :used] + 1)
            cols = [lg[:, 0:1], lg[:, 1:2] - lg[:, 0:1], lg[:, 4:4 + pn] - lg[:, 2:3], lg[:, 4:4 + pn] - lg[:, 3:4],
                    pooled[:, mx:]]
            out = torch.cat(cols, 1)
        else:
            out = pooled[:, mx:]
        ctx.meta = (B, S, H, n, use_cvm, used, mx, out.shape[1])
        ctx.save_for_backward(cvm, q_values, seg)
        E = out.shape[1]
        return out.view(S, B, E).permute(1, 0, 2).reshape(B, S * E)

    @staticmethod
    def backward(ctx, dout):
        cvm, q_values, seg = ctx.saved_tensors
        B, S, H, n, use_cvm, used, mx, E = ctx.meta
        b_of = seg % B
        dx = torch.zeros(n, H, dtype=dout.dtype, device=dout.device)
        dx[:, :4] = cvm[b_of, :4]                         # show / clk / show2 / clk2 counts ride on the gradient
        if used > 4 and q_values is not None:
            dx[:, 4:used] = q_values.reshape(B, used - 4)[b_of]
        d = dout.reshape(B, S, E)[b_of, seg // B]
        dx[:, mx:] = d[:, E - (H - mx):]
        return (dx,) + (None,) * 14


def fused_seqpool_cvm_with_pcoc(x: torch.Tensor, lod: torch.Tensor, batch_size: int, num_slots: int,
                                cvm: torch.Tensor | None = None, q_values: torch.Tensor | None = None,
                                use_cvm: bool = True, cvm_offset: int = 7, max_cvm_offset: int | None = None,
                                pad_value: float = 0.0, need_filter: bool = False, show_coeff: float = 0.2,
                                clk_coeff: float = 1.0, threshold: float = 0.96, quant_ratio: int = 0):
    """PCOC variant (fused_seqpool_cvm_with_pcoc_op.cu:36-330).  Value rows ``[N, H]`` carry
    ``max_cvm_offset`` statistics columns — ``show, clk, show2, clk2, pclk_0 .. pclk_{n-1}`` with
    ``n = cvm_offset - 4`` (``cvm_offset`` = the op's ``used`` offset; columns ``cvm_offset .. max`` are
    unused) — then the embedding.  Output per (slot, instance), ``use_cvm``:
    ``[log(show+1), log(clk+1)-log(show+1), log(pclk_j+1)-log(show2+1) (n cols), log(pclk_j+1)-log(clk2+1)
    (n cols), embedding...]``; otherwise the embedding only.  Backward writes into every pooled row:
    columns 0-3 from ``cvm`` ``[B, >=4]``, the pclk columns from ``q_values`` ``[B, n]`` (the stored q-values of
    the instances, ``store_q_value``), and the output gradient behind ``max_cvm_offset``."""
    mx = cvm_offset if max_cvm_offset is None else max_cvm_offset
    if cvm_offset < 4 or mx < cvm_offset:
        raise ValueError("with_pcoc needs cvm_offset >= 4 (show, clk, show2, clk2) and max_cvm_offset >= cvm_offset")
    if cvm is None:
        cvm = torch.zeros(batch_size, 4, dtype=x.dtype, device=x.device)
    return _SeqpoolCVMWithPCOC.apply(x, cvm, q_values, lod, batch_size, num_slots, use_cvm, cvm_offset, mx, pad_value,
                                     need_filter, show_coeff, clk_coeff, threshold, quant_ratio)


def cross_norm_layer_hadamard(fields_num: int, embed_dim: int, **kw) -> CrossNormHadamard:
    """``fluid.contrib.layers.cross_norm_layer_hadamard`` (contrib/layers/nn.py): returns the layer
    (parameters = the three data-norm summaries); call it on the ``[B, fields_num*2*embed_dim]`` input."""
    return CrossNormHadamard(fields_num, embed_dim, **kw)
