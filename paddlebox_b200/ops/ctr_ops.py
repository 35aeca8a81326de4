"""The remaining CTR operator family of the reference.

On CUDA tensors every op here runs on this repo's kernels: the GEMM-shaped ones (rank_attention,
batch_fc, scaled_fc, scaled_int8fc) on the tcgen05 GEMM through ``ops/tc_linear.py`` with their
operands built by ``csrc/cuda/ctr_ops.cu``; the memory-bound ones (with_pcoc, fused_concat,
fused_seqpool_concat, fused_seq_tensor, cross_norm_hadamard) as single-pass kernels of the same
file.  The ``*_torch`` functions are the same ops as plain PyTorch compositions: the CPU path
(no GPU in the process) and the oracles of the GPU tests.  Each cites the reference op it restates.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from .. import _native as N
from . import functional as F


def _on_gpu(*ts) -> bool:
    return any(t is not None and torch.is_tensor(t) and t.is_cuda for t in ts)


# ---------------------------------------------------------------- rank_attention
def rank_attention(x: torch.Tensor, rank_offset: torch.Tensor, rank_param: torch.Tensor, max_rank: int = 3,
                   max_size: int = 0) -> torch.Tensor:
    """PV-aware attention over the ads of one page view (see ``rank_attention_torch`` for the contract).

    CUDA: ``k_rank_expand`` gathers the instance's page-view rows into a block-sparse operand
    ``A[i, (own_i*R + rank_k)*C : +C] = x[index_k]`` (bf16, plus its transpose), so that the whole op is
    ONE tcgen05 GEMM ``A[N, R*R*C] . rank_param[R*R*C, out]`` instead of the reference's N per-instance
    products through BatchedGEMM (rank_attention_op.cu:110); the weight gradient is ``A^T . dOut`` on the
    same kernel and the input gradient is scattered back by ``k_rank_dx``."""
    if not _on_gpu(x):
        return rank_attention_torch(x, rank_offset, rank_param, max_rank, max_size)
    return _RankAttention.apply(x.contiguous().float(), rank_offset.contiguous().int(), rank_param, max_rank)


class _RankAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ro, param, R):
        from .tc_linear import _bf16_pair, _LinearTC
        Nn, Cc = x.shape
        K = R * R * Cc
        assert param.shape[0] == K, f"rank_param has {param.shape[0]} rows, expected max_rank^2 * C = {K}"
        pad8 = lambda v: (v + 7) // 8 * 8                                   # noqa: E731
        ab = torch.empty(Nn, pad8(K), dtype=torch.bfloat16, device=x.device)
        at = torch.zeros(K, pad8(Nn), dtype=torch.bfloat16, device=x.device)
        if pad8(K) != K:
            ab[:, K:].zero_()
        N.check(N.cuda().pbx_rank_attention_expand(N.ptr(x), x.stride(0), N.ptr(ro), Nn, Cc, R, N.ptr(ab),
                                                   ab.stride(0), N.ptr(at), at.stride(0), N.stream_ptr()),
                "pbx_rank_attention_expand")
        ctx.meta = (Nn, Cc, R, K)
        ctx.save_for_backward(ro, ab, at, param)
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        wb, wt = _bf16_pair(param)
        cast_transpose_multi([(param.detach().float().contiguous(), wb, wt)])
        ctx.wb = wb
        out = torch.empty(Nn, param.shape[1], dtype=torch.float32, device=x.device)
        gemm_bf16_tn(ab, wt, Nn, param.shape[1], K, out_f32=out)
        return out

    @staticmethod
    def backward(ctx, g):
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        from .tc_linear import _bf16_pair
        ro, ab, at, param = ctx.saved_tensors
        Nn, Cc, R, K = ctx.meta
        O = param.shape[1]
        g = g.contiguous().float()
        gb, gt = _bf16_pair(g)
        cast_transpose_multi([(g, gb, gt)])
        dx = dparam = None
        if ctx.needs_input_grad[2]:
            dparam = torch.empty(K, O, dtype=torch.float32, device=g.device)
            gemm_bf16_tn(at, gt, K, O, Nn, out_f32=dparam)                     # A^T [K, N] . dOut^T [O, N]^T
        if ctx.needs_input_grad[0]:
            dA = torch.empty(Nn, K, dtype=torch.float32, device=g.device)
            gemm_bf16_tn(gb, ctx.wb, Nn, K, O, out_f32=dA)                      # dOut [N, O] . param [K, O]^T
            dx = torch.zeros(Nn, Cc, dtype=torch.float32, device=g.device)
            N.check(N.cuda().pbx_rank_attention_dx(N.ptr(dA), dA.stride(0), N.ptr(ro), Nn, Cc, R, N.ptr(dx),
                                                   dx.stride(0), N.stream_ptr()), "pbx_rank_attention_dx")
        return dx, None, dparam, None


def rank_attention_torch(x: torch.Tensor, rank_offset: torch.Tensor, rank_param: torch.Tensor, max_rank: int = 3,
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
    # every layout is a batch of independent [ins, in] x [in, out] products: on CUDA they run as grouped
    # launches of the tcgen05 GEMM (bias and ReLU in the epilogue), on CPU through torch.baddbmm
    def bmm(xb, wb, bb):
        if _on_gpu(xb):
            from .tc_linear import grouped_linear_tc
            return grouped_linear_tc(xb.float(), wb.float(), bb.float(), relu=(act == "relu"))
        y_ = torch.baddbmm(bb[:, None, :], xb, wb)
        return torch.relu(y_) if act == "relu" else y_

    if transpose_weight:
        if batchcount <= 0:
            raise ValueError("with transpose_weight, batchcount should be > 0")
        bc, out = batchcount, w.shape[1] // batchcount
        wb = w.view(w.shape[0], bc, out).permute(1, 0, 2)                       # [bc, in, out]
        return bmm(x, wb, b.reshape(bc, out))
    if batchcount > 0:
        bc, ins = batchcount, x.shape[0]
        fin, out = x.shape[1] // bc, w.shape[1] // bc
        xb = x.view(ins, bc, fin).permute(1, 0, 2)                              # [bc, ins, in]
        wb = w.view(fin, bc, out).permute(1, 0, 2)                              # [bc, in, out]
        return bmm(xb, wb, b.reshape(bc, out)).permute(1, 0, 2).reshape(ins, bc * out)
    return bmm(x, w, b)


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
    if _on_gpu(x):
        return _ScaledFC.apply(x.float(), w.float(), bias.float(), s_in, s_b, s_g)

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


class _ScaledFC(torch.autograd.Function):
    """scaled_fc on the tcgen05 GEMM: operands rounded to bf16 by one cast launch, the scale factors in the
    epilogue (``alpha`` and the bias factor), forward / dX / dW each one GEMM."""

    @staticmethod
    def forward(ctx, x, w, bias, s_in, s_b, s_g):
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        from .tc_linear import _bf16_pair
        M, K = x.shape
        Nn = w.shape[1]
        xb, xt = _bf16_pair(x)
        wb, wt = _bf16_pair(w)
        cast_transpose_multi([(x.contiguous(), xb, xt), (w.contiguous(), wb, wt)])
        y = torch.empty(M, Nn, dtype=torch.float32, device=x.device)
        # (s_in * (x16 . w16) + s_b * bias16) / s_in
        b_eff = bias.reshape(-1).to(torch.bfloat16).float() * (s_b / s_in)
        gemm_bf16_tn(xb, wt, M, Nn, K, bias=b_eff, out_f32=y, alpha=1.0)
        ctx.save_for_backward(xt, wb)
        ctx.meta = (M, K, Nn, s_in, s_g)
        return torch.where(torch.isinf(y), torch.full_like(y, float("nan")), y)

    @staticmethod
    def backward(ctx, g):
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        from .tc_linear import _bf16_pair
        xt, wb = ctx.saved_tensors
        M, K, Nn, s_in, s_g = ctx.meta
        gs = (g * (s_g / s_in)).contiguous()               # the operand the reference rounds to 16 bits
        gb, gt = _bf16_pair(gs)
        cast_transpose_multi([(gs, gb, gt)])
        dx = torch.empty(M, K, dtype=torch.float32, device=g.device)
        dw = torch.empty(K, Nn, dtype=torch.float32, device=g.device)
        gemm_bf16_tn(gb, wb, M, K, Nn, out_f32=dx, alpha=s_in / s_g)
        gemm_bf16_tn(xt, gt, K, Nn, M, out_f32=dw, alpha=s_in / s_g)
        return dx, dw, g.sum(0), None, None, None


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
    if _on_gpu(x):
        scale = in_int / (input_expand_factor * weight_expand_factor) * (w_int if dequant_weights else 1.0)
        return _ScaledInt8FC.apply(x.float(), w.float(), bias.float(), float(input_expand_factor),
                                   float(input_clip_factor), float(weight_expand_factor), float(weight_clip_factor),
                                   float(int8_range), float(scale))

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


class _ScaledInt8FC(torch.autograd.Function):
    """``k_int8_quant`` writes both operands as integer-valued bf16 (|q| <= 128: exact), the tcgen05 GEMM forms
    the integer product in fp32 (exact while in_dim * 128^2 < 2^24) and applies the rescale as ``alpha`` and the
    bias in its epilogue.  Backward: the plain FC gradient on the un-quantised operands (straight-through), each
    product one bf16 GEMM."""

    @staticmethod
    def forward(ctx, x, w, bias, xe, xc, we, wc, rng, scale):
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        from .tc_linear import _bf16_pair
        M, K = x.shape
        Nn = w.shape[1]
        xq, _ = _bf16_pair(x, True, False)
        _, wqt = _bf16_pair(w, False, True)
        lib = N.cuda()
        N.check(lib.pbx_int8_quant_bf16(N.ptr(x), x.stride(0), M, K, C.c_float(xe), C.c_float(xc), C.c_float(rng),
                                        N.ptr(xq), xq.stride(0), None, 0, N.stream_ptr()), "pbx_int8_quant_bf16")
        N.check(lib.pbx_int8_quant_bf16(N.ptr(w), w.stride(0), K, Nn, C.c_float(we), C.c_float(wc), C.c_float(rng),
                                        None, 0, N.ptr(wqt), wqt.stride(0), N.stream_ptr()), "pbx_int8_quant_bf16")
        y = torch.empty(M, Nn, dtype=torch.float32, device=x.device)
        gemm_bf16_tn(xq, wqt, M, Nn, K, bias=bias.reshape(-1).contiguous(), out_f32=y, alpha=scale)
        # operands of the straight-through backward
        _, xt = _bf16_pair(x, False, True)
        wb, _ = _bf16_pair(w, True, False)
        cast_transpose_multi([(x.contiguous(), None, xt), (w.contiguous(), wb, None)])
        ctx.save_for_backward(xt, wb)
        ctx.meta = (M, K, Nn)
        return y

    @staticmethod
    def backward(ctx, g):
        from .mlp import cast_transpose_multi, gemm_bf16_tn
        from .tc_linear import _bf16_pair
        xt, wb = ctx.saved_tensors
        M, K, Nn = ctx.meta
        g = g.contiguous()
        gb, gt = _bf16_pair(g)
        cast_transpose_multi([(g, gb, gt)])
        dx = torch.empty(M, K, dtype=torch.float32, device=g.device)
        dw = torch.empty(K, Nn, dtype=torch.float32, device=g.device)
        gemm_bf16_tn(gb, wb, M, K, Nn, out_f32=dx)
        gemm_bf16_tn(xt, gt, K, Nn, M, out_f32=dw)
        return dx, dw, g.sum(0), None, None, None, None, None, None


# ---------------------------------------------------------------- cross_norm_hadamard
class _CrossHadamard(torch.autograd.Function):
    """[a, b] field pairs -> [a, b, a*b, <a, b>] blocks; one kernel forward, one backward (ctr_ops.cu)."""

    @staticmethod
    def forward(ctx, x, f, d):
        x = x.contiguous()
        B = x.shape[0]
        y = torch.empty(B, f * (3 * d + 1), dtype=torch.float32, device=x.device)
        N.check(N.cuda().pbx_cross_hadamard_fwd(N.ptr(x), x.stride(0), B, f, d, N.ptr(y), y.stride(0), N.stream_ptr()),
                "pbx_cross_hadamard_fwd")
        ctx.save_for_backward(x)
        ctx.meta = (f, d)
        return y

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        f, d = ctx.meta
        gy = gy.contiguous()
        dx = torch.empty_like(x)
        N.check(N.cuda().pbx_cross_hadamard_bwd(N.ptr(x), x.stride(0), N.ptr(gy), gy.stride(0), x.shape[0], f, d,
                                                N.ptr(dx), dx.stride(0), N.stream_ptr()), "pbx_cross_hadamard_bwd")
        return dx, None, None


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
        if x.is_cuda:                                       # expansion kernel, then the data_norm kernels
            return self.norm(_CrossHadamard.apply(x.float(), f, d))
        pairs = x.view(B, f, 2, d)
        a, b = pairs[:, :, 0], pairs[:, :, 1]
        had = a * b
        y = torch.cat([a, b, had, had.sum(2, keepdim=True)], dim=2).reshape(B, f * (3 * d + 1))
        return self.norm(y)


# ---------------------------------------------------------------- concat / slice utilities
def fused_concat(inputs, start_index: int = 0, length: int = -1, axis: int = 1) -> torch.Tensor:
    """Column slice ``[start, start+length)`` of every input, concatenated (reference
    fused_concat_op.cu:34-258)."""
    inputs = list(inputs)
    s0 = max(start_index, 0)                          # the layer clamps a negative start to 0 (nn.py:2822)
    if _on_gpu(*inputs) and axis == 1 and len(inputs) <= 64 and all(t.dim() == 2 for t in inputs) and \
            (length > 0 or len({t.shape[1] for t in inputs}) == 1):
        ln = length if length > 0 else inputs[0].shape[1] - s0
        return _FusedConcat.apply(s0, ln, *[t.float() for t in inputs])
    outs = []
    for t in inputs:
        s = s0
        e = t.shape[1] if length < 0 else s + length
        outs.append(t[:, s:e])
    return torch.cat(outs, dim=axis)


class _FusedConcat(torch.autograd.Function):
    """One launch gathers the column slice of up to 64 inputs into the output; one launch splits the
    gradient back (zero outside the slice) — fused_concat_op.cu:34-258."""

    @staticmethod
    def _tables(ts):
        n = len(ts)
        return ((C.c_void_p * n)(*[t.data_ptr() for t in ts]), (C.c_int * n)(*[t.stride(0) for t in ts]),
                (C.c_int * n)(*[t.shape[1] for t in ts]))

    @staticmethod
    def forward(ctx, start, length, *xs):
        xs = [t.contiguous() for t in xs]
        B, n = xs[0].shape[0], len(xs)
        out = torch.empty(B, n * length, dtype=torch.float32, device=xs[0].device)
        ptrs, ld, wd = _FusedConcat._tables(xs)
        N.check(N.cuda().pbx_fused_concat(ptrs, None, ld, wd, n, B, start, length, N.ptr(out), out.stride(0), 0,
                                          N.stream_ptr()), "pbx_fused_concat")
        ctx.meta = (start, length, [tuple(t.shape) for t in xs])
        return out

    @staticmethod
    def backward(ctx, g):
        start, length, shapes = ctx.meta
        g = g.contiguous()
        dxs = [torch.empty(sh, dtype=torch.float32, device=g.device) for sh in shapes]
        ptrs, ld, wd = _FusedConcat._tables(dxs)
        N.check(N.cuda().pbx_fused_concat(None, ptrs, ld, wd, len(dxs), shapes[0][0], start, length, N.ptr(g),
                                          g.stride(0), 1, N.stream_ptr()), "pbx_fused_concat(bwd)")
        return (None, None, *dxs)


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
    if _on_gpu(inputs[0]):
        geo = []
        for g, x in enumerate(inputs):
            start = max(0, offsets[g]) if isinstance(offsets, (list, tuple)) and g < len(offsets) else 0
            dim = dims[g] if isinstance(dims, (list, tuple)) and g < len(dims) else 0
            geo.append((start, dim if dim > 0 else x.shape[1] - start))
        return _SeqpoolConcat.apply(inputs[0].float(), inputs[1].float(), lod.int(), B, S, geo[0][0], geo[0][1],
                                    geo[1][0], geo[1][1])
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


class _SeqpoolConcat(torch.autograd.Function):
    """Sum-pool two per-key groups by one lod and lay their column ranges side by side, in one kernel;
    the backward writes every key row of both groups (zero outside the ranges) in one kernel."""

    @staticmethod
    def forward(ctx, x0, x1, lod, B, S, off0, d0, off1, d1):
        x0, x1, lod = x0.contiguous(), x1.contiguous(), lod.contiguous()
        out = torch.empty(B, S * (d0 + d1), dtype=torch.float32, device=x0.device)
        N.check(N.cuda().pbx_seqpool_concat_fwd(N.ptr(x0), x0.stride(0), off0, d0, N.ptr(x1), x1.stride(0), off1, d1,
                                                N.ptr(lod), B, S, N.ptr(out), out.stride(0), N.stream_ptr()),
                "pbx_seqpool_concat_fwd")
        ctx.save_for_backward(lod)
        ctx.meta = (B, S, off0, d0, off1, d1, tuple(x0.shape), tuple(x1.shape))
        return out

    @staticmethod
    def backward(ctx, g):
        (lod,) = ctx.saved_tensors
        B, S, off0, d0, off1, d1, sh0, sh1 = ctx.meta
        g = g.contiguous()
        dx0 = torch.zeros(sh0, dtype=torch.float32, device=g.device)     # rows beyond lod[-1] stay zero
        dx1 = torch.zeros(sh1, dtype=torch.float32, device=g.device)
        N.check(N.cuda().pbx_seqpool_concat_bwd(N.ptr(g), g.stride(0), N.ptr(lod), B, S, N.ptr(dx0), dx0.stride(0),
                                                sh0[1], off0, d0, N.ptr(dx1), dx1.stride(0), sh1[1], off1, d1,
                                                N.stream_ptr()), "pbx_seqpool_concat_bwd")
        return dx0, dx1, None, None, None, None, None, None, None


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
    if x.is_cuda and not (x.requires_grad or ad_input.requires_grad):
        # forward-only in the reference too (fused_seq_tensor_op.cc registers no gradient op): one kernel
        # writes all four outputs; inputs that need gradients take the differentiable composition below
        xc, ac = x.contiguous().float(), ad_input.contiguous().float()
        din = torch.empty(bc, ins, L, 4 * an * D, dtype=torch.float32, device=x.device)
        sess = torch.empty(bc, ins, L, an * D, dtype=torch.float32, device=x.device)
        side = torch.empty(bc, ins, L, (slot_num - an) * D, dtype=torch.float32, device=x.device)
        mask = torch.empty(bc, ins, L, dtype=torch.float32, device=x.device)
        N.check(N.cuda().pbx_fused_seq_tensor(N.ptr(xc), N.ptr(ac), ins, bc, slot_num, L, D, an, ad_slot_offset,
                                              N.ptr(din), N.ptr(sess), N.ptr(side), N.ptr(mask), N.stream_ptr()),
                "pbx_fused_seq_tensor")
        return din, mask.to(x.dtype), side, sess
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
        if x.is_cuda:                 # single-pass kernel: pool + filter + quant + PCOC transform (ctr_ops.cu)
            x = x.contiguous().float()
            lod32 = lod.contiguous().int()
            E = (2 + 2 * (used - 4) if use_cvm else 0) + (H - mx)
            out = torch.empty(B, S * E, dtype=torch.float32, device=x.device)
            N.check(N.cuda().pbx_seqpool_pcoc_fwd(N.ptr(x), x.stride(0), N.ptr(lod32), B, S, H, int(use_cvm), used, mx,
                                                  int(need_filter), int(quant_ratio), C.c_float(pad_value),
                                                  C.c_float(show_coeff), C.c_float(clk_coeff), C.c_float(threshold),
                                                  N.ptr(out), out.stride(0), N.stream_ptr()), "pbx_seqpool_pcoc_fwd")
            ctx.meta = (B, S, H, x.shape[0], use_cvm, used, mx, E)
            ctx.gpu = True
            ctx.save_for_backward(cvm.contiguous().float(), q_values.contiguous().float() if q_values is not None
                                  else None, lod32)
            return out
        ctx.gpu = False
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
        ctx.rows = x.shape[0]
        ctx.save_for_backward(cvm, q_values, seg)
        E = out.shape[1]
        return out.view(S, B, E).permute(1, 0, 2).reshape(B, S * E)

    @staticmethod
    def backward(ctx, dout):
        cvm, q_values, seg = ctx.saved_tensors
        B, S, H, n, use_cvm, used, mx, E = ctx.meta
        if ctx.gpu:
            lod32 = seg
            dout = dout.contiguous()
            dx = torch.zeros(n, H, dtype=torch.float32, device=dout.device)    # rows beyond lod[-1] stay zero
            N.check(N.cuda().pbx_seqpool_pcoc_bwd(N.ptr(dout), dout.stride(0), N.ptr(cvm), cvm.stride(0),
                                                  N.ptr(q_values.reshape(B, used - 4)) if (q_values is not None and
                                                                                          used > 4) else None,
                                                  N.ptr(lod32), B, S, H, int(use_cvm), used, mx, N.ptr(dx),
                                                  dx.stride(0), N.stream_ptr()), "pbx_seqpool_pcoc_bwd")
            return (dx,) + (None,) * 14
        b_of = seg % B
        dx = torch.zeros(n, H, dtype=dout.dtype, device=dout.device)
        dx[:, :4] = cvm[b_of, :4]                         # show / clk / show2 / clk2 counts ride on the gradient
        if used > 4 and q_values is not None:
            dx[:, 4:used] = q_values.reshape(B, used - 4)[b_of]
        d = dout.reshape(B, S, E)[b_of, seg // B]
        dx[:, mx:] = d[:, E - (H - mx):]
        if ctx.rows > n:                                  # rows of x beyond lod[-1] belong to no bag
            dx = torch.cat([dx, torch.zeros(ctx.rows - n, H, dtype=dx.dtype, device=dx.device)], 0)
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
