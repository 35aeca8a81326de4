"""Generic linear / grouped-linear autograd nodes on the tcgen05 GEMM.

The CTR operator family has several GEMM-shaped ops that the reference runs on cuBLAS
(``GEMM`` / ``BatchedGEMM`` call sites: scaled_fc_op.cu:214, :310-312; batch_fc_op.cu:243-539;
rank_attention_op.cu:110-202).  Here they share two nodes built on ``pbx_gemm_bf16_tn`` /
``pbx_gemm_bf16_tn_grouped`` (csrc/cuda/gemm_tcgen05.cu): bf16 operands, fp32 accumulation, the
bias / scale / ReLU in the epilogue, forward + input-gradient + weight-gradient all on the tensor
cores.  Operands are cast (and, where the other GEMM of the pair needs it, transposed) by ONE
``cast_transpose_multi`` launch.
"""
from __future__ import annotations

import torch

from .mlp import _pad8, cast_transpose_multi, gemm_bf16_tn, gemm_bf16_tn_grouped, gemm_problem


def _bf16_pair(x: torch.Tensor, want: bool = True, want_t: bool = True):
    """fp32 [.., M, K] -> bf16 copies [.., M, pad8(K)] and [.., K, pad8(M)] (allocated, not yet filled)."""
    *lead, M, K = x.shape
    o = torch.zeros(*lead, M, _pad8(K), dtype=torch.bfloat16, device=x.device) if want else None
    t = torch.zeros(*lead, K, _pad8(M), dtype=torch.bfloat16, device=x.device) if want_t else None
    return o, t


def _cast_items(x, o, t):
    """(x2d, out, out_t) triples of every leading index for ``cast_transpose_multi``."""
    if x.dim() == 2:
        return [(x, o, t)]
    return [(x[g], o[g] if o is not None else None, t[g] if t is not None else None) for g in range(x.shape[0])]


class _LinearTC(torch.autograd.Function):
    """``y = alpha * (x @ w) + beta * bias`` with prebuilt bf16 operands allowed.

    ``x_ops`` / ``w_ops``: optional ``(bf16 [M, >=K], bf16^T [K, >=M])`` pairs built by a custom
    operand kernel (rank expansion, int8 quantisation); when given, ``x`` / ``w`` only carry the
    autograd edges and shapes."""

    @staticmethod
    def forward(ctx, x, w, bias, alpha, beta, relu, x_ops, w_ops, grad_alpha):
        M, K = x.shape
        N = w.shape[1]
        items = []
        if x_ops is None:
            xb, xt = _bf16_pair(x)
            items += _cast_items(x, xb, xt)
        else:
            xb, xt = x_ops
        if w_ops is None:
            wb, wt = _bf16_pair(w)                 # wb [K, N8] (B operand of dgrad), wt [N, K8] (B operand of fwd)
            items += _cast_items(w, wb, wt)
        else:
            wb, wt = w_ops
        if items:
            cast_transpose_multi(items)
        y = torch.empty(M, N, dtype=torch.float32, device=x.device)
        b_eff = None
        if bias is not None:
            b_eff = bias.reshape(-1).float() if beta == 1.0 else bias.reshape(-1).float() * beta
        gemm_bf16_tn(xb, wt, M, N, K, bias=b_eff, out_f32=y, relu=relu, alpha=alpha)
        ctx.save_for_backward(xt, wb, y if relu else None)
        ctx.meta = (M, K, N, alpha if grad_alpha is None else grad_alpha, beta, relu, bias is not None)
        return y

    @staticmethod
    def backward(ctx, g):
        xt, wb, y = ctx.saved_tensors
        M, K, N, alpha, beta, relu, has_bias = ctx.meta
        g = g.contiguous()
        if relu:
            g = g * (y > 0)
        gb, gt = _bf16_pair(g)
        cast_transpose_multi([(g, gb, gt)])
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=torch.float32, device=g.device)
            gemm_bf16_tn(gb, wb, M, K, N, out_f32=dx, alpha=alpha)            # dout [M,N] . w[K,N]^T
        if ctx.needs_input_grad[1]:
            dw = torch.empty(K, N, dtype=torch.float32, device=g.device)
            gemm_bf16_tn(xt, gt, K, N, M, out_f32=dw, alpha=alpha)            # x^T [K,M] . dout^T[N,M]^T
        if has_bias and ctx.needs_input_grad[2]:
            db = g.sum(0) * beta
        return dx, dw, db, None, None, None, None, None, None


def linear_tc(x, w, bias=None, alpha: float = 1.0, beta: float = 1.0, relu: bool = False, x_ops=None, w_ops=None,
              grad_alpha=None):
    """``alpha * x[M,K] @ w[K,N] + beta * bias`` on the tcgen05 GEMM (CUDA tensors, fp32 in / out)."""
    return _LinearTC.apply(x, w, bias, float(alpha), float(beta), bool(relu), x_ops, w_ops, grad_alpha)


class _GroupedLinearTC(torch.autograd.Function):
    """``y[g] = x[g] @ w[g] + b[g]`` for G independent problems of one shape in grouped launches
    (8 problems per launch): forward, input gradients and weight gradients."""

    @staticmethod
    def forward(ctx, x, w, b, relu):
        G, M, K = x.shape
        N = w.shape[2]
        xb, xt = _bf16_pair(x)
        wb, wt = _bf16_pair(w)
        cast_transpose_multi(_cast_items(x, xb, xt) + _cast_items(w, wb, wt))
        y = torch.empty(G, M, N, dtype=torch.float32, device=x.device)
        bias = b.contiguous().float() if b is not None else None
        gemm_bf16_tn_grouped([gemm_problem(xb[g], wt[g], M, N, K, bias=bias[g] if bias is not None else None,
                                           out_f32=y[g], relu=relu) for g in range(G)])
        ctx.save_for_backward(xt, wb, y if relu else None)
        ctx.meta = (G, M, K, N, relu, b is not None)
        return y

    @staticmethod
    def backward(ctx, gy):
        xt, wb, y = ctx.saved_tensors
        G, M, K, N, relu, has_bias = ctx.meta
        gy = gy.contiguous()
        if relu:
            gy = gy * (y > 0)
        gb, gt = _bf16_pair(gy)
        cast_transpose_multi(_cast_items(gy, gb, gt))
        dx = dw = db = None
        probs = []
        if ctx.needs_input_grad[0]:
            dx = torch.empty(G, M, K, dtype=torch.float32, device=gy.device)
            probs += [gemm_problem(gb[g], wb[g], M, K, N, out_f32=dx[g]) for g in range(G)]
        if ctx.needs_input_grad[1]:
            dw = torch.empty(G, K, N, dtype=torch.float32, device=gy.device)
            probs += [gemm_problem(xt[g], gt[g], K, N, M, out_f32=dw[g]) for g in range(G)]
        if probs:
            gemm_bf16_tn_grouped(probs)
        if has_bias and ctx.needs_input_grad[2]:
            db = gy.sum(1)
        return dx, dw, db, None


def grouped_linear_tc(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor | None = None, relu: bool = False):
    """``x`` [G, M, K], ``w`` [G, K, N], ``b`` [G, N] -> [G, M, N] (CUDA tensors, fp32 in / out)."""
    return _GroupedLinearTC.apply(x.contiguous(), w.contiguous(), b, bool(relu))
