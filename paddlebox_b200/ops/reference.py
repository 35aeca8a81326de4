"""Plain PyTorch (fp32/fp64, CPU or GPU) restatements of the operator math.

These are the correctness oracles for the CUDA kernels and the CPU execution path of the
framework (config 1 of BASELINE.json runs entirely on them).  Each function cites the
reference code whose behaviour it restates; none of it is a translation of that code.
"""
from __future__ import annotations

import math
from typing import Sequence

import torch


# --------------------------------------------------------------------------
# fused_seqpool_cvm family
# reference: operators/fused/fused_seqpool_cvm_op.cc:271-460 (CPU kernels),
#            fused_seqpool_cvm_op.cu:34-1151 (GPU variants: filter / embed filter / concat)
# --------------------------------------------------------------------------
def _quant(v: torch.Tensor, q: int) -> torch.Tensor:
    return torch.trunc(v * q + 0.5) / q


def seqpool_cvm_fwd(x: torch.Tensor, lod: torch.Tensor, B: int, S: int, *, use_cvm=True, cvm_offset=2,
                    pad_value=0.0, need_filter=False, show_coeff=0.2, clk_coeff=1.0, threshold=0.96,
                    quant_ratio=0, clk_filter=False, embed_threshold_filter=False, embed_threshold=0.0,
                    embed_thres_size=0, embedx_concate_size=1, embedx_concate_filter=False, variant=0,
                    show_filter=False, slot_threshold=None, trade_w=None) -> torch.Tensor:
    """x: [N, hidden] value rows in slot-major bag order; lod: [S*B+1]; returns [B, S*Eo]."""
    N_, H = x.shape
    lod_l = lod.tolist()
    concat = embedx_concate_size
    if use_cvm:
        drop = clk_filter if variant == 0 else show_filter
        Eo1 = H - (1 if drop else 0)
    else:
        Eo1 = H - cvm_offset
    out = torch.zeros(B, S * Eo1 * concat, dtype=torch.float32, device=x.device)
    xd = x.double()
    for s in range(S):
        thr = float(slot_threshold[s]) if slot_threshold is not None else threshold
        for b in range(B):
            g = s * B + b
            k0, k1 = lod_l[g], lod_l[g + 1]
            for ci in range(concat):
                if concat > 1:
                    a0 = k0 + ci
                    a1 = min(a0 + 1, k1)
                else:
                    a0, a1 = k0, k1
                rows = xd[a0:a1] if a1 > a0 else xd[0:0]
                filtering = need_filter or embed_threshold_filter
                if concat > 1 and embed_threshold_filter and not embedx_concate_filter:
                    filtering = False
                if filtering and rows.shape[0]:
                    keep = torch.ones(rows.shape[0], dtype=torch.bool, device=x.device)
                    if need_filter:
                        sc = (rows[:, 0] - rows[:, 1]) * show_coeff + rows[:, 1] * clk_coeff
                        # kernel evaluates this in fp32
                        sc32 = ((x[a0:a1, 0] - x[a0:a1, 1]) * show_coeff + x[a0:a1, 1] * clk_coeff)
                        keep &= ~(sc32 < thr)
                    if embed_threshold_filter:
                        e = x[a0:a1, cvm_offset:]
                        score = torch.sqrt((e[:, 1:embed_thres_size] ** 2).sum(1)) + e[:, 0].abs()
                        keep &= ~(score < embed_threshold)
                    rows = rows[keep]
                    tw = trade_w[a0:a1][keep].double() if trade_w is not None else None
                else:
                    tw = trade_w[a0:a1].double() if trade_w is not None else None
                vals = rows.clone()
                if vals.shape[0]:
                    emb = vals[:, cvm_offset:]
                    if tw is not None:
                        emb = emb * tw[:, None]
                    if quant_ratio > 0:
                        emb = _quant(emb.float(), quant_ratio).double()
                    vals = torch.cat([vals[:, :cvm_offset], emb], 1)
                pooled = vals.sum(0) + pad_value if vals.shape[0] else torch.full((H,), pad_value, dtype=torch.float64,
                                                                               device=x.device)
                pooled = pooled.float()
                if use_cvm:
                    res = pooled.clone()
                    if variant == 0:
                        res[0] = torch.log(pooled[0] + 1)
                        res[1] = torch.log(pooled[1] + 1) - torch.log(pooled[0] + 1)
                        if clk_filter:
                            res = torch.cat([res[:1], res[2:]])
                    elif variant == 1:  # with_conv: show, clk, conv
                        res[0] = torch.log(pooled[0] + 1)
                        res[1] = torch.log(pooled[1] + 1)
                        res[2] = torch.log(pooled[2] + 1) - torch.log(pooled[1] + 1)
                        if show_filter:
                            res = res[1:]
                    else:  # with_credit: every cvm column log(x+1)
                        res[:cvm_offset] = torch.log(pooled[:cvm_offset] + 1)
                        if show_filter:
                            res = res[1:]
                else:
                    res = pooled[cvm_offset:]
                base = s * Eo1 * concat + ci * Eo1
                out[b, base:base + Eo1] = res
    return out


def seqpool_plain(kw: dict) -> bool:
    """True when the attributes need none of the per-bag special cases (filters, quantisation,
    embedx concat, trade weights, per-slot thresholds): the vectorised CPU path applies."""
    return (not kw.get("need_filter") and not kw.get("embed_threshold_filter") and not kw.get("quant_ratio")
            and kw.get("embedx_concate_size", 1) == 1 and kw.get("slot_threshold") is None
            and kw.get("trade_w") is None)


def seqpool_cvm_fwd_vec(x: torch.Tensor, lod: torch.Tensor, B: int, S: int, *, use_cvm=True, cvm_offset=2,
                        pad_value=0.0, clk_filter=False, variant=0, show_filter=False, **_) -> torch.Tensor:
    """Vectorised ``seqpool_cvm_fwd`` for plain attributes (one index_add instead of a loop over bags):
    the CPU execution path of the framework; the loop above stays the oracle."""
    H = x.shape[1]
    lens = (lod[1:] - lod[:-1]).long()
    n = int(lod[-1])
    bag = torch.repeat_interleave(torch.arange(S * B, device=x.device), lens, output_size=n)
    pooled = torch.zeros(S * B, H, dtype=torch.float64, device=x.device).index_add_(0, bag, x[:n].double())
    pooled = (pooled + pad_value).float()
    if use_cvm:
        res = pooled.clone()
        if variant == 0:
            res[:, 0] = torch.log(pooled[:, 0] + 1)
            res[:, 1] = torch.log(pooled[:, 1] + 1) - torch.log(pooled[:, 0] + 1)
            if clk_filter:
                res = torch.cat([res[:, :1], res[:, 2:]], 1)
        elif variant == 1:
            res[:, 0] = torch.log(pooled[:, 0] + 1)
            res[:, 1] = torch.log(pooled[:, 1] + 1)
            res[:, 2] = torch.log(pooled[:, 2] + 1) - torch.log(pooled[:, 1] + 1)
            if show_filter:
                res = res[:, 1:]
        else:
            res[:, :cvm_offset] = torch.log(pooled[:, :cvm_offset] + 1)
            if show_filter:
                res = res[:, 1:]
    else:
        res = pooled[:, cvm_offset:]
    Eo = res.shape[1]
    return res.reshape(S, B, Eo).permute(1, 0, 2).reshape(B, S * Eo).contiguous()


def seqpool_cvm_bwd_vec(dout: torch.Tensor, cvm: torch.Tensor, lod: torch.Tensor, B: int, S: int, hidden: int, *,
                        use_cvm=True, cvm_offset=2, clk_filter=False, variant=0, show_filter=False, **_) -> torch.Tensor:
    """Vectorised ``seqpool_cvm_bwd`` (embedx_concate_size 1)."""
    lens = (lod[1:] - lod[:-1]).long()
    n = int(lod[-1])
    dim_off = (1 if (clk_filter if variant == 0 else show_filter) else 0) if use_cvm else cvm_offset
    Eo = hidden - dim_off
    bag = torch.repeat_interleave(torch.arange(S * B, device=dout.device), lens, output_size=n)
    s_of, b_of = bag // B, bag % B
    dx = torch.empty(n, hidden, dtype=torch.float32, device=dout.device)
    dx[:, :cvm_offset] = cvm[b_of, :cvm_offset]
    dx[:, cvm_offset:] = dout[:, :S * Eo].reshape(B, S, Eo)[b_of, s_of, cvm_offset - dim_off:]
    return dx


def seqpool_cvm_bwd(dout: torch.Tensor, cvm: torch.Tensor, lod: torch.Tensor, B: int, S: int, hidden: int, *,
                    use_cvm=True, cvm_offset=2, clk_filter=False, embedx_concate_size=1, variant=0,
                    show_filter=False) -> torch.Tensor:
    """dX [N, hidden]: first cvm_offset columns carry the CVM input (show/click labels), the
    rest is dOut broadcast over the bag (fused_seqpool_cvm_op.cc:394-460)."""
    lod_l = lod.tolist()
    N_ = lod_l[-1]
    concat = embedx_concate_size
    if use_cvm:
        dim_off = 1 if (clk_filter if variant == 0 else show_filter) else 0
    else:
        dim_off = cvm_offset
    Eo1 = hidden - dim_off
    dx = torch.zeros(N_, hidden, dtype=torch.float32, device=dout.device)
    for s in range(S):
        for b in range(B):
            g = s * B + b
            for k in range(lod_l[g], lod_l[g + 1]):
                ci = min(k - lod_l[g], concat - 1)
                base = s * Eo1 * concat + ci * Eo1
                dx[k, :cvm_offset] = cvm[b, :cvm_offset]
                dx[k, cvm_offset:] = dout[b, base + cvm_offset - dim_off: base + Eo1]
    return dx


# --------------------------------------------------------------------------
# cvm / data_norm  (operators/cvm_op.h, data_norm_op.cu:39-110)
# --------------------------------------------------------------------------
def cvm_fwd(x: torch.Tensor, use_cvm: bool) -> torch.Tensor:
    if use_cvm:
        y = x.clone()
        y[:, 0] = torch.log(x[:, 0] + 1)
        y[:, 1] = torch.log(x[:, 1] + 1) - y[:, 0]
        return y
    return x[:, 2:].clone()


def cvm_bwd(dy: torch.Tensor, cvm: torch.Tensor, W: int, use_cvm: bool) -> torch.Tensor:
    dx = torch.empty(dy.shape[0], W, dtype=dy.dtype, device=dy.device)
    dx[:, :2] = cvm[:, :2]
    dx[:, 2:] = dy[:, 2:] if use_cvm else dy
    return dx


def data_norm_fwd(x, batch_size, batch_sum, batch_square_sum):
    mean = batch_sum / batch_size
    scale = torch.sqrt(batch_size / batch_square_sum)
    return (x - mean) * scale, mean, scale


def data_norm_bwd(x, dy, mean, scale, eps=1e-4):
    n = x.shape[0]
    dx = dy * scale
    d_size = torch.ones_like(mean)
    d_sum = x.sum(0) / n
    d_sq = ((x - mean) ** 2).sum(0) / n + eps
    return dx, d_size, d_sum, d_sq


# --------------------------------------------------------------------------
# sparse optimizers (heter_ps/optimizer.cuh.h:30-472) on one row, python floats
# --------------------------------------------------------------------------
def sparse_row_update(row: list, push: list, layout, cfg, key: int = 0, seed: int = 0, uniform=None) -> list:
    """Apply one merged push record to one value row (lists of python floats).
    ``layout`` is a RowLayout, ``cfg`` a SparseOptConfig.  ``uniform(key, i)`` supplies the
    lazy-creation random numbers (defaults to zeros so that tests are deterministic)."""
    L, c = layout, cfg
    row = list(row)
    D = L.embedx_dim
    g_show, g_clk = push[1], push[2]
    if g_show <= 0:
        return row
    row[L.off_tail + 1] = push[0]
    row[0] += g_show
    row[1] += g_clk
    row[L.off_tail + 0] += c.nonclk_coeff * (g_show - g_clk) + c.clk_coeff * g_clk
    row[L.off_tail + 2] = 0.0

    def adagrad(w_off, st_off, g, n, lr, g2_0, lo, hi):
        ratio = lr * math.sqrt(g2_0 / (g2_0 + row[st_off]))
        add = 0.0
        for i in range(n):
            sg = g[i] / g_show
            row[w_off + i] = min(max(row[w_off + i] + sg * ratio, lo), hi)
            add += sg * sg
        row[st_off] += add / n

    def adam(w_off, m_off, v_off, b1p_off, b2p_off, g, n, lr, b1, b2, eps, lo, hi, shared):
        b1p, b2p = row[b1p_off], row[b2p_off]
        ratio = lr * math.sqrt(1.0 - b2p) / (1.0 - b1p)
        sm = sv = 0.0
        m0, v0 = row[m_off], row[v_off]
        for i in range(n):
            sg = g[i] / g_show
            mi = row[m_off + i] if not shared else m0
            vi = row[v_off + i] if not shared else v0
            nm = b1 * mi + (1 - b1) * sg
            nv = b2 * vi + (1 - b2) * sg * sg
            row[w_off + i] = min(max(row[w_off + i] + ratio * nm / (math.sqrt(nv) + eps), lo), hi)
            if not shared:
                row[m_off + i], row[v_off + i] = nm, nv
            sm += nm
            sv += nv
        if shared:
            row[m_off], row[v_off] = sm / n, sv / n
        row[b1p_off] = b1p * b1
        row[b2p_off] = b2p * b2

    es, xs = L.off_es, L.off_xs
    if L.optimizer == 0:
        adagrad(2, es, push[3:4], 1, c.learning_rate, c.initial_g2sum, c.min_bound, c.max_bound)
    else:
        adam(2, es, es + 1, es + 2, es + 3, push[3:4], 1, c.learning_rate, c.beta1_decay_rate, c.beta2_decay_rate,
             c.ada_epsilon, c.min_bound, c.max_bound, False)
    if row[L.off_mf] == 0:
        if c.mf_create_thresholds <= c.nonclk_coeff * (row[0] - row[1]) + c.clk_coeff * row[1]:
            row[L.off_mf] = float(D)
            for i in range(D):
                row[3 + i] = (uniform(key, i) if uniform else 0.0) * c.mf_initial_range
            if L.optimizer == 1:
                row[xs + 2 * D], row[xs + 2 * D + 1] = c.mf_beta1_decay_rate, c.mf_beta2_decay_rate
            elif L.optimizer == 2:
                row[xs + 2], row[xs + 3] = c.mf_beta1_decay_rate, c.mf_beta2_decay_rate
    else:
        g = push[4:4 + D]
        if L.optimizer == 0:
            adagrad(3, xs, g, D, c.mf_learning_rate, c.mf_initial_g2sum, c.mf_min_bound, c.mf_max_bound)
        elif L.optimizer == 1:
            adam(3, xs, xs + D, xs + 2 * D, xs + 2 * D + 1, g, D, c.mf_learning_rate, c.mf_beta1_decay_rate,
                 c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound, False)
        else:
            adam(3, xs, xs + 1, xs + 2, xs + 3, g, D, c.mf_learning_rate, c.mf_beta1_decay_rate,
                 c.mf_beta2_decay_rate, c.mf_ada_epsilon, c.mf_min_bound, c.mf_max_bound, True)
    return row


def new_row(layout, cfg) -> list:
    row = [0.0] * layout.stride
    if layout.optimizer != 0:
        row[layout.off_es + 2] = cfg.beta1_decay_rate
        row[layout.off_es + 3] = cfg.beta2_decay_rate
    return row


# --------------------------------------------------------------------------
# dedup contract (heter_ps/heter_comm_inl.h:2231-2343)
# --------------------------------------------------------------------------
def dedup_keys_and_fillidx(keys: torch.Tensor):
    """merged_keys (sorted unique), restore_idx, sorted_idx, offset, merged_cnts."""
    merged, restore, cnts = torch.unique(keys, sorted=True, return_inverse=True, return_counts=True)
    sorted_idx = torch.argsort(keys, stable=True)
    offset = torch.cumsum(cnts, 0) - cnts
    return merged, restore, sorted_idx, offset, cnts


def merge_push_sorted(restore: torch.Tensor, n_unique: int, grads: torch.Tensor, key2slot: torch.Tensor,
                      cvm_offset: int, grad_scale: float) -> torch.Tensor:
    """Deterministic twin of ``merge_push``: the occurrences are sorted by unique row (stable) and every row's
    gradients are summed in fp64 in that fixed order — the reference's segmented merge (PushMergeCopy,
    box_wrapper.cu:417-475: sorted index, fp64 accumulation), chosen there for batches with hot keys.  The slot id
    of a row is that of its first occurrence.  Works on any device; no atomics."""
    H = grads.shape[1]
    r = restore.long().reshape(-1)
    out = torch.zeros(n_unique, H + 1, dtype=torch.float64, device=grads.device)
    if r.numel() == 0:
        return out.float()
    order = torch.argsort(r, stable=True)
    rs = r[order]
    g = grads.double()[order]
    g[:, cvm_offset:] *= grad_scale
    uniq, counts = torch.unique_consecutive(rs, return_counts=True)
    sums = torch.segment_reduce(g, "sum", lengths=counts, axis=0)
    starts = torch.cumsum(counts, 0) - counts
    out[uniq, 1:] = sums
    out[uniq, 0] = key2slot.reshape(-1)[order][starts].double()
    return out.float()


def merge_push(restore: torch.Tensor, n_unique: int, grads: torch.Tensor, key2slot: torch.Tensor, cvm_offset: int,
               grad_scale: float) -> torch.Tensor:
    """Push records [U, 4 + D] = [slot, show, clk, embed_g, embedx_g] merged by key
    (box_wrapper.cu:417-512)."""
    H = grads.shape[1]
    g = grads.clone().double()
    g[:, cvm_offset:] *= grad_scale
    out = torch.zeros(n_unique, H + 1, dtype=torch.float64, device=grads.device)
    out[:, 1:].index_add_(0, restore.long(), g)
    out[restore.long(), 0] = key2slot.double()
    return out.float()


# --------------------------------------------------------------------------
# FM head, rank_attention, batch_fc oracles
# --------------------------------------------------------------------------
def fm_fwd(x: torch.Tensor, S: int, E: int, wcol: int, vcol: int, D: int) -> torch.Tensor:
    xs = x[:, :S * E].reshape(x.shape[0], S, E)
    v = xs[:, :, vcol:vcol + D]
    first = xs[:, :, wcol].sum(1)
    second = 0.5 * ((v.sum(1) ** 2) - (v ** 2).sum(1)).sum(1)
    return first + second
