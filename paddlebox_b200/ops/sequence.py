"""Ragged-sequence (LoD) operators and the ranking / distillation losses CTR job scripts call around the sparse pull.

Reference: ``python/paddle/fluid/layers/sequence_lod.py`` (``sequence_pool`` ... ``sequence_reverse``),
``python/paddle/fluid/layers/loss.py`` (``rank_loss``, ``margin_rank_loss``, ``bpr_loss``,
``teacher_student_sigmoid_loss``), ``layers/nn.py`` (``log_loss``, ``hash``), ``contrib/layers/nn.py``
(``match_matrix_tensor``, ``sequence_topk_avg_pooling``) and the kernels under
``paddle/fluid/operators/sequence_ops/``.  The reference attaches the sequence boundaries to the tensor (LoD); here a
ragged tensor is the pair ``(x, lod)``: ``x`` is ``[N, ...]`` with the rows of all sequences back to back and ``lod``
the ``[n_seq + 1]`` offsets (what ``SlotBatch.lod`` / ``SlotBatch.slot_lod(s)`` hold).  An op whose result is ragged
returns ``(out, out_lod)``.  Everything is index arithmetic on the offsets plus ``index_add`` / ``scatter_reduce`` /
gathers, so it runs on the device the data is on and is differentiable through autograd; the fused
``fused_seqpool_cvm`` family (ops/seqpool.py) stays the fast path for the pull -> pool -> CVM chain.
"""
from __future__ import annotations

import numpy as np
import torch

__all__ = [
    "sequence_pool", "sequence_first_step", "sequence_last_step", "sequence_softmax", "sequence_concat",
    "sequence_expand", "sequence_expand_as", "sequence_pad", "sequence_unpad", "sequence_reverse", "sequence_slice",
    "sequence_mask", "sequence_enumerate", "sequence_reshape", "sequence_scatter", "sequence_conv",
    "sequence_topk_avg_pooling", "match_matrix_tensor", "rank_loss", "margin_rank_loss", "bpr_loss", "log_loss",
    "teacher_student_sigmoid_loss", "hash",
]


def _lod(lod, device=None) -> torch.Tensor:
    t = torch.as_tensor(lod)
    if t.dtype != torch.int64:
        t = t.long()
    return t.to(device) if device is not None and t.device != device else t


def _seg(lod: torch.Tensor, n: int) -> torch.Tensor:
    """Row -> sequence index."""
    return torch.repeat_interleave(torch.arange(lod.numel() - 1, device=lod.device), lod[1:] - lod[:-1], output_size=n)


def _pos(lod: torch.Tensor, seg: torch.Tensor) -> torch.Tensor:
    """Row -> position inside its sequence."""
    return torch.arange(seg.numel(), device=seg.device) - lod[:-1][seg]


# ---- pooling ---------------------------------------------------------------------------------------------------
def sequence_pool(x: torch.Tensor, lod, pool_type: str = "sum", is_test: bool = False, pad_value: float = 0.0):
    """``[n_seq, ...]``: every sequence reduced over its rows; an empty sequence yields ``pad_value``
    (sequence_lod.py ``sequence_pool``; pool types sum / average / sqrt / max / min / first / last)."""
    lod = _lod(lod, x.device)
    n_seq, n = lod.numel() - 1, x.shape[0]
    lens = lod[1:] - lod[:-1]
    kind = pool_type.lower()
    shape1 = (n_seq,) + (1,) * (x.dim() - 1)
    if kind in ("sum", "average", "sqrt"):
        out = torch.zeros((n_seq,) + x.shape[1:], dtype=x.dtype, device=x.device).index_add(0, _seg(lod, n), x)
        if kind == "average":
            out = out / lens.clamp(min=1).to(x.dtype).view(shape1)
        elif kind == "sqrt":
            out = out / lens.clamp(min=1).to(x.dtype).sqrt().view(shape1)
    elif kind in ("max", "min"):
        idx = _seg(lod, n).view((n,) + (1,) * (x.dim() - 1)).expand_as(x)
        out = torch.zeros((n_seq,) + x.shape[1:], dtype=x.dtype, device=x.device).scatter_reduce(
            0, idx, x, "amax" if kind == "max" else "amin", include_self=False)
    elif kind in ("first", "last"):
        at = lod[:-1] if kind == "first" else lod[1:] - 1
        out = x[at.clamp(0, max(n - 1, 0))] if n else torch.zeros((n_seq,) + x.shape[1:], dtype=x.dtype, device=x.device)
    else:
        raise ValueError(f"unknown pool_type {pool_type!r}")
    empty = (lens == 0).view(shape1)
    if bool(empty.any()) or kind in ("first", "last"):
        out = torch.where(empty, torch.full_like(out, pad_value), out)
    return out


def sequence_first_step(x, lod):
    return sequence_pool(x, lod, "first")


def sequence_last_step(x, lod):
    return sequence_pool(x, lod, "last")


def sequence_softmax(x: torch.Tensor, lod):
    """Softmax inside every sequence of a ``[N]`` / ``[N, 1]`` tensor."""
    lod = _lod(lod, x.device)
    flat = x.reshape(-1)
    seg = _seg(lod, flat.numel())
    n_seq = lod.numel() - 1
    mx = torch.full((n_seq,), -torch.inf, dtype=x.dtype, device=x.device).scatter_reduce(0, seg, flat.detach(), "amax")
    e = (flat - mx[seg]).exp()
    den = torch.zeros(n_seq, dtype=x.dtype, device=x.device).index_add(0, seg, e)
    return (e / den[seg]).view_as(x)


# ---- re-arranging ---------------------------------------------------------------------------------------------
def sequence_concat(xs, lods):
    """Sequence i of the result = sequence i of every input, one after the other."""
    lods = [_lod(l, x.device) for x, l in zip(xs, lods)]
    n_seq = lods[0].numel() - 1
    lens = torch.stack([l[1:] - l[:-1] for l in lods])                       # [n_in, n_seq]
    out_lod = torch.zeros(n_seq + 1, dtype=torch.int64, device=xs[0].device)
    out_lod[1:] = lens.sum(0).cumsum(0)
    before = lens.cumsum(0) - lens                                           # rows of earlier inputs in sequence i
    out = torch.empty((int(out_lod[-1]),) + xs[0].shape[1:], dtype=xs[0].dtype, device=xs[0].device)
    dst = []
    for k, (x, l) in enumerate(zip(xs, lods)):
        seg = _seg(l, x.shape[0])
        dst.append(out_lod[:-1][seg] + before[k][seg] + _pos(l, seg))
    return out.index_copy(0, torch.cat(dst), torch.cat(list(xs))), out_lod


def sequence_expand(x: torch.Tensor, y_lod, x_lod=None):
    """Repeat by the sequence lengths of ``y``: without ``x_lod`` row i is repeated ``len_y[i]`` times; with it,
    the whole sequence i is repeated ``len_y[i]`` times (and the result is ragged: returns ``(out, out_lod)``)."""
    y_lod = _lod(y_lod, x.device)
    reps = y_lod[1:] - y_lod[:-1]
    if x_lod is None:
        return x.repeat_interleave(reps, dim=0)
    x_lod = _lod(x_lod, x.device)
    lens = x_lod[1:] - x_lod[:-1]
    out_lens = lens.repeat_interleave(reps)                                  # one entry per produced sequence
    starts = x_lod[:-1].repeat_interleave(reps)
    out_lod = torch.zeros(out_lens.numel() + 1, dtype=torch.int64, device=x.device)
    out_lod[1:] = out_lens.cumsum(0)
    seg = _seg(out_lod, int(out_lod[-1]))
    return x[starts[seg] + _pos(out_lod, seg)], out_lod


def sequence_expand_as(x: torch.Tensor, y_lod):
    """Row i of ``x`` repeated to the length of sequence i of ``y``."""
    y_lod = _lod(y_lod, x.device)
    return x.repeat_interleave(y_lod[1:] - y_lod[:-1], dim=0)


def sequence_pad(x: torch.Tensor, lod, pad_value=0.0, maxlen: int | None = None):
    """``([n_seq, maxlen, ...] padded, [n_seq] lengths)``."""
    lod = _lod(lod, x.device)
    lens = lod[1:] - lod[:-1]
    L = int(lens.max()) if maxlen is None and lens.numel() else int(maxlen or 0)
    if lens.numel() and int(lens.max()) > L:
        raise ValueError(f"maxlen {L} is shorter than the longest sequence ({int(lens.max())})")
    n_seq = lens.numel()
    pad = torch.as_tensor(pad_value, dtype=x.dtype, device=x.device)
    out = pad.expand((n_seq, L) + x.shape[1:]).clone() if pad.dim() == 0 or pad.numel() == 1 else \
        pad.expand(x.shape[1:]).expand((n_seq, L) + x.shape[1:]).clone()
    seg = _seg(lod, x.shape[0])
    return out.index_put((seg, _pos(lod, seg)), x), lens


def sequence_unpad(x: torch.Tensor, length):
    """The inverse of ``sequence_pad``: ``(rows, lod)``."""
    length = _lod(length, x.device).reshape(-1)
    lod = torch.zeros(length.numel() + 1, dtype=torch.int64, device=x.device)
    lod[1:] = length.cumsum(0)
    seg = _seg(lod, int(lod[-1]))
    return x[seg, _pos(lod, seg)], lod


def sequence_reverse(x: torch.Tensor, lod):
    lod = _lod(lod, x.device)
    seg = _seg(lod, x.shape[0])
    return x[lod[1:][seg] - 1 - _pos(lod, seg)]


def sequence_slice(x: torch.Tensor, lod, offset, length):
    """Rows ``[offset[i], offset[i] + length[i])`` of every sequence: ``(out, out_lod)``."""
    lod = _lod(lod, x.device)
    offset, length = _lod(offset, x.device).reshape(-1), _lod(length, x.device).reshape(-1)
    if bool(((offset < 0) | (length < 0) | (offset + length > lod[1:] - lod[:-1])).any()):
        raise ValueError("sequence_slice: offset + length runs past a sequence")
    out_lod = torch.zeros(length.numel() + 1, dtype=torch.int64, device=x.device)
    out_lod[1:] = length.cumsum(0)
    seg = _seg(out_lod, int(out_lod[-1]))
    return x[(lod[:-1] + offset)[seg] + _pos(out_lod, seg)], out_lod


def sequence_mask(x, maxlen: int | None = None, dtype=torch.int64):
    """``y[..., j] = j < x[...]``."""
    x = torch.as_tensor(x)
    L = int(x.max()) if maxlen is None else int(maxlen)
    return (torch.arange(L, device=x.device) < x.unsqueeze(-1)).to(dtype)


def sequence_enumerate(x: torch.Tensor, lod, win_size: int, pad_value: int = 0):
    """``[N, win_size]``: row t holds ``x[t : t + win_size]`` of its sequence, ``pad_value`` past the end."""
    lod = _lod(lod, x.device)
    flat = x.reshape(-1)
    n = flat.numel()
    seg = _seg(lod, n)
    at = torch.arange(n, device=x.device).unsqueeze(1) + torch.arange(win_size, device=x.device)
    inside = at < lod[1:][seg].unsqueeze(1)
    return torch.where(inside, flat[at.clamp(max=max(n - 1, 0))], torch.full_like(at, pad_value).to(flat.dtype))


def sequence_reshape(x: torch.Tensor, lod, new_dim: int):
    """Rows re-cut to width ``new_dim``; every sequence must hold a whole number of new rows."""
    lod = _lod(lod, x.device)
    d = x.shape[1]
    if bool(((lod * d) % new_dim != 0).any()):
        raise ValueError("sequence_reshape: a sequence does not divide into rows of the new width")
    return x.reshape(-1, new_dim), lod * d // new_dim


def sequence_scatter(x: torch.Tensor, index: torch.Tensor, updates: torch.Tensor, lod):
    """``out[i, index[j]] += updates[j]`` for every j of sequence i (``x`` is ``[n_seq, D]``)."""
    lod = _lod(lod, x.device)
    idx = index.reshape(-1).long()
    return x.index_put((_seg(lod, idx.numel()), idx), updates.reshape(-1).to(x.dtype), accumulate=True)


# ---- context / matching ---------------------------------------------------------------------------------------
def sequence_conv(x: torch.Tensor, lod, weight: torch.Tensor, bias: torch.Tensor | None = None, filter_size: int = 3,
                  padding_start: int | None = None, act: str | None = None):
    """Context projection: row t becomes ``concat(x[t + padding_start + k] for k in range(filter_size))`` (zeros
    outside its sequence) times ``weight [filter_size * D, num_filters]``."""
    lod = _lod(lod, x.device)
    n, d = x.shape
    start = -(filter_size // 2) if padding_start is None else int(padding_start)
    seg = _seg(lod, n)
    src = torch.arange(n, device=x.device).unsqueeze(1) + (start + torch.arange(filter_size, device=x.device))
    inside = (src >= lod[:-1][seg].unsqueeze(1)) & (src < lod[1:][seg].unsqueeze(1))
    ctx = x[src.clamp(0, max(n - 1, 0))] * inside.unsqueeze(-1).to(x.dtype)                # [N, K, D]
    out = ctx.reshape(n, filter_size * d) @ weight
    if bias is not None:
        out = out + bias
    return getattr(torch, act)(out) if act else out


def match_matrix_tensor(x: torch.Tensor, x_lod, y: torch.Tensor, y_lod, w: torch.Tensor, act: str | None = None):
    """``(out, out_lod, tmp)``: for pair i and channel c the ``[len_x, len_y]`` matrix ``x_i W_c y_i^T`` with
    ``w [D, channel_num, D]``; ``out`` is ``[sum_i channel_num * len_x * len_y, 1]`` in (pair, channel, x, y)
    order, ``tmp = x W`` (match_matrix_tensor_op.cc:239-330)."""
    x_lod, y_lod = _lod(x_lod, x.device), _lod(y_lod, x.device)
    d, c = w.shape[0], w.shape[1]
    tmp = x @ w.reshape(d, c * d)
    xp, lx = sequence_pad(tmp, x_lod)                                        # [B, Lx, c*D]
    yp, ly = sequence_pad(y, y_lod)                                          # [B, Ly, D]
    B, Lx, Ly = xp.shape[0], xp.shape[1], yp.shape[1]
    full = torch.einsum("bxcd,byd->bcxy", xp.view(B, Lx, c, d), yp)
    valid = (torch.arange(Lx, device=x.device).view(1, 1, Lx, 1) < lx.view(B, 1, 1, 1)) & \
            (torch.arange(Ly, device=x.device).view(1, 1, 1, Ly) < ly.view(B, 1, 1, 1))
    out = full[valid.expand(B, c, Lx, Ly)].unsqueeze(1)
    out_lod = torch.zeros(B + 1, dtype=torch.int64, device=x.device)
    out_lod[1:] = (c * lx * ly).cumsum(0)
    if act:
        out = getattr(torch, act)(out)
    return out, out_lod, tmp


def sequence_topk_avg_pooling(x: torch.Tensor, x_lod, row_lod, col_lod, topks, channel_num: int):
    """For the ``match_matrix_tensor`` layout (pair, channel, row, column): per row and channel the mean of its
    top-k columns for every k of ``topks`` (fewer than k columns: the sum of what is there, still divided by k).
    ``([sum rows, channel_num * len(topks)], row_lod)`` (sequence_topk_avg_pooling_op.h:118-179)."""
    x_lod, row_lod, col_lod = _lod(x_lod, x.device), _lod(row_lod, x.device), _lod(col_lod, x.device)
    lr, lc = row_lod[1:] - row_lod[:-1], col_lod[1:] - col_lod[:-1]
    if bool(((x_lod[1:] - x_lod[:-1]) != channel_num * lr * lc).any()):
        raise ValueError("sequence_topk_avg_pooling: input length != channel_num * rows * columns")
    B, R = lr.numel(), int(row_lod[-1])
    Lc = int(lc.max()) if B else 0
    max_k, ks = int(topks[-1]), torch.as_tensor(list(topks), device=x.device)
    flat = x.reshape(-1)
    # address of (row r of pair b, channel j, column q): x_lod[b] + j * lr*lc + r_local * lc + q
    b_of = _seg(row_lod, R)
    r_loc = _pos(row_lod, b_of)
    base = (x_lod[:-1][b_of] + r_loc * lc[b_of]).view(R, 1, 1) + \
        (torch.arange(channel_num, device=x.device).view(1, -1, 1) * (lr * lc)[b_of].view(R, 1, 1))
    q = torch.arange(max(Lc, 1), device=x.device).view(1, 1, -1)
    inside = q < lc[b_of].view(R, 1, 1)
    vals = torch.where(inside, flat[(base + q).clamp(max=max(flat.numel() - 1, 0))] if flat.numel() else
                       torch.zeros((), dtype=x.dtype, device=x.device), torch.full((), -torch.inf, dtype=x.dtype, device=x.device))
    k_eff = min(max_k, vals.shape[-1])
    top = vals.topk(k_eff, dim=-1).values
    top = torch.where(torch.isinf(top), torch.zeros_like(top), top)          # missing columns add nothing
    run = top.cumsum(-1)
    if k_eff < max_k:
        run = torch.cat([run, run[..., -1:].expand(R, channel_num, max_k - k_eff)], dim=-1)
    out = run[..., ks - 1] / ks.to(x.dtype)
    return out.reshape(R, channel_num * len(topks)), row_lod


# ---- losses -----------------------------------------------------------------------------------------------------
def rank_loss(label: torch.Tensor, left: torch.Tensor, right: torch.Tensor):
    """RankNet: ``log(1 + exp(o)) - label * o`` with ``o = left - right`` (loss.py ``rank_loss``)."""
    o = left - right
    return torch.nn.functional.softplus(o) - label * o


def margin_rank_loss(label: torch.Tensor, left: torch.Tensor, right: torch.Tensor, margin: float = 0.1):
    """``max(0, -label * (left - right) + margin)``."""
    return torch.relu(-label * (left - right) + margin)


def bpr_loss(input: torch.Tensor, label: torch.Tensor):
    """Bayesian personalised ranking: ``-mean_{j != y} log sigmoid(x_y - x_j)`` per row, ``[N, 1]``."""
    y = label.reshape(-1, 1).long()
    pos = input.gather(1, y)
    ls = torch.nn.functional.logsigmoid(pos - input)
    ls = ls.masked_fill(torch.zeros_like(ls, dtype=torch.bool).scatter(1, y, True), 0.0)
    return -ls.sum(1, keepdim=True) / (input.shape[1] - 1)


def log_loss(input: torch.Tensor, label: torch.Tensor, epsilon: float = 1e-4):
    """``-label * log(p + eps) - (1 - label) * log(1 - p + eps)`` on probabilities (nn.py ``log_loss``)."""
    return -label * torch.log(input + epsilon) - (1.0 - label) * torch.log(1.0 - input + epsilon)


class _TeacherStudent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, label, up, low):
        ctx.save_for_backward(x, label)
        ctx.bounds = (up, low)
        ce0 = torch.relu(x) + torch.log1p(torch.exp(-x.abs()))               # click = 0 cross entropy on the logit
        ce1 = ce0 - x
        soft = lambda q: ce0 - x * q                                        # noqa: E731  teacher term
        return torch.where(label < -1.0, ce0, torch.where(label < 0.0, ce1, torch.where(
            label < 1.0, ce0 + soft(label), ce1 + soft(label - 1.0))))

    @staticmethod
    def backward(ctx, dy):
        x, label = ctx.saved_tensors
        up, low = ctx.bounds
        z = x.clamp(low, up)
        pred = torch.sigmoid(z)
        g = torch.where(label < -1.0, -pred, torch.where(label < 0.0, 1.0 - pred, label - 2.0 * pred))
        g = torch.where((z >= up) | (z <= low), torch.zeros_like(g), g)
        return -g * dy, None, None, None


def teacher_student_sigmoid_loss(input: torch.Tensor, label: torch.Tensor, soft_max_up_bound: float = 15.0,
                                 soft_max_lower_bound: float = -15.0):
    """Click loss plus distillation from a teacher score ``q``, both on the student's logit.  ``label``: -2 = no
    teacher, no click; -1 = no teacher, click; ``q`` in [0, 1) = teacher, no click; ``1 + q`` = teacher, click.
    The gradient is the reference's: zero where the logit is outside the soft-max bounds
    (teacher_student_sigmoid_loss_op.h:34-121)."""
    return _TeacherStudent.apply(input, label, float(soft_max_up_bound), float(soft_max_lower_bound))


# ---- hash -------------------------------------------------------------------------------------------------------
_P1, _P2, _P3, _P4, _P5 = (np.uint64(v) for v in (11400714785074694791, 14029467366897019727, 1609587929392839161,
                                                  9650029242287828579, 2870177450012600261))


def _rotl(v, r):
    return (v << np.uint64(r)) | (v >> np.uint64(64 - r))


def _xx_round(acc, lane):
    return _rotl(acc + lane * _P2, 31) * _P1


def xxh64_rows(rows: np.ndarray, seed: int) -> np.ndarray:
    """XXH64 of every row of a C-contiguous 2-D integer array (its bytes, little endian), vectorised over rows."""
    rows = np.ascontiguousarray(rows)
    n, nbytes = rows.shape[0], rows.shape[1] * rows.dtype.itemsize
    raw = rows.view(np.uint8).reshape(n, nbytes)
    seed = np.uint64(seed)
    p = 0
    with np.errstate(over="ignore"):
        if nbytes >= 32:
            v = [np.full(n, seed + _P1 + _P2), np.full(n, seed + _P2), np.full(n, seed), np.full(n, seed - _P1)]
            while p + 32 <= nbytes:
                lanes = raw[:, p:p + 32].copy().view(np.uint64)
                v = [_xx_round(v[i], lanes[:, i]) for i in range(4)]
                p += 32
            h = _rotl(v[0], 1) + _rotl(v[1], 7) + _rotl(v[2], 12) + _rotl(v[3], 18)
            for i in range(4):
                h = (h ^ _xx_round(np.uint64(0), v[i])) * _P1 + _P4
        else:
            h = np.full(n, seed + _P5)
        h = h + np.uint64(nbytes)
        while p + 8 <= nbytes:
            h = _rotl(h ^ _xx_round(np.uint64(0), raw[:, p:p + 8].copy().view(np.uint64)[:, 0]), 27) * _P1 + _P4
            p += 8
        if p + 4 <= nbytes:
            h = _rotl(h ^ (raw[:, p:p + 4].copy().view(np.uint32)[:, 0].astype(np.uint64) * _P1), 23) * _P2 + _P3
            p += 4
        while p < nbytes:
            h = _rotl(h ^ (raw[:, p].astype(np.uint64) * _P5), 11) * _P1
            p += 1
        h ^= h >> np.uint64(33)
        h *= _P2
        h ^= h >> np.uint64(29)
        h *= _P3
        h ^= h >> np.uint64(32)
    return h


def hash(input: torch.Tensor, hash_size: int, num_hash: int = 1):
    """``out[i, k, 0] = XXH64(bytes of row i, seed k) % hash_size`` for an integer ``[N, W]`` tensor
    (hash_op.h:44-73); host arithmetic, the result goes back to the input's device."""
    if input.dtype not in (torch.int32, torch.int64):
        raise TypeError("hash: the input must be int32 or int64")
    rows = input.detach().cpu().reshape(input.shape[0], -1).numpy()
    cols = [(xxh64_rows(rows, k) % np.uint64(hash_size)).astype(np.int64) for k in range(num_hash)]
    out = torch.from_numpy(np.stack(cols, 1)).to(input.dtype).unsqueeze(-1)
    return out.to(input.device)
