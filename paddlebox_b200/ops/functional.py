"""Differentiable CTR building blocks backed by csrc/cuda/dense_ops.cu (CUDA tensors) or
the PyTorch oracles in ``ops/reference.py`` (CPU tensors).

  continuous_value_model  — reference ``cvm`` op (operators/cvm_op.cu:29-75; layers/nn.py:14947)
  data_norm / DataNorm    — reference ``data_norm`` op (operators/data_norm_op.cu:39-265)
  fm_interaction          — DeepFM first + second order head
  sigmoid_log_loss        — sigmoid + log_loss (mean) with the gradient produced in the same pass
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from .. import _native as N
from . import reference as R


# ------------------------------------------------------------------ cvm
class _CVM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cvm, use_cvm):
        ctx.save_for_backward(cvm)
        ctx.meta = (x.shape[1], use_cvm)
        if not x.is_cuda:
            return R.cvm_fwd(x, use_cvm)
        x = x.contiguous()
        rows, W = x.shape
        y = torch.empty(rows, W if use_cvm else W - 2, dtype=x.dtype, device=x.device)
        N.check(N.cuda().pbx_cvm_fwd(N.ptr(x), C.c_int64(rows), W, int(use_cvm), N.ptr(y), N.stream_ptr()))
        return y

    @staticmethod
    def backward(ctx, dy):
        (cvm,) = ctx.saved_tensors
        W, use_cvm = ctx.meta
        if not dy.is_cuda:
            return R.cvm_bwd(dy, cvm, W, use_cvm), None, None
        dy = dy.contiguous()
        dx = torch.empty(dy.shape[0], W, dtype=dy.dtype, device=dy.device)
        N.check(N.cuda().pbx_cvm_bwd(N.ptr(dy), N.ptr(cvm.contiguous()), C.c_int64(dy.shape[0]), W, int(use_cvm),
                                     N.ptr(dx), N.stream_ptr()))
        return dx, None, None


def continuous_value_model(x: torch.Tensor, cvm: torch.Tensor, use_cvm: bool = True) -> torch.Tensor:
    """x: [rows, W] with (show, click) in columns 0..1; cvm: [rows, 2] show/click labels
    that replace the first two gradient columns in backward."""
    return _CVM.apply(x, cvm, use_cvm)


# ------------------------------------------------------------------ data_norm
class _DataNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bsize, bsum, bsq, eps, stat_sink):
        if not x.is_cuda:
            y, mean, scale = R.data_norm_fwd(x, bsize, bsum, bsq)
        else:
            x = x.contiguous()
            rows, Cc = x.shape
            y = torch.empty_like(x)
            mean = torch.empty(Cc, dtype=torch.float32, device=x.device)
            scale = torch.empty(Cc, dtype=torch.float32, device=x.device)
            N.check(N.cuda().pbx_data_norm_fwd(N.ptr(x), C.c_int64(rows), Cc, N.ptr(bsize), N.ptr(bsum), N.ptr(bsq),
                                               N.ptr(y), N.ptr(mean), N.ptr(scale), N.stream_ptr()))
        ctx.save_for_backward(x, mean, scale)
        ctx.eps = eps
        ctx.sink = stat_sink
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mean, scale = ctx.saved_tensors
        if not dy.is_cuda:
            dx, d_size, d_sum, d_sq = R.data_norm_bwd(x, dy, mean, scale, ctx.eps)
        else:
            dy = dy.contiguous()
            rows, Cc = x.shape
            dx = torch.empty_like(x)
            stats = torch.zeros(3, Cc, dtype=torch.float32, device=x.device)
            N.check(N.cuda().pbx_data_norm_bwd(N.ptr(x), N.ptr(dy), C.c_int64(rows), Cc, N.ptr(mean), N.ptr(scale),
                                               C.c_float(ctx.eps), N.ptr(dx), N.ptr(stats[0]), N.ptr(stats[1]),
                                               N.ptr(stats[2]), N.stream_ptr()))
            d_size, d_sum, d_sq = stats[0], stats[1], stats[2]
        if ctx.sink is not None:
            ctx.sink(d_size, d_sum, d_sq)
        return dx, None, None, None, None, None


class _DataNormSlot(torch.autograd.Function):
    """data_norm with ``slot_dim``: the input row is a run of slot blocks of ``slot_dim`` columns whose first
    column is the slot's show count; a block with show == 0 (|x| < 1e-7) yields zeros, receives no gradient
    and does not count in the batch statistics (CPU kernels data_norm_op.cc:334-362, :592-720).  Returns
    ``(y, mask)``."""

    @staticmethod
    def forward(ctx, x, bsize, bsum, bsq, eps, stat_sink, slot_dim):
        rows, Cc = x.shape
        mean = bsum / bsize
        scale = torch.sqrt(bsize / bsq)
        show = x.reshape(rows, Cc // slot_dim, slot_dim)[:, :, 0]
        live = (show.abs() >= 1e-7)                                            # [rows, blocks]
        mask = live[:, :, None].expand(rows, Cc // slot_dim, slot_dim).reshape(rows, Cc).to(x.dtype)
        y = (x - mean) * scale * mask
        ctx.save_for_backward(x, mask, mean, scale)
        ctx.eps, ctx.sink = eps, stat_sink
        ctx.mark_non_differentiable(mask)
        return y, mask

    @staticmethod
    def backward(ctx, dy, _dmask):
        x, mask, mean, scale = ctx.saved_tensors
        if ctx.sink is not None:
            cnt = mask.sum(0)
            has = (cnt >= 1).to(x.dtype)
            denom = cnt.clamp(min=1.0)
            d_sum = (x * mask).sum(0) / denom * has
            d_sq = ((((x - mean) ** 2) * mask).sum(0) / denom + cnt * ctx.eps) * has
            ctx.sink(has, d_sum, d_sq)
        return dy * scale * mask, None, None, None, None, None, None


class DataNorm(nn.Module):
    """``y = (x - batch_sum/batch_size) * sqrt(batch_size/batch_square_sum)``.

    The three summaries are *not* trained by the dense optimizer: their "gradients" are
    the batch statistics and they are updated with a decay
    (``summary = summary * decay + stat``, data_norm_op.cu:90-110).  With
    ``sync_stats`` the statistics are first averaged over the data-parallel group
    (reference: 3 ncclAllReduce calls, data_norm_op.cu:203-231)."""

    def __init__(self, num_features: int, epsilon: float = 1e-4, summary_decay_rate: float = 0.9999999,
                 sync_stats: bool = False, name: str = "data_norm", slot_dim: int = -1,
                 enable_scale_and_shift: bool = False):
        """``slot_dim`` > 0: show-aware normalisation per slot block (see ``_DataNormSlot``);
        ``enable_scale_and_shift``: trainable ``scale_w`` / ``bias`` applied after the normalisation
        (``y * scale_w + bias``; data_norm_op.cc:369-386) — both are options of the reference layer
        (layers/nn.py:3490-3504)."""
        super().__init__()
        self.name = name
        self.epsilon = epsilon
        self.decay = summary_decay_rate
        self.sync_stats = sync_stats
        self.slot_dim = slot_dim
        if slot_dim > 0 and num_features % slot_dim:
            raise ValueError(f"num_features {num_features} is not a multiple of slot_dim {slot_dim}")
        self.scale_w = nn.Parameter(torch.ones(num_features)) if enable_scale_and_shift else None
        self.bias = nn.Parameter(torch.zeros(num_features)) if enable_scale_and_shift else None
        self.register_buffer("batch_size", torch.full((num_features,), 1e4))
        self.register_buffer("batch_sum", torch.zeros(num_features))
        self.register_buffer("batch_square_sum", torch.full((num_features,), 1e4))
        self.reduce_fn = None  # set by the trainer: f(tensor[3, C]) -> averaged in place
        self.stat_sync = None  # set by the trainer on GPUs: f(module, d_size, d_sum, d_sq) = fused reduce + update

    def _apply_stats(self, d_size, d_sum, d_sq):
        with torch.no_grad():
            if self.sync_stats and self.stat_sync is not None:
                self.stat_sync(self, d_size, d_sum, d_sq)       # one kernel: allreduce of the 3 vectors + decay update
                return
            stats = torch.stack([d_size, d_sum, d_sq])
            if self.sync_stats and self.reduce_fn is not None:
                stats = self.reduce_fn(stats)
            self.batch_size.mul_(self.decay).add_(stats[0])
            self.batch_sum.mul_(self.decay).add_(stats[1])
            self.batch_square_sum.mul_(self.decay).add_(stats[2])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sink = self._apply_stats if self.training else None
        if self.slot_dim > 0:
            y, mask = _DataNormSlot.apply(x, self.batch_size, self.batch_sum, self.batch_square_sum, self.epsilon,
                                          sink, self.slot_dim)
            return y if self.scale_w is None else (y * self.scale_w + self.bias) * mask
        y = _DataNorm.apply(x, self.batch_size, self.batch_sum, self.batch_square_sum, self.epsilon, sink)
        return y if self.scale_w is None else y * self.scale_w + self.bias


def data_norm(x, batch_size, batch_sum, batch_square_sum, epsilon: float = 1e-4, stat_sink=None):
    return _DataNorm.apply(x, batch_size, batch_sum, batch_square_sum, epsilon, stat_sink)


# ------------------------------------------------------------------ FM head
class _FM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, S, E, wcol, vcol, D):
        ctx.meta = (S, E, wcol, vcol, D)
        ctx.save_for_backward(x)
        B = x.shape[0]
        out = torch.empty(B, dtype=torch.float32, device=x.device)
        N.check(N.cuda().pbx_fm_fwd(N.ptr(x), x.stride(0), B, S, E, wcol, vcol, D, N.ptr(out), N.stream_ptr()))
        return out

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        S, E, wcol, vcol, D = ctx.meta
        dx = torch.zeros_like(x)
        N.check(N.cuda().pbx_fm_bwd(N.ptr(x), x.stride(0), N.ptr(g.contiguous()), x.shape[0], S, E, wcol, vcol, D,
                                    N.ptr(dx), dx.stride(0), 0, N.stream_ptr()))
        return dx, None, None, None, None, None


def fm_interaction(x: torch.Tensor, num_slots: int, slot_width: int, wcol: int, vcol: int, dim: int) -> torch.Tensor:
    """DeepFM head over pooled slot embeddings ``x`` [B, >= S*E]:
    ``sum_s w_s + 0.5 * sum_d ((sum_s v_sd)^2 - sum_s v_sd^2)`` -> [B]."""
    if not x.is_cuda:
        return R.fm_fwd(x, num_slots, slot_width, wcol, vcol, dim)
    assert x.stride(1) == 1
    return _FM.apply(x, num_slots, slot_width, wcol, vcol, dim)


# ------------------------------------------------------------------ loss
class _SigmoidLogLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logit, label):
        B = logit.numel()
        logit = logit.contiguous().view(-1)
        label = label.contiguous().view(-1).float()
        pred = torch.empty(B, dtype=torch.float32, device=logit.device)
        loss = torch.zeros((), dtype=torch.float32, device=logit.device)
        dlogit = torch.empty(B, dtype=torch.float32, device=logit.device)
        N.check(N.cuda().pbx_sigmoid_logloss(N.ptr(logit), N.ptr(label), B, N.ptr(pred), N.ptr(loss), N.ptr(dlogit),
                                             C.c_float(1.0 / B), N.stream_ptr()))
        ctx.save_for_backward(dlogit)
        ctx.mark_non_differentiable(pred)
        return loss, pred

    @staticmethod
    def backward(ctx, gloss, gpred):
        (dlogit,) = ctx.saved_tensors
        return dlogit * gloss, None


def sigmoid_log_loss(logit: torch.Tensor, label: torch.Tensor):
    """(mean log-loss, sigmoid(logit)); the logit gradient is produced in the forward kernel."""
    if not logit.is_cuda:
        z = logit.view(-1)
        y = label.view(-1).float()
        loss = torch.nn.functional.binary_cross_entropy_with_logits(z, y)
        return loss, torch.sigmoid(z).detach()
    return _SigmoidLogLoss.apply(logit, label)


class _CtrHead(torch.autograd.Function):
    """logit = tower + extra + bias -> (mean log-loss, pred) as one kernel forward, one backward.
    The bias gradient is added straight into ``bias.grad`` when that exists (as the towers do
    with their weights), so autograd launches no accumulation kernel for it."""

    @staticmethod
    def forward(ctx, tower, extra, bias, label, tower_mod=None):
        B = tower.shape[0]
        assert tower.dim() in (1, 2) and (tower.dim() == 1 or tower.shape[1] == 1)
        ctx.tower_mod = tower_mod
        dev = tower.device
        pred = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dlogit = torch.empty(B, dtype=torch.float32, device=dev)
        if extra is not None:
            extra = extra.contiguous()
        assert label.dtype == torch.float32
        N.check(N.cuda().pbx_ctr_head_fwd(N.ptr(tower), tower.stride(0), N.ptr(extra), N.ptr(bias), N.ptr(label),
                                          label.stride(0), B, N.ptr(pred), N.ptr(loss), N.ptr(dlogit),
                                          N.stream_ptr()), "pbx_ctr_head_fwd")
        ctx.save_for_backward(dlogit)
        ctx.shape = tuple(tower.shape)
        ctx.has_extra = extra is not None
        ctx.bias = bias if (bias is not None and bias.requires_grad and bias.is_leaf) else None
        ctx.bias_nonleaf = bias is not None and ctx.bias is None
        ctx.mark_non_differentiable(pred)
        ctx.set_materialize_grads(False)
        return loss, pred

    @staticmethod
    def backward(ctx, gloss, gpred):
        (dlogit,) = ctx.saved_tensors
        B = dlogit.numel()
        dev = dlogit.device
        if gloss is None:
            return None, None, None, None, None
        need_t = ctx.needs_input_grad[0]
        need_e = ctx.needs_input_grad[1] and ctx.has_extra
        d_tower = torch.empty(ctx.shape, dtype=torch.float32, device=dev) if need_t else None
        d_extra = torch.empty(B, dtype=torch.float32, device=dev) if need_e else None
        d_bias_ret, d_bias_ptr, acc = None, None, 0
        if ctx.bias is not None:
            if ctx.bias.grad is not None:
                d_bias_ptr, acc = ctx.bias.grad, 1          # in place: nothing returned to autograd
            else:
                d_bias_ret = d_bias_ptr = torch.empty(1, dtype=torch.float32, device=dev)
        elif ctx.bias_nonleaf and ctx.needs_input_grad[2]:
            d_bias_ret = d_bias_ptr = torch.empty(1, dtype=torch.float32, device=dev)
        # the tower that produced `tower` (a FusedMLP ending in one linear unit) takes its output
        # gradient as bf16 operands: write them here instead of in a cast launch of its own
        dz = dzt = None
        slot = ctx.tower_mod.output_grad_slot(B) if (ctx.tower_mod is not None and need_t) else None
        if slot is not None:
            dz, dzt = slot
        N.check(N.cuda().pbx_ctr_head_bwd(N.ptr(dlogit), N.ptr(gloss.contiguous()), B, N.ptr(d_tower), N.ptr(d_extra),
                                          N.ptr(d_bias_ptr), acc, N.ptr(dz), dz.stride(0) if dz is not None else 0,
                                          N.ptr(dzt), N.stream_ptr()), "pbx_ctr_head_bwd")
        return d_tower, d_extra, d_bias_ret, None, None


def ctr_head(tower: torch.Tensor, label: torch.Tensor, extra: torch.Tensor | None = None,
             bias: torch.Tensor | None = None, tower_mod=None):
    """(mean log-loss, pred) of ``sigmoid(tower + extra + bias)`` against ``label``.

    ``tower`` [B] or [B, 1] (the last tower layer's output), ``extra`` [B] further logit terms
    (FM / wide part), ``bias`` [1] parameter, ``label`` [B] float32 (may be a strided column)."""
    if not tower.is_cuda:
        z = tower.reshape(-1)
        if extra is not None:
            z = z + extra
        if bias is not None:
            z = z + bias
        return sigmoid_log_loss(z, label)
    return _CtrHead.apply(tower, extra, bias, label, tower_mod)


# ------------------------------------------------------------------ misc CTR utilities
def row_gather(table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """Replica-cache gather ``out[i] = table[idx[i]]`` (pull_cache_value, box_wrapper.cu:1210-1224)."""
    if not table.is_cuda:
        return table[idx.long()]
    idx = idx.contiguous().long()
    out = torch.empty(idx.numel(), table.shape[1], dtype=torch.float32, device=table.device)
    N.check(N.cuda().pbx_row_gather(N.ptr(table.contiguous()), N.ptr(idx), C.c_int64(idx.numel()), table.shape[1],
                                    N.ptr(out), N.stream_ptr()))
    return out


# ---- masked_data_norm (operators/masked_data_norm_op.cu:39-307) ---------------------------------
class _MaskedDataNorm(torch.autograd.Function):
    """Rows with ``mask == 0`` produce zeros, receive no gradient and do not contribute to the
    batch statistics; the summary "gradients" follow the GPU kernel: size 1 (0 if no row is
    selected), mean of selected rows, mean squared deviation + epsilon."""

    @staticmethod
    def forward(ctx, x, mask, bsize, bsum, bsq, eps, stat_sink):
        mean = bsum / bsize
        scale = torch.sqrt(bsize / bsq)
        m = mask.reshape(-1, 1).to(x.dtype)
        y = (x - mean) * scale * m
        ctx.save_for_backward(x, m, mean, scale)
        ctx.eps, ctx.sink = eps, stat_sink
        return y

    @staticmethod
    def backward(ctx, dy):
        x, m, mean, scale = ctx.saved_tensors
        dx = dy * scale * m
        if ctx.sink is not None:
            cnt = m.sum()
            has = (cnt > 0).to(x.dtype)
            denom = cnt.clamp(min=1.0)
            d_size = torch.ones_like(mean) * has
            d_sum = (x * m).sum(0) / denom * has
            d_sq = ((((x - mean) ** 2) * m).sum(0) / denom + ctx.eps) * has
            ctx.sink(d_size, d_sum, d_sq)
        return dx, None, None, None, None, None, None


def masked_data_norm(x, mask, batch_size, batch_sum, batch_square_sum, epsilon: float = 1e-4, stat_sink=None):
    """``y[r] = mask[r] ? (x[r] - sum/size) * sqrt(size/square_sum) : 0``."""
    return _MaskedDataNorm.apply(x, mask, batch_size, batch_sum, batch_square_sum, epsilon, stat_sink)


# ---- store_q_value (operators/store_q_value_op.cc:36-96) ----------------------------------------
_Q_STORE: dict = {}


def store_q_value(q_tensors, batch=None, device_id: int | None = None) -> list:
    """Copy the PCOC q tensors of the current batch to the host and park them on the batch
    (and in a per-device store the next pack step reads), like ``BatchGpuPackMgr().store_qvalue``."""
    host = [q.detach().to("cpu", non_blocking=False).contiguous() for q in q_tensors]
    if device_id is None:
        device_id = q_tensors[0].device.index or 0 if q_tensors[0].is_cuda else -1
    _Q_STORE[device_id] = host
    if batch is not None:
        batch.q_values = host
    return host


def stored_q_values(device_id: int = 0) -> list:
    return _Q_STORE.get(device_id, [])
