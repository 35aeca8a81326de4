"""Dense towers on the tcgen05 GEMM (csrc/cuda/gemm_tcgen05.cu).

``gemm_bf16_tn``  — thin wrapper of the kernel: C = epilogue(alpha * A[M,K] . B[N,K]^T).
``FusedMLP``      — a whole ReLU tower as ONE autograd node: bf16 operands / fp32
                    accumulate, bias+ReLU fused in the GEMM epilogue, ReLU-backward mask
                    fused in the dgrad epilogue, weight gradients accumulated straight into
                    ``param.grad`` with a split-K GEMM, transposed operand copies produced by
                    the epilogues (no transpose passes).
``GroupedMLP``    — E independent towers with the same shapes (MMoE experts; replaces the
                    reference's ``batch_fc`` BatchedGEMM, operators/batch_fc_op.cu:243).

On CPU (no GPU in the process) the same modules run in plain fp32 PyTorch so that the
framework's plumbing is testable anywhere; on a GPU box the CUDA path is mandatory.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch
import torch.nn as nn

from .. import _native as N


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def gemm_impl() -> int:
    """0 = tcgen05 (default), 1 = SIMT cross-check kernel (PBX_GEMM_IMPL=simt)."""
    return 1 if os.environ.get("PBX_GEMM_IMPL", "tcgen05") == "simt" else 0


def gemm_bf16_tn(A: torch.Tensor, B: torch.Tensor, M: int, Nn: int, K: int, *, bias=None, mask=None, out_bf16=None,
                 out_t=None, out_f32=None, relu=False, accumulate=False, alpha=1.0, splitk=1, block_n=0,
                 impl=None, col_out=None, col_split: int = -1) -> None:
    """A: bf16 [>=M, lda], B: bf16 [>=N, ldb] (row strides taken from the tensors).
    ``col_out`` / ``col_split``: output column ``col_split`` goes to ``col_out[row]`` instead of
    ``out_f32`` (bias gradient when B carries a row of ones)."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and A.is_cuda
    N.check(N.cuda().pbx_gemm_bf16_tn(
        N.ptr(A), A.stride(0), N.ptr(B), B.stride(0), M, Nn, K, N.ptr(bias),
        N.ptr(mask), mask.stride(0) if mask is not None else 0,
        N.ptr(out_bf16), out_bf16.stride(0) if out_bf16 is not None else 0,
        N.ptr(out_t), out_t.stride(0) if out_t is not None else 0,
        N.ptr(out_f32), out_f32.stride(0) if out_f32 is not None else 0,
        int(relu), int(accumulate), C.c_float(alpha), splitk, block_n,
        gemm_impl() if impl is None else impl, N.ptr(col_out), col_split, N.stream_ptr()), "pbx_gemm_bf16_tn")


def gemm_problem(A, B, M, Nn, K, *, bias=None, mask=None, out_bf16=None, out_t=None, out_f32=None, relu=False,
                 accumulate=False, alpha=1.0, splitk=1, col_out=None, col_split: int = -1) -> N.GemmProblem:
    """One entry of a grouped launch: the arguments of :func:`gemm_bf16_tn` as a record."""
    def p(t):
        return t.data_ptr() if t is not None else None

    def ld(t):
        return t.stride(0) if t is not None else 0
    return N.GemmProblem(p(A), p(B), p(bias), p(mask), p(out_bf16), p(out_t), p(out_f32), p(col_out),
                         A.stride(0), B.stride(0), M, Nn, K, ld(mask), ld(out_bf16), ld(out_t), ld(out_f32),
                         int(relu), int(accumulate), splitk, col_split, alpha)


def gemm_bf16_tn_grouped(problems, block_n: int = 64) -> None:
    """Up to 8 independent GEMMs per launch (one shared tile width)."""
    for i0 in range(0, len(problems), 8):
        chunk = problems[i0:i0 + 8]
        arr = (N.GemmProblem * len(chunk))(*chunk)
        N.check(N.cuda().pbx_gemm_bf16_tn_grouped(arr, len(chunk), block_n, N.stream_ptr()),
                "pbx_gemm_bf16_tn_grouped")


def cast_transpose(x: torch.Tensor, out: torch.Tensor | None, out_t: torch.Tensor | None) -> None:
    """fp32 [M, K] -> bf16 ``out`` [M, >=K] and/or bf16 ``out_t`` [K, >=M]."""
    M, K = x.shape
    N.check(N.cuda().pbx_cast_transpose(N.ptr(x), x.stride(0), M, K, N.ptr(out),
                                        out.stride(0) if out is not None else 0, N.ptr(out_t),
                                        out_t.stride(0) if out_t is not None else 0, N.stream_ptr()),
            "pbx_cast_transpose")


CT_MAX = 12


def _wgrad_splitk(B: int) -> int:
    """Split-K factor of the grouped weight-gradient launch (the batch is the reduction axis).
    B // 1024 keeps the whole group within one wave of two CTAs per SM for the CTR towers."""
    env = os.environ.get("PBX_WGRAD_SPLITK")
    if env:
        return max(1, int(env))
    return max(1, min(16, B // 1024))
_SIDE_STREAMS: dict = {}


def _side_stream(device) -> "torch.cuda.Stream":
    """One auxiliary stream per device for work that is off the critical path (weight gradients)."""
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    st = _SIDE_STREAMS.get(key)
    if st is None:
        st = _SIDE_STREAMS[key] = torch.cuda.Stream(device=key)
    return st


def cast_transpose_multi(items) -> None:
    """``items``: (x fp32 [M, K], out bf16 | None, out_t bf16 | None) triples, cast (and transposed)
    in ONE launch per 12 tensors — a training step is bound by the number of dependent launches."""
    for i0 in range(0, len(items), CT_MAX):
        chunk = items[i0:i0 + CT_MAX]
        n = len(chunk)
        xs = (C.c_void_p * n)(*[x.data_ptr() for x, _, _ in chunk])
        outs = (C.c_void_p * n)(*[o.data_ptr() if o is not None else None for _, o, _ in chunk])
        outts = (C.c_void_p * n)(*[t.data_ptr() if t is not None else None for _, _, t in chunk])
        ld = (C.c_int * n)(*[x.stride(0) for x, _, _ in chunk])
        M = (C.c_int * n)(*[x.shape[0] for x, _, _ in chunk])
        K = (C.c_int * n)(*[x.shape[1] for x, _, _ in chunk])
        ldo = (C.c_int * n)(*[o.stride(0) if o is not None else 0 for _, o, _ in chunk])
        ldt = (C.c_int * n)(*[t.stride(0) if t is not None else 0 for _, _, t in chunk])
        N.check(N.cuda().pbx_cast_transpose_multi(n, xs, ld, M, K, outs, ldo, outts, ldt, N.stream_ptr()),
                "pbx_cast_transpose_multi")


def colsum_bf16(x: torch.Tensor, M: int, Nn: int, out: torch.Tensor) -> None:
    N.check(N.cuda().pbx_colsum_bf16(N.ptr(x), x.stride(0), M, Nn, N.ptr(out), N.stream_ptr()), "pbx_colsum_bf16")


class _Workspace:
    """Per-batch-size activation buffers of one tower (bf16, padded leading dims)."""

    def __init__(self, dims, B, device):
        bf = torch.bfloat16
        Bp = _pad8(B)
        self.B = B
        # H[l]: [B, dims[l]] activations (l = 0 is the cast input), Ht[l] their transposes
        self.H = [torch.zeros(B, _pad8(d), dtype=bf, device=device) for d in dims[:-1]]
        # one extra row of ones: the wgrad GEMM then emits the bias gradient as column d
        self.Ht = [torch.zeros(d + 1, Bp, dtype=bf, device=device) for d in dims[:-1]]
        for t in self.Ht:
            t[-1, :B] = 1.0
        # dZ[l]: grad w.r.t. pre-activation of layer l (1-based), and transposes
        self.dZ = [torch.zeros(B, _pad8(d), dtype=bf, device=device) for d in dims[1:]]
        self.dZt = [torch.zeros(d, Bp, dtype=bf, device=device) for d in dims[1:]]


class _MLPFn(torch.autograd.Function):
    """The tower as one node.  With ``fm`` = (S, E, wcol, vcol, D) the node also evaluates the
    DeepFM first/second-order term on the same input and, in backward, adds its gradient into
    the tower's input gradient in place (no zero-fill + add passes for the second consumer)."""

    @staticmethod
    def forward(ctx, x, mod, fm, *params):
        ctx.mod = mod
        ctx.fm = fm
        ctx.x_requires_grad = x.requires_grad
        out = mod._forward_cuda(x, torch.is_grad_enabled() or True)
        if fm is None:
            return out
        S, E, wcol, vcol, D = fm
        xx = x if x.stride(1) == 1 else x.contiguous()
        fm_out = torch.empty(x.shape[0], dtype=torch.float32, device=x.device)
        N.check(N.cuda().pbx_fm_fwd(N.ptr(xx), xx.stride(0), x.shape[0], S, E, wcol, vcol, D, N.ptr(fm_out),
                                    N.stream_ptr()), "pbx_fm_fwd")
        ctx.save_for_backward(xx)
        ctx.set_materialize_grads(False)
        return out, fm_out

    @staticmethod
    def backward(ctx, dout, dfm=None):
        mod = ctx.mod
        nparam = 2 * len(mod.weights)
        need_dx = ctx.x_requires_grad
        if dout is None:                      # only the FM branch contributed
            dout = torch.zeros(mod._cur.B, mod.dims[-1], dtype=torch.float32, device=mod.weights[0].device)
        dx = mod._backward_cuda(dout.contiguous(), need_dx)
        if ctx.fm is not None and dfm is not None and need_dx:
            (xx,) = ctx.saved_tensors
            S, E, wcol, vcol, D = ctx.fm
            N.check(N.cuda().pbx_fm_bwd(N.ptr(xx), xx.stride(0), N.ptr(dfm.contiguous()), xx.shape[0], S, E, wcol, vcol,
                                        D, N.ptr(dx), dx.stride(0), 1, N.stream_ptr()), "pbx_fm_bwd")
        return (dx, None, None) + (None,) * nparam


class _EmuBF16Linear(torch.autograd.Function):
    """CPU statement of the GPU tower's numerics (``FLAGS.pbx_cpu_emulate_bf16``): every GEMM operand —
    activations, weights, output gradients — rounded to bf16, products accumulated in fp32, master weights
    and their gradients fp32.  Lets the bf16-vs-fp32 training comparison (AUC parity) run without a GPU."""

    @staticmethod
    def forward(ctx, x, w, b):
        xb, wb = x.bfloat16().float(), w.bfloat16().float()
        ctx.save_for_backward(xb, wb)
        return xb @ wb.t() + b

    @staticmethod
    def backward(ctx, g):
        xb, wb = ctx.saved_tensors
        gb = g.bfloat16().float()
        return gb @ wb, gb.t() @ xb, gb.sum(0)


class FusedMLP(nn.Module):
    """ReLU tower ``dims[0] -> dims[1] -> ... -> dims[-1]``; the last layer is linear
    unless ``final_relu``.  Parameter names follow the reference's convention
    (``fc_i.w_0`` / ``fc_i.b_0``) so dense checkpoints stay compatible."""

    def __init__(self, dims, final_relu: bool = False, name: str = "fc", seed: int | None = None):
        super().__init__()
        self.dims = list(dims)
        self.final_relu = final_relu
        self.name = name
        g = torch.Generator().manual_seed(seed) if seed is not None else None
        self.weights = nn.ParameterList()
        self.biases = nn.ParameterList()
        for k, n in zip(self.dims[:-1], self.dims[1:]):
            w = torch.empty(n, k)
            bound = 1.0 / math.sqrt(k)
            w.uniform_(-bound, bound, generator=g)
            self.weights.append(nn.Parameter(w))
            self.biases.append(nn.Parameter(torch.zeros(n)))
        self._ws: dict[int, _Workspace] = {}
        self._wb = None          # bf16 weights [N, pad8(K)]
        self._wbt = None         # bf16 transposed weights [K, pad8(N)]
        self._dirty = True
        self._versions = None
        self._cur = None

    # ---- parameter bookkeeping ------------------------------------------------
    def named_dense_params(self):
        for i, (w, b) in enumerate(zip(self.weights, self.biases)):
            yield f"{self.name}_{i}.w_0", w
            yield f"{self.name}_{i}.b_0", b

    def mark_weights_dirty(self) -> None:
        """Call after the dense optimizer changed the fp32 weights outside autograd's
        version tracking (paddlebox_b200 optimizers do)."""
        self._dirty = True

    def _refresh_bf16(self, extra=None) -> None:
        """bf16 (+ transposed) copies of the weights when they changed, and of ``extra`` (the
        tower input triple), all in one launch."""
        vers = tuple(w._version for w in self.weights)
        if not self._dirty and vers == self._versions and self._wb is not None:
            if extra is not None:
                cast_transpose_multi([extra])
            return
        dev = self.weights[0].device
        if self._wb is None:
            self._wb = [torch.zeros(w.shape[0], _pad8(w.shape[1]), dtype=torch.bfloat16, device=dev)
                        for w in self.weights]
            self._wbt = [torch.zeros(w.shape[1], _pad8(w.shape[0]), dtype=torch.bfloat16, device=dev)
                         for w in self.weights]
        items = [(w.data, wb, wbt) for w, wb, wbt in zip(self.weights, self._wb, self._wbt)]
        if extra is not None:
            items.append(extra)
        cast_transpose_multi(items)
        self._dirty = False
        self._versions = vers

    # ---- execution ------------------------------------------------------------
    def forward_with_fm(self, x: torch.Tensor, num_slots: int, slot_width: int, wcol: int, vcol: int, dim: int):
        """(tower(x), fm(x)): the tower and the DeepFM interaction term as ONE autograd node."""
        if not x.is_cuda:
            from . import reference as R
            return self.forward(x), R.fm_fwd(x, num_slots, slot_width, wcol, vcol, dim)
        return _MLPFn.apply(x, self, (num_slots, slot_width, wcol, vcol, dim), *self.weights, *self.biases)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            from ..flags import FLAGS
            lin = _EmuBF16Linear.apply if FLAGS.pbx_cpu_emulate_bf16 else torch.nn.functional.linear
            h = x
            L = len(self.weights)
            for i, (w, b) in enumerate(zip(self.weights, self.biases)):
                h = lin(h, w, b)
                if i < L - 1 or self.final_relu:
                    h = torch.relu(h)
            return h
        return _MLPFn.apply(x, self, None, *self.weights, *self.biases)

    def _forward_cuda(self, x: torch.Tensor, training: bool) -> torch.Tensor:
        B = x.shape[0]
        assert x.shape[1] == self.dims[0] and x.dtype == torch.float32
        ws = self._ws.get(B)
        if ws is None:
            ws = self._ws[B] = _Workspace(self.dims, B, x.device)
        self._cur = ws
        self._refresh_bf16((x.contiguous() if x.stride(1) != 1 else x, ws.H[0], ws.Ht[0]))
        L = len(self.weights)
        out = torch.empty(B, self.dims[-1], dtype=torch.float32, device=x.device)
        for l in range(L):
            last = l == L - 1
            K, Nn = self.dims[l], self.dims[l + 1]
            if last:
                gemm_bf16_tn(ws.H[l], self._wb[l], B, Nn, K, bias=self.biases[l].data, out_f32=out,
                             relu=self.final_relu)
            else:
                gemm_bf16_tn(ws.H[l], self._wb[l], B, Nn, K, bias=self.biases[l].data, out_bf16=ws.H[l + 1],
                             out_t=ws.Ht[l + 1], relu=True)
        self._last_out = out if self.final_relu else None
        return out

    def output_grad_slot(self, B: int):
        """(dZ [B, pad8(1)], dZt [1, pad8(B)]) of the last layer when a loss head may write the
        output gradient straight into them as bf16 (single linear output unit); else None."""
        ws = self._cur
        if ws is None or ws.B != B or self.dims[-1] != 1 or self.final_relu:
            return None
        self._dz_ready = True
        L = len(self.weights)
        return ws.dZ[L - 1], ws.dZt[L - 1]

    def _backward_cuda(self, dout: torch.Tensor, need_dx: bool):
        ws = self._cur
        B = ws.B
        L = len(self.weights)
        if self.final_relu:
            dout = dout * (self._last_out > 0)
        if getattr(self, "_dz_ready", False):
            self._dz_ready = False          # the loss head already wrote the bf16 operands
        else:
            cast_transpose(dout, ws.dZ[L - 1], ws.dZt[L - 1])
        dx = None
        wgrads = []
        from ..flags import FLAGS
        overlap = bool(FLAGS.pbx_overlap_wgrad) and gemm_impl() == 0
        main = torch.cuda.current_stream(dout.device)
        side = _side_stream(dout.device) if overlap else None
        for l in range(L - 1, -1, -1):
            K, Nn = self.dims[l], self.dims[l + 1]
            w, b = self.weights[l], self.biases[l]
            if w.grad is None:
                w.grad = torch.zeros_like(w.data)
            if b.grad is None:
                b.grad = torch.zeros_like(b.data)
            # wgrad: dW[N, K] += dZ^T[N, B] . (H^T[K, B])^T   (split-K over the batch)
            # (the ones row of Ht makes output column K the bias gradient: no separate column-sum pass)
            big = ((Nn + 127) // 128) * ((K + 128) // 128) >= 148
            if overlap:
                # Weight gradients are off the critical path: layer l's runs on the side stream as
                # soon as dZ[l] exists, concurrently with the dgrad chain on the main stream (the
                # tower GEMMs are latency-bound and leave most of the machine idle).
                ev = torch.cuda.Event()
                ev.record(main)
                side.wait_event(ev)
                with torch.cuda.stream(side):
                    gemm_bf16_tn(ws.dZt[l], ws.Ht[l], Nn, K + 1, B, out_f32=w.grad, accumulate=True,
                                 splitk=1 if big else max(1, min(16, B // 512)), col_out=b.grad, col_split=K)
            elif big:
                # already a full wave of output tiles: no split-K (no contended atomics), own
                # launch with the wide 128x256 tile
                gemm_bf16_tn(ws.dZt[l], ws.Ht[l], Nn, K + 1, B, out_f32=w.grad, accumulate=True, splitk=1,
                             col_out=b.grad, col_split=K)
            else:
                # one grouped launch once every dZ exists (two CTAs per SM: one wave)
                wgrads.append(gemm_problem(ws.dZt[l], ws.Ht[l], Nn, K + 1, B, out_f32=w.grad, accumulate=True,
                                           splitk=_wgrad_splitk(B), col_out=b.grad, col_split=K))
            if l > 0:
                # dgrad fused with the ReLU mask of the previous activation
                gemm_bf16_tn(ws.dZ[l], self._wbt[l], B, K, Nn, mask=ws.H[l], out_bf16=ws.dZ[l - 1],
                             out_t=ws.dZt[l - 1])
            elif need_dx:
                # row stride padded to 16 bytes: the epilogue then writes 16-byte vectors (with an
                # odd stride it falls back to 4-byte stores that touch 32 rows per instruction)
                dx = torch.empty(B, (K + 3) // 4 * 4, dtype=torch.float32, device=dout.device)[:, :K]
                gemm_bf16_tn(ws.dZ[0], self._wbt[0], B, K, Nn, out_f32=dx)
        if overlap:
            done = torch.cuda.Event()
            done.record(side)
            self._wgrad_done = done         # joined by the caller after its own tail work
            main.wait_event(done)
        elif not wgrads:
            pass
        elif gemm_impl() == 0:
            gemm_bf16_tn_grouped(wgrads, block_n=64)
        else:   # SIMT cross-check path: one problem at a time
            for q in wgrads:
                N.check(N.cuda().pbx_gemm_bf16_tn(q.A, q.lda, q.B, q.ldb, q.M, q.N, q.K, None, None, 0, None, 0, None, 0,
                                                  C.c_void_p(q.out_f32), q.ld_f32, 0, 1, C.c_float(1.0), 1, 0, 1,
                                                  C.c_void_p(q.col_out), q.col_split, N.stream_ptr()), "pbx_gemm_bf16_tn")
        return dx


class GroupedMLP(nn.Module):
    """``num`` towers of identical shape; outputs stacked as [num, B, out].  Expert towers of
    MMoE / the reference's ``batch_fc`` (operators/batch_fc_op.cu:243).

    With a SHARED input (x [B, in], the MMoE case) the first layers of all towers are one wide
    layer ``in -> num*h1``: one forward GEMM with N = num*h1, one dgrad GEMM with K = num*h1 that
    produces the input gradient already summed over the towers, one wgrad GEMM — instead of
    ``num`` skinny GEMMs and ``num`` input-gradient tensors of B x in floats added by autograd.
    The remaining layers are block-diagonal and run per tower on column slices of the wide
    activation.  With per-tower inputs (x [num, B, in]) the towers are independent."""

    def __init__(self, num: int, dims, final_relu: bool = True, name: str = "expert", seed: int | None = None,
                 shared_input: bool = True, extra_linear: int = 0):
        """``extra_linear`` > 0 (shared input only): that many further LINEAR output columns ride on
        the wide first layer (the MMoE gate logits) — ``forward`` then returns (stacked, extra)."""
        super().__init__()
        self.num, self.dims, self.shared_input, self.extra = num, list(dims), shared_input, extra_linear
        self.first_relu = len(dims) > 2 or final_relu
        if shared_input:
            self.first = FusedMLP([dims[0], num * dims[1] + extra_linear],
                                  final_relu=self.first_relu and extra_linear == 0, name=f"{name}_l0", seed=seed)
            # re-draw each tower's block with the fan-in bound of a single tower layer (same init law)
            rest = dims[1:]
            self.towers = nn.ModuleList([FusedMLP(rest, final_relu=final_relu, name=f"{name}{i}",
                                                  seed=None if seed is None else seed + 1 + i)
                                         for i in range(num)]) if len(rest) > 1 else None
        else:
            self.first = None
            self.towers = nn.ModuleList([FusedMLP(dims, final_relu=final_relu, name=f"{name}{i}",
                                                  seed=None if seed is None else seed + i) for i in range(num)])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.first is None:
            if x.dim() == 3:
                return torch.stack([t(x[i]) for i, t in enumerate(self.towers)], 0)
            return torch.stack([t(x) for t in self.towers], 0)
        assert x.dim() == 2, "shared-input GroupedMLP takes x [B, in]"
        h = self.first(x)                                    # [B, num*h1 (+ extra)]
        extra = None
        if self.extra:
            h, extra = torch.split(h, [self.num * self.dims[1], self.extra], dim=1)
            if self.first_relu:
                h = torch.relu(h)
        parts = torch.split(h, self.dims[1], dim=1)          # backward = one concat, no zero-fills
        if self.towers is None:
            out = torch.stack(parts, 0)
        else:
            out = torch.stack([t(p) for t, p in zip(self.towers, parts)], 0)
        return (out, extra) if self.extra else out

    def dense_towers(self):
        return ([self.first] if self.first is not None else []) + (list(self.towers) if self.towers is not None else [])

    def mark_weights_dirty(self):
        for t in self.dense_towers():
            t.mark_weights_dirty()

    def named_dense_params(self):
        for t in self.dense_towers():
            yield from t.named_dense_params()
