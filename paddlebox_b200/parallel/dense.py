"""Dense-parameter storage, optimizer and data-parallel synchronisation.

All dense parameters of a model alias ONE flat fp32 buffer and all their gradients alias ONE
flat gradient buffer (the reference does the same with ``MemoryShareTensor`` so that the
dense sync is one collective, boxps_worker.cc:373-392, :481-520).  On GPUs the gradient
buffer lives in symmetric memory, so the data-parallel allreduce is the one-kernel,
scale-fused peer-memory collective of ``parallel/symm.py`` with no packing copy
(the reference copies every grad into a fused buffer first, c_mixallgather_op.cc:209-214).

Sync modes (reference ``sync_dense_mode`` / sharding stage, boxps_worker.cc:393-395, :601-770,
:1191-1258):

* ``"allreduce"``  every step: grads summed over ranks, scaled by 1/nranks (default)
* ``"kstep"``      K-step model averaging: no per-step sync, parameters (and optionally Adam
                   moments, FLAGS_enable_sync_dense_moment) averaged every ``sync_weight_step``
* ``"sharding"``   ZeRO-1: reduce-scatter grads, each rank updates its slice, all-gather params
* ``"async"``      host-resident parameters updated by a background Adam thread
                   (``host.AsyncDenseTable``)
* ``"none"``       single process

``FLAGS.dense_use_nccl`` routes the collectives through ``torch.distributed`` (NCCL) instead —
the labelled baseline path, also used on CPU (gloo).
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable

import torch
import torch.nn as nn

from .. import _native as N
from ..flags import FLAGS


class DenseParamStore:
    def __init__(self, module: nn.Module, rank: int = 0, nranks: int = 1, group=None, use_symm: bool | None = None,
                 with_moments: bool = True, topo=None):
        self.module = module
        self.rank, self.nranks, self.group = rank, nranks, group
        # more than one node (parallel/topology.py): the symmetric buffer spans this node's GPUs only and
        # the collectives get an inter-node leg on each rank's shard
        self.topo = topo if (topo is not None and topo.multi_node) else None
        self.lrank, self.lsize, self.lgroup = (topo.local_rank, topo.local_size, topo.intra_group) \
            if self.topo else (rank, nranks, group)
        params = [p for p in module.parameters() if p.requires_grad]
        self.params = params
        self.device = params[0].device if params else torch.device("cpu")
        self.cuda = self.device.type == "cuda"
        # 16-byte aligned segments so that every tensor can be a vectorised collective operand
        self.offsets, off = [], 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + 3) // 4 * 4
        self.numel = max(off, 4)
        self.use_symm = (self.cuda and self.lsize > 1 and not FLAGS.dense_use_nccl) if use_symm is None else use_symm
        self.flat = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
        self.symm = None
        if self.use_symm:
            from .symm import SymmetricBuffer
            # [grads | params] : the params half is only used by sharding / k-step collectives
            self.symm = SymmetricBuffer(2 * self.numel, self.lrank, self.lsize, group=self.lgroup, device=self.device)
            self.gflat = self.symm.tensor[:self.numel]
            self.pscratch = self.symm.tensor[self.numel:]
        else:
            self.gflat = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            self.pscratch = None
        for p, o in zip(params, self.offsets):
            n = p.numel()
            self.flat[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat[o:o + n].view_as(p)
            p.grad = self.gflat[o:o + n].view_as(p)
        self.m = torch.zeros_like(self.flat) if with_moments else None
        self.v = torch.zeros_like(self.flat) if with_moments else None
        self.step_dev = torch.zeros(2, dtype=torch.float32, device=self.device)

    # ---- collectives --------------------------------------------------------------------
    def allreduce_grads(self, scale: float | None = None, stream=None) -> None:
        """grads <- sum over ranks * scale (default 1/nranks)."""
        if self.nranks == 1:
            return
        scale = 1.0 / self.nranks if scale is None else scale
        if self.topo:
            self.topo.hierarchical_allreduce(self.gflat, scale, self.symm, 0)
        elif self.use_symm:
            self.symm.allreduce(0, self.numel, scale, stream=stream)
        else:
            import torch.distributed as dist
            dist.all_reduce(self.gflat, group=self.group)
            self.gflat.mul_(scale)

    def average_params(self, with_moments: bool = False) -> None:
        """K-step model averaging (SyncParam, boxps_worker.cc:1191-1258)."""
        if self.nranks == 1:
            return
        bufs = [self.flat] + ([self.m, self.v] if with_moments and self.m is not None else [])
        for b in bufs:
            if self.topo:
                if self.use_symm:
                    self.pscratch.copy_(b)
                    self.topo.hierarchical_allreduce(self.pscratch, 1.0 / self.nranks, self.symm, self.numel)
                    b.copy_(self.pscratch)
                else:
                    self.topo.hierarchical_allreduce(b, 1.0 / self.nranks)
            elif self.use_symm:
                self.pscratch.copy_(b)
                self.symm.allreduce(self.numel, self.numel, 1.0 / self.nranks)
                b.copy_(self.pscratch)
            else:
                import torch.distributed as dist
                dist.all_reduce(b, group=self.group)
                b.mul_(1.0 / self.nranks)

    def shard_range(self):
        """Slice this rank's optimizer owns in sharding mode: 1/nranks of the buffer on one node; with
        several nodes 1/gpus_per_node (every node holds a full replica of the sharded state)."""
        chunk = ((self.numel + self.lsize - 1) // self.lsize + 3) // 4 * 4
        b = min(self.numel, self.lrank * chunk)
        return b, min(self.numel, b + chunk)

    def reduce_scatter_grads(self) -> None:
        if self.nranks == 1:
            return
        if self.topo:
            import torch.distributed as dist
            b, e = self.shard_range()
            if self.use_symm:
                self.symm.collective("reduce_scatter", 0, self.numel, 1.0 / self.nranks)
            else:
                dist.all_reduce(self.gflat, group=self.lgroup)
                self.gflat.mul_(1.0 / self.nranks)
            self.topo.sync_dense(self.gflat[b:e])            # inter-node leg on the owned shard only
        elif self.use_symm:
            self.symm.collective("reduce_scatter", 0, self.numel, 1.0 / self.nranks)
        else:
            import torch.distributed as dist
            dist.all_reduce(self.gflat, group=self.group)
            self.gflat.mul_(1.0 / self.nranks)

    def all_gather_params(self) -> None:
        if self.nranks == 1:
            return
        if self.use_symm:
            b, e = self.shard_range()
            self.pscratch[b:e].copy_(self.flat[b:e])
            self.symm.collective("all_gather", self.numel, self.numel, 1.0)
            self.flat.copy_(self.pscratch)
        else:
            import torch.distributed as dist
            b, e = self.shard_range()
            mask = torch.zeros_like(self.flat)
            mask[b:e] = self.flat[b:e]
            dist.all_reduce(mask, group=self.lgroup)
            self.flat.copy_(mask)

    def broadcast_params(self, root: int = 0) -> None:
        if self.nranks == 1:
            return
        import torch.distributed as dist
        dist.broadcast(self.flat, src=root, group=self.group)

    def zero_grad(self) -> None:
        self.gflat.zero_()

    def mark_dirty(self) -> None:
        fn = getattr(self.module, "mark_weights_dirty", None)
        if fn:
            fn()


class DenseOptimizer:
    """Fused flat Adam / SGD over a ``DenseParamStore`` (one kernel for every parameter;
    the reference runs one ``adam`` op per parameter)."""

    def __init__(self, store: DenseParamStore, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, kind: str = "adam"):
        self.store, self.lr, self.betas, self.eps, self.weight_decay, self.kind = store, lr, betas, eps, \
            weight_decay, kind
        self.t = 0

    def step(self, grad_scale: float = 1.0, zero_grad: bool = True, lo: int = 0, hi: int | None = None) -> None:
        s = self.store
        hi = s.numel if hi is None else hi
        n = hi - lo
        if n <= 0:
            return
        self.t += 1
        if s.cuda:
            lib = N.cuda()
            if self.kind == "adam":
                N.check(lib.pbx_adam_flat(N.ptr(s.flat[lo:]), N.ptr(s.gflat[lo:]), N.ptr(s.m[lo:]), N.ptr(s.v[lo:]),
                                          C.c_int64(n), C.c_float(self.lr), C.c_float(self.betas[0]),
                                          C.c_float(self.betas[1]), C.c_float(self.eps),
                                          C.c_float(self.weight_decay), C.c_float(grad_scale), N.ptr(s.step_dev),
                                          int(zero_grad), N.stream_ptr()), "pbx_adam_flat")
            else:
                N.check(lib.pbx_sgd_flat(N.ptr(s.flat[lo:]), N.ptr(s.gflat[lo:]), C.c_int64(n), C.c_float(self.lr),
                                         C.c_float(grad_scale), int(zero_grad), N.stream_ptr()), "pbx_sgd_flat")
        else:
            g = s.gflat[lo:hi] * grad_scale
            p = s.flat[lo:hi]
            if self.kind == "adam":
                b1, b2 = self.betas
                if self.weight_decay:
                    g = g + self.weight_decay * p
                s.m[lo:hi].mul_(b1).add_(g, alpha=1 - b1)
                s.v[lo:hi].mul_(b2).addcmul_(g, g, value=1 - b2)
                c1, c2 = 1 - b1 ** self.t, 1 - b2 ** self.t
                p.sub_(self.lr / c1 * s.m[lo:hi] / ((s.v[lo:hi] / c2).sqrt() + self.eps))
            else:
                p.sub_(self.lr * g)
            if zero_grad:
                s.gflat[lo:hi].zero_()
        s.mark_dirty()


class OverlappedGradReducer:
    """Bucketed dense-gradient allreduce that runs UNDER backward.

    The reference reduces the fused dense buffer once per step after the whole backward
    (boxps_worker.cc:1191-1258; c_mixallgather_op.cc:209-328 packs, then reduces).  For a tower with hundreds
    of MB of parameters that leaves the whole transfer exposed.  Here the flat gradient buffer is cut into
    buckets of consecutive parameters (taken in reverse registration order, the order backward produces
    them); a post-accumulate hook on every parameter counts its bucket down, and a bucket whose gradients
    are all final is reduced at once on a side stream while autograd keeps producing the earlier layers.
    Buckets start in a fixed order on every rank (bucket i never before bucket i-1), so the ranks issue the
    same collectives in the same order whatever the hook timing.

    Transport per bucket: the scale-fused one-kernel collective over symmetric memory when the store has
    one (``symm.allreduce(off, n, scale, stream=side)``), else ``torch.distributed.all_reduce(async_op=True)``
    (NCCL on GPUs, gloo on CPU) — optionally on a bf16 copy of the bucket (``cast_bf16``: half the bytes;
    cast back and scale in one pass)."""

    def __init__(self, store: "DenseParamStore", bucket_bytes: int, cast_bf16: bool = False):
        self.store, self.cast = store, bool(cast_bf16)
        self.scale = 1.0 / store.nranks
        self.side = torch.cuda.Stream(device=store.device) if store.cuda else None
        # buckets over the flat buffer: [lo, hi) element ranges of consecutive parameters, last parameter first
        self.buckets: list[dict] = []
        cur_hi, cur_lo, cnt = None, None, 0
        ends = [o + (p.numel() + 3) // 4 * 4 for p, o in zip(store.params, store.offsets)]
        self.bucket_of = {}
        for i in range(len(store.params) - 1, -1, -1):
            if cur_hi is None:
                cur_hi = ends[i]
            cur_lo = store.offsets[i]
            cnt += 1
            self.bucket_of[id(store.params[i])] = len(self.buckets)
            if (cur_hi - cur_lo) * 4 >= bucket_bytes or i == 0:
                self.buckets.append({"lo": cur_lo, "hi": cur_hi, "n_params": cnt, "pending": cnt})
                cur_hi, cnt = None, 0
        self.next = 0
        self.inflight: list = []
        self.launched_in_backward = 0          # buckets started from hooks in the last step (overlap evidence)
        self.handles = [p.register_post_accumulate_grad_hook(self._hook) for p in store.params]
        if self.cast:
            self.bf = torch.empty(store.numel, dtype=torch.bfloat16, device=store.device)

    def remove(self) -> None:
        for h in self.handles:
            h.remove()
        self.handles = []

    def begin_step(self) -> None:
        for b in self.buckets:
            b["pending"] = b["n_params"]
        self.next = 0
        self.inflight.clear()
        self.launched_in_backward = 0

    def _hook(self, p) -> None:
        b = self.buckets[self.bucket_of[id(p)]]
        b["pending"] -= 1
        started = self._start_ready()
        self.launched_in_backward += started

    def _start_ready(self, force: bool = False) -> int:
        n = 0
        while self.next < len(self.buckets) and (force or self.buckets[self.next]["pending"] <= 0):
            self._launch(self.buckets[self.next])
            self.next += 1
            n += 1
        return n

    def _launch(self, b: dict) -> None:
        s = self.store
        lo, n = b["lo"], b["hi"] - b["lo"]
        view = s.gflat[lo:lo + n]
        if s.cuda:
            self.side.wait_stream(torch.cuda.current_stream(s.device))      # the bucket's gradients are final
        if s.use_symm and s.topo is None and not self.cast:
            s.symm.allreduce(lo, n, self.scale, stream=self.side)
            self.inflight.append((None, view, None))
            return
        import torch.distributed as dist
        ctx = torch.cuda.stream(self.side) if s.cuda else _NullCtx()
        with ctx:
            if self.cast:
                t = self.bf[lo:lo + n]
                t.copy_(view)
                w = dist.all_reduce(t, group=s.group, async_op=True)
            else:
                t = None
                w = dist.all_reduce(view, group=s.group, async_op=True)
        self.inflight.append((w, view, t))

    def finish(self) -> None:
        """After backward: start what no hook started (parameters without a gradient this step), then make
        the training stream wait for every bucket; cast back / scale."""
        s = self.store
        self._start_ready(force=True)
        ctx = torch.cuda.stream(self.side) if s.cuda else _NullCtx()
        with ctx:
            for w, view, t in self.inflight:
                if w is None:
                    continue
                w.wait()
                if t is not None:
                    torch.mul(t, self.scale, out=view)         # bf16 -> fp32 and the 1/nranks scale in one pass
                else:
                    view.mul_(self.scale)
        if s.cuda:
            torch.cuda.current_stream(s.device).wait_stream(self.side)
        self.inflight.clear()


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class DenseSync:
    """One object per trainer that applies the configured sync mode around the optimizer step."""

    def __init__(self, store: DenseParamStore, opt: DenseOptimizer, mode: str = "allreduce",
                 sync_weight_step: int = 1, async_table=None):
        if store.nranks == 1 and mode in ("allreduce", "kstep", "sharding"):
            mode = "none"
        self.store, self.opt, self.mode, self.k = store, opt, mode, max(1, sync_weight_step)
        self.async_table = async_table
        self.steps = 0
        self.fold_engine = None
        self.overlap = None
        if mode == "allreduce" and FLAGS.pbx_dense_overlap_mb > 0:
            self.overlap = OverlappedGradReducer(store, int(FLAGS.pbx_dense_overlap_mb * (1 << 20)),
                                                 FLAGS.pbx_dense_overlap_bf16)
        if mode == "async":
            from ..host import AsyncDenseTable
            self.async_table = async_table or AsyncDenseTable(store.numel, lr=opt.lr)
            if store.rank == 0 or async_table is None:
                self.async_table.set_params(store.flat)

    def fold_into(self, engine) -> bool:
        """Let the sparse engine's push kernel carry the per-step gradient allreduce (it runs after
        every dense gradient of the step is final; see csrc/cuda/sparse_a2a.cu).  Only for the
        plain allreduce mode over symmetric memory with the same ranks as the sparse group."""
        s = self.store
        if (self.mode != "allreduce" or not s.use_symm or s.topo is not None or engine is None
                or not FLAGS.pbx_fold_dense_allreduce or self.overlap is not None
                or getattr(engine, "nranks", 1) != s.nranks):
            return False
        self._fold_desc = (s.symm, 0, s.numel, 1.0 / s.nranks)
        engine.fold_dense = self._fold_desc
        engine.fold_done = False
        self.fold_engine = engine
        return True

    def before_forward(self) -> None:
        if self.overlap is not None:
            self.overlap.begin_step()
        if self.fold_engine is not None:
            # several trainers may share one sparse engine (join / update models): the push of THIS
            # step must reduce THIS trainer's gradient buffer
            self.fold_engine.fold_dense = self._fold_desc
            self.fold_engine.fold_done = False
            self.fold_engine.pulls_this_step = 0
        if self.mode == "async":
            # PullDense: whole parameter vector host -> device (boxps_worker.cc:1300-1302)
            self.async_table.pull(out=self.store.flat)
            self.store.mark_dirty()

    def after_backward(self) -> None:
        s, o = self.store, self.opt
        self.steps += 1
        if FLAGS.enable_dense_nccl_barrier and s.nranks > 1 and self.mode in ("allreduce", "kstep", "sharding"):
            import torch.distributed as dist                # FLAGS_enable_dense_nccl_barrier: ranks meet first
            dist.barrier(group=s.group)
        if self.mode == "none":
            o.step()
        elif self.mode == "allreduce":
            eng = self.fold_engine
            if self.overlap is not None:
                self.overlap.finish()          # buckets were started from autograd hooks during backward
            elif eng is not None and getattr(eng, "fold_done", False):
                eng.fold_done = False          # the sparse push of this step already reduced the grads
            else:
                s.allreduce_grads()
            o.step()
        elif self.mode == "kstep":
            o.step()
            if self.steps % self.k == 0:
                s.average_params(FLAGS.enable_sync_dense_moment)
                s.mark_dirty()
        elif self.mode == "sharding":
            s.reduce_scatter_grads()
            b, e = s.shard_range()
            o.step(lo=b, hi=e, zero_grad=False)
            s.zero_grad()
            s.all_gather_params()
            s.mark_dirty()
        elif self.mode == "async":
            self.async_table.push(s.gflat)      # PushDense (D2H of the gradient vector)
            s.zero_grad()
        else:
            raise ValueError(f"unknown dense sync mode {self.mode!r}")

    def finalize(self) -> None:
        """End of pass: make replicas identical again (boxps_worker.cc:1347-1356)."""
        if self.mode == "kstep":
            self.store.average_params(FLAGS.enable_sync_dense_moment)
            self.store.mark_dirty()
        elif self.mode == "async":
            self.async_table.wait()
            self.async_table.pull(out=self.store.flat)
            if self.store.nranks > 1:
                # one process per GPU means one host table per rank (the reference shares one between the
                # GPU threads of its single process): the replicas are averaged at the end of every pass
                self.store.average_params(False)
                self.async_table.set_params(self.store.flat)
            self.store.mark_dirty()
