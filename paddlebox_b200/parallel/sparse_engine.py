"""Sharded sparse pull / push engine (one instance per rank / GPU).

Python face of ``csrc/cuda/sparse_a2a.cu``: the table is sharded by key hash over the
HBM of all ranks; ``pull`` and ``push`` are each ONE cooperative kernel that performs the
dedup, the all-to-all (stores into CUDA-IPC-mapped peer memory over NVLink/NVSwitch), the
hash lookup / fused optimizer update, and the cross-rank barriers — no NCCL call and no
host round trip on the sparse path.

Reference role: ``HeterComm::pull_sparse/push_sparse`` + ``BoxWrapper::PullSparseCaseGPU``
/ ``PushSparseGradCaseGPU`` (heter_ps/heter_comm_inl.h:1477-1825,
fleet/box_wrapper_impl.h:25-522).
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from .. import _native as N
from ..ops.sparse_table import SparseTable

STATUS_MSG = {1: "sparse mailbox overflow: more unique keys for one shard than `cap`; "
                 "rebuild SparseEngine with a larger cap / max_keys",
              2: "cross-rank barrier timed out: a peer rank died or did not launch the same pull/push"}


class SparseEngine:
    def __init__(self, table: SparseTable, max_keys: int, rank: int = 0, nranks: int = 1, cap: int = 0,
                 group=None, connect: bool = True):
        self.table = table
        self.rank, self.nranks = rank, nranks
        self.max_keys = int(max_keys)
        self.group = group
        self._h = C.c_void_p()
        lib = N.cuda()
        with torch.cuda.device(table.device):
            N.check(lib.pbx_sparse_create(C.byref(self._h), table.handle, rank, nranks, C.c_int64(self.max_keys),
                                          C.c_int64(cap)), "pbx_sparse_create")
        cap_, rows_, ps, pd = C.c_int64(), C.c_int64(), C.c_int(), C.c_int()
        N.check(lib.pbx_sparse_geometry(self._h, C.byref(cap_), C.byref(rows_), C.byref(ps), C.byref(pd)))
        self.cap, self.n_unique_rows, self.pull_stride, self.push_stride = cap_.value, rows_.value, ps.value, pd.value
        self._values = N.tensor_from_ptr(lib.pbx_sparse_pull_values(self._h), (self.n_unique_rows, self.pull_stride),
                                         torch.float32, owner=self, device=table.device)
        if nranks > 1 and connect:
            self._connect_ipc()

    # ---- wiring ---------------------------------------------------------
    def _connect_ipc(self) -> None:
        """Exchange CUDA IPC handles through torch.distributed (control plane only)."""
        import torch.distributed as dist
        buf = C.create_string_buffer(64)
        N.check(N.cuda().pbx_sparse_ipc_handle(self._h, buf), "pbx_sparse_ipc_handle")
        mine = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        gathered = [None] * self.nranks
        dist.all_gather_object(gathered, mine.numpy().tobytes(), group=self.group)
        blob = b"".join(gathered)
        N.check(N.cuda().pbx_sparse_connect(self._h, blob), "pbx_sparse_connect")
        dist.barrier(group=self.group)

    @staticmethod
    def connect_local(engines: Sequence["SparseEngine"]) -> None:
        """Wire engines that live in one process (virtual ranks on one GPU)."""
        arr = (C.c_void_p * len(engines))(*[e._h for e in engines])
        for e in engines:
            N.check(N.cuda().pbx_sparse_connect_local(e._h, arr), "pbx_sparse_connect_local")

    def __del__(self):
        try:
            if self._h:
                N.cuda().pbx_sparse_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ---- data plane -------------------------------------------------------
    @property
    def values(self) -> torch.Tensor:
        """[nranks*cap, pull_stride] unique value rows of the last pull (row = restore idx)."""
        return self._values

    def pull(self, keys: torch.Tensor, restore: torch.Tensor | None = None,
             n_dev: torch.Tensor | None = None) -> torch.Tensor:
        """Pull the value rows of ``keys`` (int64, device).  Returns the restore index
        (int32 [n]): occurrence i's values are ``self.values[restore[i]]``.

        ``n_dev`` (int32 device scalar, e.g. ``lod[S*B:]``): the real key count, read by the
        kernel; ``keys`` may then be a fixed-capacity buffer — this is what lets one captured
        CUDA graph serve batches of different lengths."""
        n = keys.numel()
        self.peek_status()
        self.pulls_this_step = getattr(self, "pulls_this_step", 0) + 1
        if restore is None:
            restore = torch.empty(max(n, 1), dtype=torch.int32, device=keys.device)
        N.check(N.cuda().pbx_sparse_pull_ex(self._h, N.ptr(keys), C.c_int64(n), N.ptr(n_dev), N.ptr(restore),
                                            N.stream_ptr()), "pbx_sparse_pull")
        return restore[:n]

    def push_pooled(self, restore: torch.Tensor, dout: torch.Tensor, cvm: torch.Tensor, lod: torch.Tensor, B: int,
                    S: int, Eo: int, cvm_offset: int, dim_off: int, grad_scale: float,
                    ld_dout: int | None = None) -> None:
        """Fused seqpool/CVM backward + merge-by-key + all-to-all + optimizer update.
        ``dout`` [B, >= S*Eo] with unit column stride (row stride ``ld_dout``, default the tensor's),
        ``cvm`` [B, cvm_offset], ``lod`` int32 [S*B+1]."""
        ld = dout.stride(0) if ld_dout is None else ld_dout
        assert dout.stride(1) == 1 and dout.shape[0] == B and dout.shape[1] >= S * Eo and ld >= S * Eo \
            and dout.dtype == torch.float32
        fold = getattr(self, "fold_dense", None)
        # (only when the step has ONE pull/push pair: with several, dense gradients that depend on
        #  a later push's inputs would not be final yet when the first push runs)
        if fold is not None and self.nranks > 1 and getattr(self, "pulls_this_step", 1) <= 1 \
                and not getattr(self, "fold_done", False):
            # the trainer's dense-gradient allreduce rides on this kernel's barriers
            symm, off, n, scale = fold
            N.check(N.cuda().pbx_sparse_push_dense(self._h, N.ptr(restore), C.c_int64(restore.numel()), N.ptr(dout), ld,
                                                   N.ptr(cvm), N.ptr(lod), B, S, Eo, cvm_offset, dim_off, None, None,
                                                   0, 0, C.c_float(grad_scale), symm._h, C.c_int64(off),
                                                   C.c_int64(n), C.c_float(scale), N.stream_ptr()),
                    "pbx_sparse_push_dense")
            self.fold_done = True
            return
        N.check(N.cuda().pbx_sparse_push_ex(self._h, N.ptr(restore), C.c_int64(restore.numel()), N.ptr(dout), ld,
                                            N.ptr(cvm), N.ptr(lod), B, S, Eo, cvm_offset, dim_off, None, None, 0,
                                            0, C.c_float(grad_scale), N.stream_ptr()), "pbx_sparse_push")

    def push_keys(self, restore: torch.Tensor, grads: torch.Tensor, key2slot: torch.Tensor, cvm_offset: int,
                  grad_scale: float) -> None:
        """Push explicit per-occurrence gradients [n, hidden] (API-parity path of
        ``push_box_sparse``; reference PushCopy/PushMergeCopy, box_wrapper.cu:344-512)."""
        assert grads.is_contiguous() and grads.dtype == torch.float32
        N.check(N.cuda().pbx_sparse_push(self._h, N.ptr(restore), C.c_int64(restore.numel()), None, None, None, 0,
                                         0, 0, cvm_offset, 0, N.ptr(grads), N.ptr(key2slot), grads.shape[1], 1,
                                         C.c_float(grad_scale), N.stream_ptr()), "pbx_sparse_push")

    def peek_status(self) -> None:
        """Non-blocking: raises if an EARLIER step (whose status copy has reached the host) overflowed
        a mailbox or timed out at a barrier — a faulty step stops training within a step or two
        instead of at the end of the pass.  Costs one read of pinned memory."""
        st = N.cuda().pbx_sparse_status_peek(self._h)
        if st:
            raise RuntimeError(STATUS_MSG.get(st, f"sparse engine status {st}"))

    def check_status(self) -> None:
        """Synchronising health check; raises on mailbox overflow / barrier timeout."""
        st = C.c_int()
        N.check(N.cuda().pbx_sparse_status(self._h, C.byref(st), 1, N.stream_ptr()))
        if st.value:
            raise RuntimeError(STATUS_MSG.get(st.value, f"sparse engine status {st.value}"))

    PULL_PHASES = ("P0 reset", "P1 dedup+route keys", "P2 restore idx", "P2 clear dedup + barrier",
                   "P3 serve rows + barrier")
    PUSH_PHASES = ("Q0 init (records go out in k_push_records before this kernel)", "Q1 merge duplicate keys", "Q2 ship + barrier", "Q3 optimizer")

    def phase_times(self) -> dict:
        """Wall time (us) of every phase of the last pull and the last push on this rank, from
        ``globaltimer`` stamps written by one thread at the phase boundaries (synchronises)."""
        ts = (C.c_uint64 * 16)()
        N.check(N.cuda().pbx_sparse_phase_times(self._h, ts, N.stream_ptr()), "pbx_sparse_phase_times")
        t = [int(x) for x in ts]
        return {"pull": {n: round((t[i + 1] - t[i]) / 1e3, 1) for i, n in enumerate(self.PULL_PHASES)},
                "push": {n: round((t[9 + i] - t[8 + i]) / 1e3, 1) for i, n in enumerate(self.PUSH_PHASES)}}


class EmulatedSparseGroup:
    """All ranks of a sparse group as *virtual ranks* of one cooperative launch on ONE GPU.

    Used to test the multi-rank pull/push logic on a single device: the per-rank code is
    the production kernel, the cross-rank barrier becomes the grid barrier (separate
    spinning launches on one GPU could deadlock — see B200_PROFILING.md)."""

    def __init__(self, tables: Sequence[SparseTable], max_keys: int, cap: int = 0):
        self.nranks = len(tables)
        self.engines = [SparseEngine(t, max_keys, rank=r, nranks=self.nranks, cap=cap, connect=False)
                        for r, t in enumerate(tables)]
        SparseEngine.connect_local(self.engines)

    def pull(self, keys: Sequence[torch.Tensor]):
        R = self.nranks
        restores = [torch.empty(max(k.numel(), 1), dtype=torch.int32, device=k.device) for k in keys]
        hs = (C.c_void_p * R)(*[e._h for e in self.engines])
        kp = (C.c_void_p * R)(*[k.data_ptr() for k in keys])
        ns = (C.c_int64 * R)(*[k.numel() for k in keys])
        rp = (C.c_void_p * R)(*[r.data_ptr() for r in restores])
        N.check(N.cuda().pbx_sparse_pull_emu(hs, R, kp, ns, rp, N.stream_ptr()), "pbx_sparse_pull_emu")
        return [r[:k.numel()] for r, k in zip(restores, keys)]

    def push_pooled(self, restores, douts, cvms, lods, B, S, Eo, cvm_offset, dim_off, grad_scale):
        R = self.nranks
        hs = (C.c_void_p * R)(*[e._h for e in self.engines])
        rp = (C.c_void_p * R)(*[r.data_ptr() for r in restores])
        ns = (C.c_int64 * R)(*[r.numel() for r in restores])
        dp = (C.c_void_p * R)(*[d.data_ptr() for d in douts])
        cp = (C.c_void_p * R)(*[c.data_ptr() for c in cvms])
        lp = (C.c_void_p * R)(*[l.data_ptr() for l in lods])
        N.check(N.cuda().pbx_sparse_push_emu(hs, R, rp, ns, dp, cp, lp, B, S, Eo, cvm_offset, dim_off,
                                             C.c_float(grad_scale), N.stream_ptr()), "pbx_sparse_push_emu")

    def check_status(self):
        for e in self.engines:
            e.check_status()
