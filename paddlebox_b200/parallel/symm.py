"""Symmetric device memory and the dense collectives that run over it.

``SymmetricBuffer`` is one fp32 region of identical size on every rank, mapped into all
peers (CUDA IPC).  ``allreduce`` & co. are single cooperative kernels doing P2P loads /
stores over NVLink (csrc/cuda/symm.cu) with the gradient scale fused in — the product
path; ``torch.distributed`` (NCCL) is used for the rendezvous only and, in
``parallel/dense.py``, as the labelled baseline.

Reference role: NCCL call sites of SURVEY.md §2.11 (boxps_worker.cc:1191-1258,
c_allreduce_x_op.cc:114-151, c_mixallgather_op.cc:221-328).
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from .. import _native as N

OPS = {"allreduce": 0, "reduce_scatter": 1, "all_gather": 2, "broadcast": 3, "barrier": 4}


class SymmetricBuffer:
    def __init__(self, numel: int, rank: int = 0, nranks: int = 1, group=None, connect: bool = True, device=None):
        self.rank, self.nranks, self.numel = rank, nranks, int(numel)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._h = C.c_void_p()
        lib = N.cuda()
        with torch.cuda.device(self.device):
            N.check(lib.pbx_symm_create(C.byref(self._h), rank, nranks, C.c_uint64(self.numel * 4)),
                    "pbx_symm_create")
        self.tensor = N.tensor_from_ptr(lib.pbx_symm_ptr(self._h), (self.numel,), torch.float32, owner=self,
                                        device=self.device)
        if nranks > 1 and connect:
            import torch.distributed as dist
            buf = C.create_string_buffer(64)
            N.check(lib.pbx_symm_ipc_handle(self._h, buf))
            gathered = [None] * nranks
            dist.all_gather_object(gathered, bytes(buf.raw), group=group)
            N.check(lib.pbx_symm_connect(self._h, b"".join(gathered)), "pbx_symm_connect")
            dist.barrier(group=group)

    def __del__(self):
        try:
            if self._h:
                N.cuda().pbx_symm_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def collective(self, op: str, offset: int = 0, numel: int | None = None, scale: float = 1.0, root: int = 0,
                   max_blocks: int = 0, stream=None) -> None:
        n = self.numel - offset if numel is None else numel
        N.check(N.cuda().pbx_symm_collective(self._h, C.c_int64(offset), C.c_int64(n), C.c_float(scale), OPS[op],
                                             root, max_blocks, N.stream_ptr(stream)), f"pbx_symm_collective[{op}]")

    def allreduce(self, offset: int = 0, numel: int | None = None, scale: float = 1.0, stream=None) -> None:
        """In-place sum over ranks times ``scale`` (pass 1/nranks for the DP gradient mean)."""
        if self.nranks == 1:
            if scale != 1.0:
                n = self.numel - offset if numel is None else numel
                self.tensor[offset:offset + n].mul_(scale)
            return
        self.collective("allreduce", offset, numel, scale, stream=stream)

    def stat_update(self, offset: int, width: int, scale: float, decay: float, size: torch.Tensor, total: torch.Tensor,
                    sqsum: torch.Tensor, stream=None) -> None:
        """``summary = summary * decay + scale * sum over ranks of stats`` for the three data-norm summaries in one
        launch: the statistics ``[3, width]`` of every rank sit at ``offset`` of the symmetric region, the summary
        vectors are ordinary local tensors (csrc/cuda/symm.cu ``k_symm_stat_update``)."""
        assert size.is_contiguous() and total.is_contiguous() and sqsum.is_contiguous()
        N.check(N.cuda().pbx_symm_stat_update(self._h, C.c_int64(offset), width, C.c_float(scale), C.c_float(decay),
                                              N.ptr(size), N.ptr(total), N.ptr(sqsum), N.stream_ptr(stream)),
                "pbx_symm_stat_update")

    @staticmethod
    def stat_update_emu(bufs: Sequence["SymmetricBuffer"], offset: int, width: int, scale: float, decay: float,
                        summaries: Sequence[Sequence[torch.Tensor]]) -> None:
        """All ranks of an emulated group in one launch (tests): ``summaries[r] = (size, sum, sqsum)`` of rank r."""
        arr = (C.c_void_p * len(bufs))(*[b._h for b in bufs])
        flat = [t.data_ptr() for trio in summaries for t in trio]
        sp = (C.c_void_p * len(flat))(*flat)
        N.check(N.cuda().pbx_symm_stat_update_emu(arr, len(bufs), C.c_int64(offset), width, C.c_float(scale),
                                                  C.c_float(decay), sp, N.stream_ptr()), "pbx_symm_stat_update_emu")

    def check_status(self) -> None:
        st = C.c_int()
        N.check(N.cuda().pbx_symm_status(self._h, C.byref(st), N.stream_ptr()))
        if st.value:
            raise RuntimeError("symmetric-memory barrier timed out (a peer rank is missing)")

    # ---- single-GPU emulation of a group (tests) -------------------------------
    @staticmethod
    def emulated_group(nranks: int, numel: int) -> list["SymmetricBuffer"]:
        bufs = [SymmetricBuffer(numel, r, nranks, connect=False) for r in range(nranks)]
        arr = (C.c_void_p * nranks)(*[b._h for b in bufs])
        for b in bufs:
            N.check(N.cuda().pbx_symm_connect_local(b._h, arr))
        return bufs

    @staticmethod
    def collective_emu(bufs: Sequence["SymmetricBuffer"], offset: int, numel: int, scale: float, op: str,
                       root: int = 0) -> None:
        arr = (C.c_void_p * len(bufs))(*[b._h for b in bufs])
        N.check(N.cuda().pbx_symm_collective_emu(arr, len(bufs), C.c_int64(offset), C.c_int64(numel),
                                                 C.c_float(scale), OPS[op], root, N.stream_ptr()),
                "pbx_symm_collective_emu")
