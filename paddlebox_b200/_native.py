"""ctypes bindings of the two in-tree native libraries.

``cuda()`` returns the handle of ``libpbx_cuda.so`` (sm_100a kernels) and ``host()`` the
handle of ``libpbx_host.so`` (C++ host runtime).  Libraries are built on first use if the
``.so`` is missing (``paddlebox_b200._build``).  On a machine with a GPU a missing or
unloadable CUDA library is a hard error — ops never fall back silently to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

import torch

from . import _build

_LOCK = threading.Lock()
_CUDA = None
_HOST = None

c_void_p, c_int, c_int64, c_uint64, c_float, c_char_p = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float, C.c_char_p


class SparseOptConfig(C.Structure):
    """Mirror of ``PbxSparseOptConfig`` (include/pbx_types.h); defaults from the reference's
    ``OptimizerConfig`` (heter_ps/optimizer_conf.h:20-45)."""
    _fields_ = [(n, C.c_float) for n in (
        "nonclk_coeff", "clk_coeff", "min_bound", "max_bound", "learning_rate", "initial_g2sum",
        "initial_range", "beta1_decay_rate", "beta2_decay_rate", "ada_epsilon", "mf_create_thresholds",
        "mf_learning_rate", "mf_initial_g2sum", "mf_initial_range", "mf_beta1_decay_rate",
        "mf_beta2_decay_rate", "mf_min_bound", "mf_max_bound", "mf_ada_epsilon", "reserved")]

    DEFAULTS = dict(nonclk_coeff=0.1, clk_coeff=1.0, min_bound=-10.0, max_bound=10.0, learning_rate=0.05,
                    initial_g2sum=3.0, initial_range=0.0, beta1_decay_rate=0.9, beta2_decay_rate=0.999,
                    ada_epsilon=1e-8, mf_create_thresholds=10.0, mf_learning_rate=0.05, mf_initial_g2sum=3.0,
                    mf_initial_range=1e-4, mf_beta1_decay_rate=0.9, mf_beta2_decay_rate=0.999,
                    mf_min_bound=-10.0, mf_max_bound=10.0, mf_ada_epsilon=1e-8, reserved=0.0)

    def __init__(self, **kw):
        super().__init__()
        vals = dict(self.DEFAULTS)
        for k, v in kw.items():
            if k not in vals:
                raise KeyError(f"unknown sparse optimizer option {k!r}")
            vals[k] = v
        for k, v in vals.items():
            setattr(self, k, float(v))

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


class RowLayout(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("embedx_dim", "optimizer", "embed_state", "embedx_state", "off_mf",
                                       "off_tail", "off_es", "off_xs", "stride", "pull_stride", "push_stride")]


def make_layout(embedx_dim: int, optimizer: int) -> RowLayout:
    """Python twin of ``pbx_make_layout`` (include/pbx_types.h)."""
    l = RowLayout()
    l.embedx_dim, l.optimizer = embedx_dim, optimizer
    l.embed_state = 1 if optimizer == 0 else 4
    l.embedx_state = 1 if optimizer == 0 else (2 * embedx_dim + 2 if optimizer == 1 else 4)
    l.off_mf = 3 + embedx_dim
    l.off_tail = 4 + embedx_dim
    l.off_es = l.off_tail + 3
    l.off_xs = l.off_es + l.embed_state
    words = l.off_xs + l.embedx_state
    l.stride = (words + 1 + 31) // 32 * 32  # whole 128-byte lines; last word = row lock (see pbx_types.h)
    l.pull_stride = (3 + embedx_dim + 3) // 4 * 4
    l.push_stride = (4 + embedx_dim + 3) // 4 * 4
    return l


class SeqpoolParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("B", "S", "hidden", "x_stride", "out_stride", "Eo", "use_cvm", "cvm_offset",
                                       "clk_filter", "need_filter", "quant_ratio", "embed_threshold_filter",
                                       "embed_thres_size", "concat", "concat_filter", "variant", "show_filter")] + \
               [(n, C.c_float) for n in ("pad_value", "show_coeff", "clk_coeff", "threshold", "embed_threshold")] + \
               [("slot_threshold", C.c_void_p), ("trade_w", C.c_void_p)]


class GemmProblem(C.Structure):
    """Mirror of ``PbxGemmProblem`` (include/pbx_types.h)."""
    _fields_ = [(n, C.c_void_p) for n in ("A", "B", "bias", "mask", "out_bf16", "out_t", "out_f32", "col_out")] + \
               [(n, C.c_int) for n in ("lda", "ldb", "M", "N", "K", "ld_mask", "ld_out", "ld_t", "ld_f32", "relu",
                                       "accumulate", "splitk", "col_split")] + [("alpha", C.c_float)]


class NativeError(RuntimeError):
    pass


def _lib_path(name: str) -> Path:
    # PBX_HOST_LIB / PBX_CUDA_LIB: load another build of a library (a sanitizer build of the host runtime,
    # benchmarks/sanitize_pytests.sh)
    override = os.environ.get(f"{name.upper()}_LIB")
    if override:
        return Path(override)
    return _build.LIB_DIR / f"lib{name}.so"


def cuda_available() -> bool:
    return torch.cuda.is_available()


def cuda():
    """Handle of libpbx_cuda.so; raises if it cannot be loaded."""
    global _CUDA
    if _CUDA is None:
        with _LOCK:
            if _CUDA is None:
                p = _lib_path("pbx_cuda")
                if not p.exists():
                    _build.build_cuda(verbose=True)
                lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
                lib.pbx_last_error.restype = C.c_char_p
                lib.pbx_sparse_pull_values.restype = C.c_void_p
                lib.pbx_symm_ptr.restype = C.c_void_p
                _CUDA = lib
    return _CUDA


def host():
    """Handle of libpbx_host.so (built on demand; needs only g++)."""
    global _HOST
    if _HOST is None:
        with _LOCK:
            if _HOST is None:
                p = _lib_path("pbx_host")
                if not p.exists():
                    _build.build_host(verbose=True)
                lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
                lib.pbh_last_error.restype = C.c_char_p
                _HOST = lib
    return _HOST


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = cuda().pbx_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def hcheck(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = host().pbh_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'host call'} failed (rc={rc}): {msg}")


def stream_ptr(stream: "torch.cuda.Stream | None" = None) -> C.c_void_p:
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def ptr(t: "torch.Tensor | None") -> C.c_void_p:
    if t is None:
        return C.c_void_p(0)
    return C.c_void_p(t.data_ptr())


class _RawCudaBuffer:
    """Zero-copy torch view of device memory owned by a native object."""

    def __init__(self, address: int, shape, typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(address), False),
                                         "version": 3, "strides": None}
        self._owner = owner


def tensor_from_ptr(address: int, shape, dtype: torch.dtype, owner=None, device=None) -> torch.Tensor:
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.int64: "<i8", torch.uint8: "|u1",
               torch.float64: "<f8"}[dtype]
    buf = _RawCudaBuffer(address, shape, typestr, owner)
    t = torch.as_tensor(buf, device=device or torch.device("cuda", torch.cuda.current_device()))
    t._pbx_owner = buf  # keep the native owner alive as long as the view
    return t
