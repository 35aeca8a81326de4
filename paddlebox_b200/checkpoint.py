"""Dense-parameter checkpoints in the reference's on-disk layout.

``fluid.io.save_persistables`` writes one file per variable, each a serialized LoDTensor
(framework/lod_tensor.cc:205-231, framework/tensor_util.cc:1056-1110):

    u32  lod-tensor version (0)
    u64  lod_level, then per level: u64 byte size + raw offsets
    u32  tensor version (0)
    i32  size of the TensorDesc protobuf, then the protobuf:
         field 1 (varint)  data_type   (FP32 = 5, INT64 = 3, INT32 = 2, FP64 = 6, BF16 = 22, FP16 = 4)
         field 2 (varint, repeated, unpacked) dims
    raw  tensor data (little endian)

This module reads and writes that format without protobuf (the TensorDesc message is two
varint fields) and names the files with the reference's parameter names (``fc_0.w_0`` ...;
SURVEY §9.6), so dense checkpoints interchange with a Paddle program of the same topology.
Sparse-table checkpoints are handled by ``BoxWrapper.save_base / save_delta``.
"""
from __future__ import annotations

import os
import struct

import numpy as np
import torch

_DTYPES = {5: np.float32, 6: np.float64, 3: np.int64, 2: np.int32, 4: np.float16, 0: np.bool_, 1: np.int16,
           20: np.uint8, 21: np.int8}
_CODES = {np.dtype(v): k for k, v in _DTYPES.items()}
_BF16 = 22


def _varint(n: int) -> bytes:
    out = bytearray()
    n &= (1 << 64) - 1
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int):
    shift, val = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def serialize_lod_tensor(t: torch.Tensor, lod=()) -> bytes:
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        code, raw = _BF16, t.view(torch.int16).numpy().tobytes()
    else:
        arr = t.numpy()
        code, raw = _CODES[arr.dtype], arr.tobytes()
    desc = b"\x08" + _varint(code) + b"".join(b"\x10" + _varint(d) for d in t.shape)
    out = [struct.pack("<I", 0), struct.pack("<Q", len(lod))]
    for level in lod:
        lv = np.asarray(level, dtype=np.uint64)
        out += [struct.pack("<Q", lv.nbytes), lv.tobytes()]
    out += [struct.pack("<I", 0), struct.pack("<i", len(desc)), desc, raw]
    return b"".join(out)


def deserialize_lod_tensor(buf: bytes):
    pos = 0
    (ver,) = struct.unpack_from("<I", buf, pos)
    pos += 4
    if ver != 0:
        raise ValueError(f"unsupported LoDTensor version {ver}")
    (levels,) = struct.unpack_from("<Q", buf, pos)
    pos += 8
    lod = []
    for _ in range(levels):
        (nbytes,) = struct.unpack_from("<Q", buf, pos)
        pos += 8
        lod.append(np.frombuffer(buf, dtype=np.uint64, count=nbytes // 8, offset=pos).tolist())
        pos += nbytes
    (tver,) = struct.unpack_from("<I", buf, pos)
    pos += 4
    (dsize,) = struct.unpack_from("<i", buf, pos)
    pos += 4
    end = pos + dsize
    code, dims = None, []
    while pos < end:
        tag, pos = _read_varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
            if field == 1:
                code = v
            elif field == 2:
                dims.append(v if v < (1 << 63) else v - (1 << 64))
        elif wire == 2:      # packed dims
            ln, pos = _read_varint(buf, pos)
            sub_end = pos + ln
            while pos < sub_end:
                v, pos = _read_varint(buf, pos)
                dims.append(v)
        else:
            raise ValueError("unexpected wire type in TensorDesc")
    n = int(np.prod(dims)) if dims else 1
    if code == _BF16:
        arr = np.frombuffer(buf, dtype=np.int16, count=n, offset=pos).copy()
        t = torch.from_numpy(arr).view(torch.bfloat16).reshape(dims)
    else:
        arr = np.frombuffer(buf, dtype=_DTYPES[code], count=n, offset=pos).copy()
        t = torch.from_numpy(arr).reshape(dims)
    return t, lod


def named_persistables(model) -> list:
    if hasattr(model, "named_dense_params"):
        return list(model.named_dense_params())
    return [(n, p) for n, p in model.named_parameters()] + [(n, b) for n, b in model.named_buffers()]


def _is_fc_weight(name: str, t: torch.Tensor) -> bool:
    """FC weights live as [out, in] here (the layout the tcgen05 towers read) and as [in, out] in a
    reference checkpoint (``mul``/``fc`` ops compute X @ W): ``*.w_0`` tensors, and their Adam
    moments, are written and read transposed in the last two dimensions."""
    base = name
    for suf in ("_moment1_0", "_moment2_0"):
        if base.endswith(suf):
            base = base[:-len(suf)]
    return base.endswith(".w_0") and t.dim() >= 2


def _to_file_layout(name: str, t: torch.Tensor) -> torch.Tensor:
    return t.transpose(-1, -2).contiguous() if _is_fc_weight(name, t) else t


def _from_file_layout(name: str, src: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
    """``src`` as read from the file -> the layout of ``like``; a shape that does not fit raises
    (no silent reshape of a tensor that merely has the right number of elements)."""
    want = tuple(like.shape)
    if _is_fc_weight(name, like):
        want_file = want[:-2] + (want[-1], want[-2])
        if tuple(src.shape) != want_file:
            raise ValueError(f"{name}: checkpoint holds shape {tuple(src.shape)}, expected {want_file} ([in, out])")
        return src.transpose(-1, -2)
    if tuple(src.shape) != want:
        if src.numel() == like.numel() and (src.dim() <= 1 or like.dim() <= 1):
            return src.reshape(want)             # [n] <-> [n, 1] / [1, n] vectors
        raise ValueError(f"{name}: checkpoint holds shape {tuple(src.shape)}, expected {want}")
    return src


def _moment_items(items, optimizer_store):
    """(name, view of m, view of v) per trainable parameter; the store lays parameters out in
    ``module.parameters()`` order, which is NOT the order of ``named_dense_params`` — match by identity."""
    off_of = {id(p): o for p, o in zip(optimizer_store.params, optimizer_store.offsets)}
    out = []
    for name, p in items:
        if not getattr(p, "requires_grad", False) or id(p) not in off_of:
            continue
        off, n = off_of[id(p)], p.numel()
        out.append((name, p, optimizer_store.m[off:off + n].view(p.shape), optimizer_store.v[off:off + n].view(p.shape)))
    return out


def save_persistables(model, dirname: str, optimizer_store=None, filename: str | None = None) -> list:
    """One file per variable (or one combined file when ``filename`` is given — the reference's
    ``save_combine`` layout: the serialized tensors back to back in name order).  FC weights are
    written in the reference's [in, out] layout (see ``_is_fc_weight``)."""
    os.makedirs(dirname, exist_ok=True)
    items = named_persistables(model)
    if optimizer_store is not None and getattr(optimizer_store, "m", None) is not None:
        # Adam moments with the reference's naming: <param>_moment1_0 / _moment2_0
        for name, p, m, v in _moment_items(list(items), optimizer_store):
            items.append((f"{name}_moment1_0", m))
            items.append((f"{name}_moment2_0", v))
    names = []
    if filename:
        with open(os.path.join(dirname, filename), "wb") as f:
            for name, t in sorted(items, key=lambda kv: kv[0]):
                f.write(serialize_lod_tensor(_to_file_layout(name, t.detach())))
                names.append(name)
        with open(os.path.join(dirname, filename + ".names"), "w") as f:
            f.write("\n".join(names))
        return names
    for name, t in items:
        with open(os.path.join(dirname, name), "wb") as f:
            f.write(serialize_lod_tensor(_to_file_layout(name, t.detach())))
        names.append(name)
    return names


def load_persistables(model, dirname: str, optimizer_store=None, strict: bool = True) -> list:
    loaded = []
    items = named_persistables(model)
    with torch.no_grad():
        for name, t in items:
            path = os.path.join(dirname, name)
            if not os.path.exists(path):
                if strict:
                    raise FileNotFoundError(path)
                continue
            src, _ = deserialize_lod_tensor(open(path, "rb").read())
            t.copy_(_from_file_layout(name, src.to(t.dtype), t))
            loaded.append(name)
        if optimizer_store is not None and getattr(optimizer_store, "m", None) is not None:
            for name, p, m, v in _moment_items(items, optimizer_store):
                for suffix, buf in (("_moment1_0", m), ("_moment2_0", v)):
                    path = os.path.join(dirname, name + suffix)
                    if os.path.exists(path):
                        src, _ = deserialize_lod_tensor(open(path, "rb").read())
                        buf.copy_(_from_file_layout(name + suffix, src.to(buf.dtype), buf))
                        loaded.append(name + suffix)
    fn = getattr(model, "mark_weights_dirty", None)
    if fn:
        fn()
    return loaded
