"""Per-batch NaN/Inf guard (reference FLAGS_check_nan_inf -> CheckBatchNanOrInfRet +
DumpAllScope + hard failure, boxps_worker.cc:1325-1336; nan_inf_utils_detail.cc:720-806)."""
from __future__ import annotations

import os

import torch


def has_nan_inf(tensors: dict) -> list:
    """Names of tensors that contain NaN or Inf (one fused device reduction per tensor,
    one host sync for all)."""
    names, flags = [], []
    for name, t in tensors.items():
        if t is None or not t.is_floating_point():
            continue
        names.append(name)
        flags.append((~torch.isfinite(t)).any())
    if not flags:
        return []
    bad = torch.stack(flags).cpu().tolist()
    return [n for n, b in zip(names, bad) if b]


def check_batch_nan_inf(tensors: dict, dump_dir: str | None = None, tag: str = "batch") -> None:
    bad = has_nan_inf(tensors)
    if not bad:
        return
    if dump_dir:
        os.makedirs(dump_dir, exist_ok=True)
        torch.save({k: v.detach().cpu() for k, v in tensors.items() if v is not None},
                   os.path.join(dump_dir, f"nan_inf_{tag}.pt"))
    raise FloatingPointError(f"NaN/Inf found in {bad} ({tag}); scope dumped to {dump_dir!r}")
