"""Correctness self-check of the sharded sparse path on the ranks and GPUs it really runs on.

``bench.py`` calls this after its timed regions (never inside one) at every N, so that the
driver's scaling run carries evidence that the multi-GPU kernels — system-scope flag barriers,
IPC-mapped mailboxes, remote stores, cross-GPU row locks — compute the right thing, not only that
they finish.  Two checks, both against owner-side oracles gathered over ``torch.distributed``:

* pull: every rank pulls a key sample (keys shared by all ranks, keys private to the rank, keys
  absent from the table, the padding key 0, duplicates) through the engine and compares with
  ``SparseTable.lookup`` executed on the owning rank;
* push: every rank pushes gradients for a batch of keys private to it (duplicates inside the
  batch exercise the merge) and the owners' rows are compared with
  ``ops/reference.py::sparse_row_update`` applied to the rows read before the push.

Reference semantics checked: HeterComm::pull_sparse / push_sparse (heter_ps/heter_comm_inl.h:1477-1825),
PushMergeCopy (fleet/box_wrapper.cu:417-475), SparseAdagradOptimizer (heter_ps/optimizer.cuh.h:30-145).
"""
from __future__ import annotations

import numpy as np
import torch


def _owner_rows(table, keys_all: torch.Tensor, rank: int, world: int, full_rows: bool, group=None) -> torch.Tensor:
    """Rows of ``keys_all`` (identical on every rank) read on the owning rank, summed to all ranks."""
    import torch.distributed as dist
    from ..box_wrapper import owner_of
    own = torch.from_numpy(owner_of(keys_all.cpu().numpy().astype(np.uint64), world) == rank).to(keys_all.device)
    width = table.layout.stride if full_rows else table.hidden
    out = torch.zeros(keys_all.numel(), width, dtype=torch.float32, device=keys_all.device)
    if bool(own.any()):
        mine = keys_all[own]
        out[own] = table.fetch_rows(mine) if full_rows else table.lookup(mine)
    if world > 1:
        dist.all_reduce(out, group=group)
    return out


def _gather_keys(keys: torch.Tensor, world: int, group=None) -> torch.Tensor:
    import torch.distributed as dist
    if world == 1:
        return keys[None]
    out = [torch.empty_like(keys) for _ in range(world)]
    dist.all_gather(out, keys, group=group)
    return torch.stack(out)


def sparse_selfcheck(box, n_total_keys: int, rank: int, world: int, n: int = 2048, group=None) -> dict:
    """Returns ``{"pull_max_abs", "push_max_abs", ..., "ok"}`` (identical on all ranks); ``push_max_abs`` is
    max |got - expected| / (1 + |expected|) over every word of every updated row."""
    import torch.distributed as dist
    from ..ops import reference as R
    from ..ops.sparse_table import synthetic_key
    eng, table = box.engine, box.table
    dev = table.device
    H = table.hidden
    fold = getattr(eng, "fold_dense", None)
    eng.fold_dense = None                      # these pushes must not carry a dense allreduce
    try:
        # ------------------------------------------------------------------ pull
        g = torch.Generator().manual_seed(77 + rank)
        shared = synthetic_key(torch.arange(n // 2))                                   # hot keys, every rank
        priv = synthetic_key(torch.randint(0, n_total_keys, (n // 2,), generator=g))    # this rank's own draw
        absent = synthetic_key(torch.arange(n_total_keys + 5, n_total_keys + 5 + 64))  # not in the table
        keys = torch.cat([shared, priv, absent, torch.zeros(8, dtype=torch.int64), shared[:32], priv[:32]]).to(dev)
        restore = eng.pull(keys)
        got = eng.values[restore.long()][:, :H].clone()
        all_keys = _gather_keys(keys, world, group)
        exp = _owner_rows(table, all_keys.reshape(-1), rank, world, False, group).view(world, keys.numel(), H)[rank]
        pull_err = float((got - exp).abs().max())
        n_nonzero = int((exp.abs().sum(1) > 0).sum())
        # ------------------------------------------------------------------ push (private keys)
        B, S, bag = 64, 4, 2
        idx = (torch.randint(0, max(n_total_keys // world - 1, 1), (200,), generator=g) * world + rank) % n_total_keys
        pk = synthetic_key(idx)
        occ = pk[torch.randint(0, pk.numel(), (B * S * bag,), generator=g)].to(dev)
        lod = (torch.arange(S * B + 1, dtype=torch.int32) * bag).to(dev)
        dout = torch.randn(B, S * H, generator=g).to(dev)
        cvm = torch.stack([torch.ones(B), (torch.rand(B, generator=g) < 0.3).float()], 1).to(dev)
        uk, inv = torch.unique(occ, return_inverse=True)
        all_uk_n = torch.tensor([uk.numel()], device=dev)
        cnts = [torch.zeros_like(all_uk_n) for _ in range(world)]
        if world > 1:
            dist.all_gather(cnts, all_uk_n, group=group)
        else:
            cnts = [all_uk_n]
        m = int(max(int(c) for c in cnts))
        uk_pad = torch.zeros(m, dtype=torch.int64, device=dev)
        uk_pad[:uk.numel()] = uk
        all_uk = _gather_keys(uk_pad, world, group).reshape(-1)
        before = _owner_rows(table, all_uk, rank, world, True, group).view(world, m, -1)[rank][:uk.numel()]
        r2 = eng.pull(occ)
        eng.push_pooled(r2, dout, cvm, lod, B, S, H, 2, 0, -float(B))
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier(group=group)
        after = _owner_rows(table, all_uk, rank, world, True, group).view(world, m, -1)[rank][:uk.numel()]
        dx = R.seqpool_cvm_bwd_vec(dout.cpu(), cvm.cpu(), lod.cpu(), B, S, H)
        k2s = torch.repeat_interleave(torch.arange(S), torch.full((S,), B * bag))
        merged = R.merge_push(inv.cpu(), uk.numel(), dx, k2s, 2, -float(B))
        L, cfg = table.layout, table.config
        slot_col = L.off_tail + 1
        push_err, n_found = 0.0, 0
        b_cpu, a_cpu = before.cpu(), after.cpu()
        for i in range(uk.numel()):
            if float(b_cpu[i].abs().sum()) == 0.0:
                continue                       # key absent from the table: the push must be ignored
            n_found += 1
            e = torch.tensor(R.sparse_row_update(b_cpu[i].tolist(), merged[i].tolist(), L, cfg))
            e[slot_col] = a_cpu[i][slot_col]    # slot id of a key seen in several slots: last writer wins
            e[L.stride - 1] = 0.0
            push_err = max(push_err, float(((a_cpu[i] - e).abs() / (1.0 + e.abs())).max()))
        t = torch.tensor([pull_err, push_err], dtype=torch.float64, device=dev)
        cnt = torch.tensor([n_nonzero, n_found], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            dist.all_reduce(cnt, op=dist.ReduceOp.MIN, group=group)
        eng.check_status()
        pull_err, push_err = t.tolist()
        ok = pull_err == 0.0 and push_err < 2e-3 and int(cnt[0]) > n // 2 and int(cnt[1]) > 50
        return {"pull_max_abs": pull_err, "push_max_abs": round(push_err, 7), "pull_keys_found_min": int(cnt[0]),
                "push_keys_found_min": int(cnt[1]), "ok": bool(ok)}
    finally:
        eng.fold_dense = fold
