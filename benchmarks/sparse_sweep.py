#!/usr/bin/env python
"""BASELINE config 5: sparse pull/push all-to-all sweep (keys/batch 64K-16M x dim 8-128) at
2/4/8 GPUs, device-timed, max over ranks, against an NCCL all-to-all implementation of the
same algorithm (the labelled baseline) and against the NVLink roofline.

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/sparse_sweep.py \
        --keys 65536,1048576 --dims 8,16,64 --json out.json

Per (keys, dim) the script reports, for pull and push:
  us            kernel time (max over ranks, median of iterations)
  nvlink_MB     bytes this rank must move over NVLink (keys out + values in for pull, merged
                records out for push), counted from the routed unique keys
  roofline_us   max(nvlink bytes / 900 GB/s (NVLink 5 per direction), local HBM bytes / hbm_gbs)
  frac          roofline_us / us
  nccl_us       the same pull done as torch.unique + all_to_all_single(keys) + table lookup +
                all_to_all_single(values) + index_select  (NCCL baseline)
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NVLINK_GBS = 900.0   # NVLink 5 per direction per GPU: the denominator BASELINE.md / SURVEY.md §6 name


def hbm_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6650.0


def timed(fn, iters, warmup, dev):
    for _ in range(warmup):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--keys", default="65536,1048576,4194304,16777216")   # config 5: 64 K - 16 M keys per batch
    ap.add_argument("--dims", default="8,16,64,128")
    ap.add_argument("--table-keys-per-gpu", type=int, default=20_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default="")
    ap.add_argument("--no-nccl", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from paddlebox_b200.box_wrapper import owner_of
    from paddlebox_b200.ops.sparse_table import SparseTable, synthetic_key
    from paddlebox_b200.parallel.sparse_engine import SparseEngine
    out = []
    for dim in [int(x) for x in a.dims.split(",")]:
        for nk in [int(x) for x in a.keys.split(",")]:
            table = SparseTable(int(a.table_keys_per_gpu / 0.75), dim, "adagrad")   # reference load factor
            table.fill_synthetic(a.table_keys_per_gpu * world, 1, rank, world)
            eng = SparseEngine(table, nk, rank, world, cap=int(2.0 * nk / world) + 4096)
            g = torch.Generator().manual_seed(rank)
            n_total = a.table_keys_per_gpu * world
            idx = (torch.rand(nk, generator=g, dtype=torch.float64) * n_total).long().clamp_(max=n_total - 1)
            keys = synthetic_key(idx, 1).to(dev)
            restore = torch.empty(nk, dtype=torch.int32, device=dev)
            H = 3 + dim
            # one bag per key: B*S = nk with S = 1 keeps the push path simple
            lod = torch.arange(nk + 1, dtype=torch.int32, device=dev)
            dout = torch.randn(nk, H, device=dev)
            cvm = torch.ones(nk, 2, device=dev)
            pull = lambda: eng.pull(keys, restore)
            push = lambda: eng.push_pooled(restore, dout, cvm, lod, nk, 1, H, 2, 0, -1.0)
            t_pull = timed(pull, a.iters, 3, dev)
            t_push = timed(push, a.iters, 3, dev)
            eng.check_status()
            pull(); push()
            phases = eng.phase_times()
            uk = torch.unique(keys)
            own = owner_of(uk.cpu().numpy().view("uint64"), world)
            remote = int((own != rank).sum())
            u = uk.numel()
            ps, pd, rs = eng.pull_stride * 4, eng.push_stride * 4, table.layout.stride * 4
            nv_pull = remote * (8 + ps)
            nv_push = remote * pd
            hbm = hbm_gbs()
            local_pull = nk * 12 + u * (8 + 128 + ps)          # keys, restore, probe line, row line, value write
            local_push = nk * (4 + H * 4) + u * (2 * pd + 2 * rs)
            roof_pull = max(nv_pull / NVLINK_GBS / 1e3, local_pull / hbm / 1e3)
            roof_push = max(nv_push / NVLINK_GBS / 1e3, local_push / hbm / 1e3)
            rec = {"n_gpus": world, "keys_per_batch": nk, "dim": dim, "unique": u, "remote_unique": remote,
                   "pull_us": round(t_pull, 1), "push_us": round(t_push, 1),
                   "pull_nvlink_MB": round(nv_pull / 1e6, 2), "push_nvlink_MB": round(nv_push / 1e6, 2),
                   "pull_roofline_us": round(roof_pull, 1), "push_roofline_us": round(roof_push, 1),
                   "pull_frac": round(roof_pull / t_pull, 3), "push_frac": round(roof_push / t_push, 3),
                   "pull_nvlink_GBps": round(nv_pull / t_pull / 1e3, 1), "push_nvlink_GBps": round(nv_push / t_push / 1e3, 1)}
            rec["phases_us_rank0"] = phases
            if not a.no_nccl:
                # ---- NCCL baseline of the same pull: unique -> a2a keys -> lookup -> a2a values -> gather
                def nccl_pull():
                    ukk, inv = torch.unique(keys, return_inverse=True)
                    h = _mix64(ukk)
                    dst = ((h >> 40) & 0xFFFFFF) % world
                    order = torch.argsort(dst)
                    sk = ukk[order]
                    counts = torch.bincount(dst, minlength=world)
                    rc = torch.empty_like(counts)
                    dist.all_to_all_single(rc, counts)
                    cl, rl = counts.tolist(), rc.tolist()
                    rk = torch.empty(sum(rl), dtype=torch.int64, device=dev)
                    dist.all_to_all_single(rk, sk, rl, cl)
                    vals = table.lookup(rk)
                    back = torch.empty(sk.numel(), H, device=dev)
                    dist.all_to_all_single(back, vals, [c for c in cl], [r for r in rl])
                    res = torch.empty_like(back)
                    res[order] = back
                    return res[inv]
                rec["nccl_pull_us"] = round(timed(nccl_pull, max(3, a.iters // 2), 2, dev), 1)
                rec["pull_speedup_vs_nccl"] = round(rec["nccl_pull_us"] / t_pull, 2)
            out.append(rec)
            if rank == 0:
                print(json.dumps(rec), flush=True)
            del eng, table
            torch.cuda.empty_cache()
    if rank == 0 and a.json:
        json.dump(out, open(a.json, "w"), indent=1)
    dist.destroy_process_group()


def _mix64(k: torch.Tensor) -> torch.Tensor:
    def c(v):
        return torch.tensor(v - (1 << 64) if v >= (1 << 63) else v, dtype=torch.int64, device=k.device)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xff51afd7ed558ccd)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xc4ceb9fe1a85ec53)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    return k


if __name__ == "__main__":
    main()
