#!/usr/bin/env python
"""B-nccl: the library-composed baseline of BASELINE.md §4 for config 3 (DeepFM).

Same model, data shape, table size and optimizers as ``bench.py``, but every hot path is the
"call a library" version the north star names as *the baseline, not the product*:

  sparse pull   torch.unique (sort-based dedup, as the reference's cub path) -> bucket by owner
                (argsort) -> NCCL all_to_all of keys -> hash lookup on the owner -> NCCL all_to_all
                of values -> index_select                       (HeterComm flow with NCCL exchange)
  seqpool+CVM   torch segment sum (index_add_) + log transforms
  towers        torch.nn.Linear under bf16 autocast (cuBLAS), PyTorch autograd
  FM / loss     plain torch ops
  sparse push   autograd gradient of the gathered rows -> index_add_ merge by unique key ->
                NCCL all_to_all of merged records -> optimizer update on the owner
  dense sync    one flattened NCCL all_reduce + separate scale (c_allreduce_sum + scale),
                torch.optim.Adam(fused=True)

The owner-side hash lookup and optimizer update use this repo's table kernels (the reference's
equivalents are its hashtable kernels); everything between them is library code.  Launched like
bench.py:  torchrun --nproc-per-node N benchmarks/baseline_nccl.py --gpus N --steps K --warmup W
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def mix64(k: torch.Tensor) -> torch.Tensor:
    def c(v):
        return torch.tensor(v - (1 << 64) if v >= (1 << 63) else v, dtype=torch.int64, device=k.device)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xff51afd7ed558ccd)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * c(0xc4ceb9fe1a85ec53)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    return k


class Exchange:
    """dedup + route + NCCL all-to-all; keeps what the push needs."""

    def __init__(self, table, world, rank, dev):
        self.table, self.world, self.rank, self.dev = table, world, rank, dev

    def pull(self, keys):
        uk, inv = torch.unique(keys, return_inverse=True)
        if self.world == 1:
            self.state = (uk, inv, None, None, None, uk)
            return self.table.lookup(uk), inv
        dst = ((mix64(uk) >> 40) & 0xFFFFFF) % self.world
        order = torch.argsort(dst)
        sk = uk[order]
        counts = torch.bincount(dst, minlength=self.world)
        rc = torch.empty_like(counts)
        dist.all_to_all_single(rc, counts)
        cl, rl = counts.tolist(), rc.tolist()                      # host sync, as the reference's shard sizes
        rk = torch.empty(sum(rl), dtype=torch.int64, device=self.dev)
        dist.all_to_all_single(rk, sk, rl, cl)
        vals = self.table.lookup(rk)
        back = torch.empty(sk.numel(), vals.shape[1], device=self.dev)
        dist.all_to_all_single(back, vals, cl, rl)
        res = torch.empty_like(back)
        res[order] = back
        self.state = (uk, inv, order, cl, rl, rk)
        return res, inv

    def push(self, merged):
        """merged: [n_unique, push_stride] records in unique-key order."""
        uk, inv, order, cl, rl, rk = self.state
        if self.world == 1:
            self.table.update(uk, merged)
            return
        send = merged[order]
        recv = torch.empty(sum(rl), merged.shape[1], device=self.dev)
        dist.all_to_all_single(recv, send, rl, cl)
        # a key requested by several ranks arrives several times: apply per source rank, in order
        off = 0
        for r in range(self.world):
            n = rl[r]
            if n:
                self.table.update(rk[off:off + n], recv[off:off + n])
            off += n


class TorchDeepFM(nn.Module):
    def __init__(self, S, D, dense, hidden):
        super().__init__()
        self.S, self.D, self.E = S, D, 3 + D
        dims = [S * self.E + dense, *hidden, 1]
        layers = []
        for i, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            layers.append(nn.Linear(a, b))
            if i < len(dims) - 2:
                layers.append(nn.ReLU())
        self.tower = nn.Sequential(*layers)
        self.bias = nn.Parameter(torch.zeros(1))

    def forward(self, pooled, dense_feats, label):
        B = pooled.shape[0]
        p = pooled.view(B, self.S, self.E)
        show, clk = p[..., 0], p[..., 1]
        cv = torch.stack([torch.log(show + 1), torch.log(clk + 1) - torch.log(show + 1)], -1)
        x3 = torch.cat([cv, p[..., 2:]], -1)
        x = torch.cat([x3.reshape(B, -1), dense_feats], 1)
        v = x3[..., 3:]
        fm = x3[..., 2].sum(1) + 0.5 * ((v.sum(1) ** 2) - (v ** 2).sum(1)).sum(1)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            deep = self.tower(x)
        logit = deep.float().view(-1) + fm + self.bias
        return nn.functional.binary_cross_entropy_with_logits(logit, label)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--slots", type=int, default=26)
    ap.add_argument("--dim", type=int, default=16)
    ap.add_argument("--dense", type=int, default=13)
    ap.add_argument("--max-bag", type=int, default=5)
    ap.add_argument("--hidden", default="512,256,128")
    ap.add_argument("--keys-per-gpu", type=int, default=125_000_000)
    ap.add_argument("--n-batches", type=int, default=8)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    import ctypes as C
    import bench as BM
    from paddlebox_b200 import _native as N
    from paddlebox_b200.ops.sparse_table import SparseTable
    lib = N.cuda()
    B, S, D = args.batch, args.slots, args.dim
    H = 3 + D
    table = SparseTable(int(args.keys_per_gpu / 0.6) + 1024, D, "adagrad",
                        config=N.SparseOptConfig(mf_create_thresholds=10.0))
    table.fill_synthetic(args.keys_per_gpu * world, 1, rank, world, create_mf=True)
    ex = Exchange(table, world, rank, dev)
    torch.manual_seed(0)
    model = TorchDeepFM(S, D, args.dense, tuple(int(x) for x in args.hidden.split(","))).to(dev)
    params = [p for p in model.parameters()]
    if world > 1:
        for p in params:
            dist.broadcast(p.data, 0)
    opt = torch.optim.Adam(params, lr=1e-3, fused=True)
    PS = table.layout.push_stride
    scratch = torch.empty(1024, dtype=torch.int32, device=dev)
    batches = []
    for keys_h, off_h, dense_h, n in BM.make_host_batches(args, rank, world, args.n_batches):
        dk, do, dd = keys_h.to(dev), off_h.to(dev), dense_h.to(dev)
        keys = torch.empty(n, dtype=torch.int64, device=dev)
        lod = torch.empty(S * B + 1, dtype=torch.int32, device=dev)
        N.check(lib.pbx_batch_build(N.ptr(dk), N.ptr(do), B, S, N.ptr(keys), N.ptr(lod), N.ptr(scratch),
                                    N.stream_ptr()))
        batches.append((keys, lod, dd, n))
    torch.cuda.synchronize()
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    bags = torch.arange(S * B, device=dev)

    def step(batch):
        keys, lod, dense, n = batch
        vals_u, inv = ex.pull(keys)
        vals_u = vals_u[:, :H].contiguous().requires_grad_(True)
        rows = vals_u[inv]
        lens = (lod[1:] - lod[:-1]).long()
        bag_id = torch.repeat_interleave(bags, lens, output_size=n)
        pooled = torch.zeros(S * B, H, device=dev).index_add_(0, bag_id, rows)
        pooled = pooled.view(S, B, H).permute(1, 0, 2).reshape(B, S * H)
        label = dense[:, 0]
        loss = model(pooled, dense[:, 1:], label)
        opt.zero_grad(set_to_none=False)
        loss.backward()
        # merged push records [slot, show, clk, embed_g, embedx_g]
        U = vals_u.shape[0]
        merged = torch.zeros(U, PS, device=dev)
        ins = bag_id % B
        merged[:, 0] = torch.zeros(U, device=dev).index_copy_(0, inv, (bag_id // B).float())
        merged[:, 1] = torch.bincount(inv, minlength=U).float()
        merged[:, 2] = torch.zeros(U, device=dev).index_add_(0, inv, label[ins])
        merged[:, 3:3 + H - 2] = vals_u.grad[:, 2:] * (-float(B))
        ex.push(merged)
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat.mul_(1.0 / world)
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(batches[i % len(batches)])
    barrier()
    evs = []
    for i in range(args.steps):
        N.check(lib.pbx_l2_flush(N.ptr(flush_buf), C.c_int64(flush_buf.numel()), N.stream_ptr()))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        loss = step(batches[(args.warmup + i) % len(batches)])
        b.record()
        evs.append((a, b))
    barrier()
    ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"impl": "b-nccl (torch.unique + NCCL all_to_all + cuBLAS towers + NCCL allreduce)",
                          "metric": "instances/sec (DeepFM sparse-CTR training, device-timed, max over ranks)",
                          "value": round(B * world * args.steps / (ms / 1e3), 1), "unit": "instances/s",
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": round(ms / args.steps, 4), "loss": float(loss),
                          "config": {"model": "DeepFM", "batch_per_gpu": B, "sparse_slots": S, "emb_dim": D,
                                     "table_keys_per_gpu": args.keys_per_gpu, "l2": "flushed between timed steps"}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
