#!/usr/bin/env python
"""Concurrent pinned host->device copy rate with all ranks copying at once (what the pipelined
feed does every step), unbound vs bound to the GPU's NUMA-local cores (NVML cpu affinity).
    torchrun --nproc-per-node 8 benchmarks/probes/h2d_concurrent.py"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


def gpu_cores(idx):
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cores = [64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1]
        return cores
    except Exception as e:
        return []


def run(tag, dev, nbytes=3_228_419, iters=200):
    src = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    src.fill_(1)
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    st = torch.cuda.Stream()
    for _ in range(5):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        a.record()
        for _ in range(iters):
            dst.copy_(src, non_blocking=True)
        b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / iters
    t = torch.tensor([us], device=dev, dtype=torch.float64)
    mx = t.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    if dist.get_rank() == 0:
        print(f"{tag}: {nbytes/1e6:.2f} MB copy, slowest rank {float(mx):.1f} us/copy = {nbytes/float(mx)/1e3:.1f} GB/s per GPU",
              flush=True)


def main():
    rank, lr = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    cores = gpu_cores(lr)
    if rank == 0:
        print("cpu count", os.cpu_count(), "affinity now", len(os.sched_getaffinity(0)), "gpu0 local cores", cores[:8], "...",
              len(cores), flush=True)
    run("unbound", dev)
    if cores:
        os.sched_setaffinity(0, set(cores))
        run("bound to NVML-local cores (new pinned buffer)", dev)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
