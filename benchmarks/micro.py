#!/usr/bin/env python
"""Kernel micro-benchmarks with roofline fractions (device-timed with CUDA events, L2 flushed
between iterations, >= 3 warm-ups).  Peaks come from MEASURED_PEAKS.json (driver-written).

    python benchmarks/micro.py [--what gemm,sparse,seqpool] [--json out.json]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from paddlebox_b200 import _native as N  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


def timeit(fn, iters=20, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def bench_gemm(out, flush):
    from paddlebox_b200.ops.mlp import _pad8, gemm_bf16_tn
    hbm, tf, src = peaks()
    dev = torch.device("cuda")
    shapes = [("fwd L0", 4096, 512, 507), ("fwd L1", 4096, 256, 512), ("fwd L2", 4096, 128, 256),
              ("dgrad L1", 4096, 512, 256), ("wgrad L0 (split-K 8)", 512, 508, 4096),
              ("wgrad L1 (split-K 8)", 256, 513, 4096), ("big", 8192, 8192, 8192), ("mmoe expert", 4096, 1024, 20100)]
    for name, M, Nn, K in shapes:
        A = torch.randn(M, _pad8(K), device=dev).to(torch.bfloat16)
        B = torch.randn(Nn, _pad8(K), device=dev).to(torch.bfloat16)
        wgrad = name.startswith("wgrad")
        if wgrad:
            o32 = torch.zeros(M, Nn, device=dev)
            fn = lambda: gemm_bf16_tn(A, B, M, Nn, K, out_f32=o32, accumulate=True, splitk=8)
        else:
            ob = torch.zeros(M, _pad8(Nn), dtype=torch.bfloat16, device=dev)
            ot = torch.zeros(Nn, _pad8(M), dtype=torch.bfloat16, device=dev)
            bias = torch.zeros(Nn, device=dev)
            fn = lambda: gemm_bf16_tn(A, B, M, Nn, K, bias=bias, out_bf16=ob, out_t=ot, relu=True)
        us = timeit(fn, flush=flush if M * K < (1 << 26) else None)
        Ak, Bk = A[:, :K].contiguous(), B[:, :K].contiguous()
        us_lib = timeit(lambda: torch.matmul(Ak, Bk.t()), flush=flush if M * K < (1 << 26) else None)
        fl = 2.0 * M * Nn * K
        out.append({"kernel": "gemm_tcgen05", "case": name, "M": M, "N": Nn, "K": K, "us": round(us, 2),
                    "tflops": round(fl / us / 1e6, 1), "frac_of_peak": round(fl / us / 1e6 / tf, 3),
                    "cublas_us": round(us_lib, 2), "vs_cublas": round(us_lib / us, 2), "peak_src": src})
        print(out[-1], flush=True)
        if M * K < (1 << 26):
            # in-kernel timeline of CTA (0,0,0), cold caches (debug stamps; not a timing of record)
            lib = N.cuda()
            N.check(lib.pbx_gemm_debug_enable(1))
            flush.zero_()
            fn()
            ts = (C.c_uint64 * 8)()
            N.check(lib.pbx_gemm_debug_ts(ts))
            N.check(lib.pbx_gemm_debug_enable(0))
            t = [int(x) for x in ts]
            names = ["setup", "first stage landed", "mainloop issue", "accumulator complete", "epilogue regs->smem",
                     "stores issued", "exit"]
            tl = {n: round((t[i + 1] - t[i]) / 1e3, 2) for i, n in enumerate(names)}
            out.append({"kernel": "gemm_timeline_us", "case": name, **tl})
            print(out[-1], flush=True)


def bench_big_tiles(out):
    """large GEMMs (MMoE wide expert layer): tile width 128 vs 256."""
    from paddlebox_b200.ops.mlp import _pad8, gemm_bf16_tn
    hbm, tf, src = peaks()
    dev = torch.device("cuda")
    for name, M, Nn, K in [("big", 8192, 8192, 8192), ("mmoe wide fwd", 4096, 8192, 20100),
                           ("mmoe wide dgrad", 4096, 20100, 8192), ("mmoe wide wgrad", 8192, 20100, 4096)]:
        A = torch.randn(M, _pad8(K), device=dev).to(torch.bfloat16)
        B = torch.randn(Nn, _pad8(K), device=dev).to(torch.bfloat16)
        ob = torch.zeros(M, _pad8(Nn), dtype=torch.bfloat16, device=dev)
        for bn in (128, 256):
            us = timeit(lambda: gemm_bf16_tn(A, B, M, Nn, K, out_bf16=ob, block_n=bn), iters=5, warmup=2)
            fl = 2.0 * M * Nn * K
            out.append({"kernel": "gemm_big", "case": name, "bn": bn, "us": round(us, 1),
                        "tflops": round(fl / us / 1e6, 1), "frac_of_peak": round(fl / us / 1e6 / tf, 3)})
            print(out[-1], flush=True)
        us = timeit(lambda: torch.matmul(A, B.t()), iters=5, warmup=2)
        print({"case": name, "cublas_us": round(us, 1), "cublas_tflops": round(2.0 * M * Nn * K / us / 1e6, 1)}, flush=True)


def bench_fwd_sweep(out, flush):
    """forward / dgrad shapes of the DeepFM tower: tile width sweep (run under PBX_GEMM_PIPE=deep|shallow)."""
    from paddlebox_b200.ops.mlp import _pad8, gemm_bf16_tn
    dev = torch.device("cuda")
    pipe = os.environ.get("PBX_GEMM_PIPE", "auto")
    for name, M, Nn, K in [("fwd L0", 4096, 512, 507), ("fwd L1", 4096, 256, 512), ("fwd L2", 4096, 128, 256),
                           ("dgrad L1", 4096, 512, 256), ("dgrad L2", 4096, 256, 128)]:
        A = torch.randn(M, _pad8(K), device=dev).to(torch.bfloat16)
        B = torch.randn(Nn, _pad8(K), device=dev).to(torch.bfloat16)
        ob = torch.zeros(M, _pad8(Nn), dtype=torch.bfloat16, device=dev)
        ot = torch.zeros(Nn, _pad8(M), dtype=torch.bfloat16, device=dev)
        bias = torch.zeros(Nn, device=dev)
        for bn in (32, 64, 128):
            if bn > Nn:
                continue
            us = timeit(lambda: gemm_bf16_tn(A, B, M, Nn, K, bias=bias, out_bf16=ob, out_t=ot, relu=True, block_n=bn),
                        flush=flush)
            out.append({"kernel": "gemm_fwd_sweep", "pipe": pipe, "case": name, "bn": bn,
                        "ctas": ((M + 127) // 128) * ((Nn + bn - 1) // bn), "us": round(us, 2)})
            print(out[-1], flush=True)


def bench_wgrad_sweep(out, flush):
    """wgrad shapes of the DeepFM tower: tile width x split-K sweep (atomics contention vs main-loop length)."""
    from paddlebox_b200.ops.mlp import _pad8, gemm_bf16_tn
    dev = torch.device("cuda")
    for name, M, Nn, K in [("wgrad L0", 512, 508, 4096), ("wgrad L1", 256, 513, 4096), ("wgrad L2", 128, 257, 4096),
                           ("dgrad L0 f32", 4096, 507, 512)]:
        A = torch.randn(M, _pad8(K), device=dev).to(torch.bfloat16)
        B = torch.randn(Nn, _pad8(K), device=dev).to(torch.bfloat16)
        o32 = torch.zeros(M, Nn, device=dev)
        for bn in (32, 64, 128):
            for sk in (1, 2, 4, 8, 16):
                if name.startswith("dgrad") and sk > 1:
                    continue
                fn = lambda: gemm_bf16_tn(A, B, M, Nn, K, out_f32=o32, accumulate=sk > 1 or name.startswith("wgrad"),
                                          splitk=sk, block_n=bn)
                us = timeit(fn, flush=flush)
                ctas = ((M + 127) // 128) * ((Nn + bn - 1) // bn) * sk
                out.append({"kernel": "gemm_sweep", "case": name, "bn": bn, "splitk": sk, "ctas": ctas, "us": round(us, 2)})
                print(out[-1], flush=True)


def bench_sparse(out, flush, keys_per_gpu=125_000_000, dim=16, B=4096, S=26, max_bag=5):
    from paddlebox_b200.ops.seqpool import SeqpoolAttrs, seqpool_cvm_forward
    from paddlebox_b200.ops.sparse_table import SparseTable, synthetic_key
    from paddlebox_b200.parallel.sparse_engine import SparseEngine
    hbm, tf, src = peaks()
    dev = torch.device("cuda")
    t = SparseTable(int(keys_per_gpu / 0.6), dim, "adagrad")
    t.fill_synthetic(keys_per_gpu, key_seed=1)
    eng = SparseEngine(t, B * S * max_bag)
    g = torch.Generator().manual_seed(0)
    batches = []
    for _ in range(8):
        lens = torch.randint(1, max_bag + 1, (S * B,), generator=g)
        lod = torch.zeros(S * B + 1, dtype=torch.int32)
        lod[1:] = torch.cumsum(lens, 0)
        n = int(lod[-1])
        u = torch.rand(n, generator=g, dtype=torch.float64)
        keys = synthetic_key((u * u * keys_per_gpu).long().clamp_(max=keys_per_gpu - 1), 1)
        batches.append((keys.to(dev), lod.to(dev), n, keys.unique().numel()))
    H = 3 + dim
    it = {"i": 0}
    restore = torch.empty(B * S * max_bag, dtype=torch.int32, device=dev)

    def pull():
        k, l, n, u = batches[it["i"] % 8]
        it["i"] += 1
        eng.pull(k, restore)
    us = timeit(pull, flush=flush)
    n_avg = sum(b[2] for b in batches) / 8
    u_avg = sum(b[3] for b in batches) / 8
    # algorithmic bytes: keys in (8N) + restore out (4N) + per unique key: key probe (8) + row read
    # (pull_stride*4) + value write (pull_stride*4) + dedup table traffic (~16/key)
    PS = eng.pull_stride * 4
    by = 12 * n_avg + u_avg * (8 + 2 * PS + 16)
    out.append({"kernel": "k_sparse_pull", "keys": int(n_avg), "unique": int(u_avg), "us": round(us, 2),
                "GBps": round(by / us / 1e3, 1), "frac_of_hbm": round(by / us / 1e3 / hbm, 3), "peak_src": src})
    print(out[-1], flush=True)
    attrs = SeqpoolAttrs()
    k, l, n, u = batches[0]
    r0 = eng.pull(k)
    pooled = torch.empty(B, S * H, device=dev)
    us = timeit(lambda: seqpool_cvm_forward(eng.values, l, B, S, attrs, H, restore=r0, out=pooled), flush=flush)
    by = n * (4 + PS) + B * S * H * 4 + (S * B + 1) * 4
    out.append({"kernel": "k_seqpool_cvm_fwd_fast", "us": round(us, 2), "GBps": round(by / us / 1e3, 1),
                "frac_of_hbm": round(by / us / 1e3 / hbm, 3)})
    print(out[-1], flush=True)
    dout = torch.randn(B, S * H, device=dev)
    cvm = torch.ones(B, 2, device=dev)

    def push():
        k, l, n, u = batches[0]
        eng.push_pooled(r0, dout, cvm, l, B, S, H, 2, 0, -float(B))
    us = timeit(push, flush=flush)
    PD = eng.push_stride * 4
    RS = t.layout.stride * 4
    by = n * 4 + B * S * H * 4 + u * (2 * PD + 2 * RS + PD)
    out.append({"kernel": "k_sparse_push", "keys": n, "unique": u, "us": round(us, 2),
                "GBps": round(by / us / 1e3, 1), "frac_of_hbm": round(by / us / 1e3 / hbm, 3)})
    print(out[-1], flush=True)
    eng.check_status()
    # per-phase wall time inside the two persistent kernels (globaltimer stamps by one thread)
    ts = (C.c_uint64 * 16)()
    flush.zero_()
    pull()
    flush.zero_()
    push()
    N.check(N.cuda().pbx_sparse_phase_times(eng._h, ts, N.stream_ptr()))
    t = [int(x) for x in ts]
    names_pull = ["P0 reset", "P1 dedup+route keys", "P2 restore idx", "P2 clear dedup + barrier", "P3 serve rows + barrier"]
    names_push = ["Q0 init (records go out in k_push_records before this kernel)", "Q1 merge duplicate keys", "Q2 ship + barrier", "Q3 optimizer"]
    ph = {"kernel": "phases_us",
          "pull": {n: round((t[i + 1] - t[i]) / 1e3, 1) for i, n in enumerate(names_pull)},
          "push": {n: round((t[9 + i] - t[8 + i]) / 1e3, 1) for i, n in enumerate(names_push)}}
    out.append(ph)
    print(ph, flush=True)
    # load balance across the CTAs of the persistent kernels: when did each CTA FINISH a phase's work
    # (before waiting at the barrier that ends it)?  first / median / last finisher after the phase began
    import numpy as np
    grid = C.c_int()
    N.check(N.cuda().pbx_sparse_debug_stamps(eng._h, 1, None, C.byref(grid), N.stream_ptr()))
    flush.zero_()
    pull()
    flush.zero_()
    push()
    buf = (C.c_uint64 * (grid.value * 16))()
    N.check(N.cuda().pbx_sparse_debug_stamps(eng._h, 1, buf, C.byref(grid), N.stream_ptr()))
    N.check(N.cuda().pbx_sparse_phase_times(eng._h, ts, N.stream_ptr()))
    t = [int(x) for x in ts]
    st = np.frombuffer(buf, dtype=np.uint64).reshape(grid.value, 16).astype(np.int64)
    imb = {"kernel": "phase_finish_us_over_ctas (first / median / last finisher, since the phase began)"}
    for name, slot, begin in [("P0", 0, t[0]), ("P1", 1, t[1]), ("P2a", 2, t[2]), ("P2b", 3, t[3]), ("P3", 4, t[4]),
                              ("Q0", 8, t[8]), ("Q1", 9, t[9]), ("Q2", 10, t[10]), ("Q3", 11, t[11])]:
        d = (st[:, slot] - begin) / 1e3
        imb[name] = [round(float(d.min()), 1), round(float(np.median(d)), 1), round(float(d.max()), 1)]
    out.append(imb)
    print(imb, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="gemm,sparse")
    ap.add_argument("--json", default="")
    ap.add_argument("--keys-per-gpu", type=int, default=125_000_000)
    a = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = []
    if "gemm" in a.what:
        bench_gemm(out, flush)
    if "big_tiles" in a.what:
        bench_big_tiles(out)
    if "fwd_sweep" in a.what:
        bench_fwd_sweep(out, flush)
    if "wgrad_sweep" in a.what:
        bench_wgrad_sweep(out, flush)
    if "sparse" in a.what:
        bench_sparse(out, flush, a.keys_per_gpu)
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
