#!/usr/bin/env python
"""BASELINE config 5, reference side: the sparse pull / push of the reference's own HeterPS (``baseline/_ref``,
compiled unmodified, see baseline/README.md) on the sweep shapes of ``benchmarks/sparse_sweep.py`` — the
denominator of "vs the reference's HeterComm".

    python benchmarks/sparse_sweep_reference.py --gpus 8 --dims 8,16,64,128 --keys 65536,1048576,4194304,16777216 \
        --json profiles/sparse_sweep_reference_8gpu.json

Process model as in ``baseline/reference_arm.py``: ONE process, one worker thread per GPU (the reference's model),
each worker pulls / pushes its own batch of ``keys`` random table keys (one slot, every key its own bag, the same
batch shape as the product sweep).  Per iteration all workers meet at a barrier, each times its call with CUDA
events on its stream; reported is the median over iterations of the max over workers.  One sub-process per
embedding width (the table's value layout depends on it, and tearing a managed-memory table down is slow).
NOT RUN in round 2 (written after the GPU budget was spent); it only uses entry points ``reference_arm.py``
exercises.  Always run it under a ``timeout``.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))


def sweep_one_dim(a) -> int:
    import torch
    import reference_arm as RA
    ndev, dim = a.gpus, a.dim
    if not RA.LIB.exists():
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref/libref_heterps.so is not built"}))
        return 0
    if torch.cuda.device_count() < ndev:
        print(json.dumps({"impl": "reference", "unavailable": f"needs {ndev} GPUs"}))
        return 0
    lib = C.CDLL(str(RA.LIB))
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ref_last_error.restype = C.c_char_p
    lib.ref_build_synthetic.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
    lib.ref_pull_sparse.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.ref_push_sparse.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.ref_set_compute_stream.argtypes = [C.c_int, C.c_void_p]
    for d in range(ndev):
        torch.cuda.set_device(d)
        torch.zeros(1, device=f"cuda:{d}")
    torch.cuda.set_device(0)
    h = lib.ref_create(ndev, dim, 1, 1)
    counts = (C.c_int64 * ndev)()
    n_total = a.table_keys_per_gpu * ndev
    if lib.ref_build_synthetic(h, n_total, 1, 1, counts) != 0:
        print(json.dumps({"impl": "reference", "unavailable": "table build failed: " + lib.ref_last_error().decode()[:200]}))
        return 0
    E = 3 + dim
    keys_list = [int(x) for x in a.keys.split(",")]
    results = {nk: [None] * ndev for nk in keys_list}
    barrier = threading.Barrier(ndev)
    errors: list = []

    def work(d: int) -> None:
        try:
            dev = torch.device("cuda", d)
            torch.cuda.set_device(dev)
            stream = torch.cuda.Stream(device=dev)
            torch.cuda.set_stream(stream)
            lib.ref_set_compute_stream(d, C.c_void_p(stream.cuda_stream))
            g = torch.Generator().manual_seed(d)
            for nk in keys_list:
                idx = (torch.rand(nk, generator=g, dtype=torch.float64) * n_total).long().clamp_(max=n_total - 1)
                keys = RA.synthetic_key(idx, 1).to(dev)
                vals = torch.empty(nk, E, device=dev)
                grads = torch.randn(nk, E, device=dev)
                grads[:, 0] = 1.0
                grads[:, 1] = 0.0
                kp, vp, gp = (C.c_void_p * 1)(keys.data_ptr()), (C.c_void_p * 1)(vals.data_ptr()), (C.c_void_p * 1)(grads.data_ptr())
                sl = (C.c_int64 * 1)(nk)

                def pull():
                    if lib.ref_pull_sparse(h, d, kp, vp, sl, 1, E) != 0:
                        raise RuntimeError(lib.ref_last_error().decode())

                def push():
                    if lib.ref_push_sparse(h, d, gp, sl, 1, E, nk) != 0:
                        raise RuntimeError(lib.ref_last_error().decode())

                t = {"pull": [], "push": []}
                for it in range(a.warmup + a.iters):
                    for name, fn in (("pull", pull), ("push", push)):
                        torch.cuda.synchronize(dev)
                        barrier.wait(timeout=300)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        fn()
                        e1.record()
                        torch.cuda.synchronize(dev)
                        if it >= a.warmup:
                            t[name].append(e0.elapsed_time(e1) * 1e3)
                results[nk][d] = t
                del keys, vals, grads
                torch.cuda.empty_cache()
        except BaseException as e:  # noqa: BLE001
            errors.append(f"worker {d}: {e}")
            barrier.abort()

    ths = [threading.Thread(target=work, args=(d,)) for d in range(ndev)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    if errors:
        print(json.dumps({"impl": "reference", "dim": dim, "unavailable": errors[0][:300]}), flush=True)
        os._exit(0)
    for nk in keys_list:
        rec = {"impl": "reference", "n_gpus": ndev, "keys_per_batch": nk, "dim": dim}
        for name in ("pull", "push"):
            per_iter = [max(results[nk][d][name][i] for d in range(ndev)) for i in range(a.iters)]
            per_iter.sort()
            rec[name + "_us"] = round(per_iter[len(per_iter) // 2], 1)
        print(json.dumps(rec), flush=True)
    os._exit(0)          # skip the destructors of the managed-memory tables


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--dims", default="8,16,64,128")
    ap.add_argument("--dim", type=int, default=0, help="(internal) one width in this process")
    ap.add_argument("--keys", default="65536,1048576,4194304,16777216")
    ap.add_argument("--table-keys-per-gpu", type=int, default=20_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--per-dim-timeout", type=int, default=600)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    if a.dim:
        return sweep_one_dim(a)
    out = []
    for dim in [int(x) for x in a.dims.split(",")]:
        cmd = [sys.executable, os.path.abspath(__file__), "--gpus", str(a.gpus), "--dim", str(dim), "--keys", a.keys,
               "--table-keys-per-gpu", str(a.table_keys_per_gpu), "--iters", str(a.iters), "--warmup", str(a.warmup)]
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=a.per_dim_timeout)
            lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        except subprocess.TimeoutExpired:
            lines = [json.dumps({"impl": "reference", "dim": dim, "unavailable": f"no result within {a.per_dim_timeout} s"})]
        for ln in lines:
            print(ln, flush=True)
            out.append(json.loads(ln))
    if a.json:
        with open(a.json, "w") as f:
            json.dump(out, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
