"""Builds the reference arm's native library, ``baseline/_ref/libref_heterps.so``.

The reference (mark914/PaddleBox) cannot be pip-installed or CMake-built offline (DESIGN.md
"Reference arm").  What CAN be built is its open sparse engine: the HeterPS sources are copied
UNMODIFIED from /root/reference into ``baseline/_ref/src`` (git-ignored; sha256 of every copy is
recorded in ``baseline/_ref/MANIFEST.json`` and checked against the originals) and compiled against
the thin shim headers in ``baseline/shims`` — one small header per Paddle-internal include — plus
the driver in ``baseline/harness``.  The SIMT kernels are compiled for sm_100 as they are.

    python baseline/build_ref.py            # copy (if /root/reference is present) + compile
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path(os.environ.get("PBX_REFERENCE_ROOT", "/root/reference"))
OUT = HERE / "_ref"
SRC = OUT / "src"
HPS = "paddle/fluid/framework/fleet/heter_ps"
FILES = [
    "heter_comm.h", "heter_comm_inl.h", "heter_comm_kernel.h", "heter_comm_kernel.cu", "hashtable.h",
    "hashtable_kernel.cu", "feature_value.h", "feature_value.cu", "optimizer.cuh.h", "optimizer_conf.h",
    "heter_resource.h", "heter_resource.cc", "mem_pool.h", "gpu_graph_utils.h", "heter_ps.h", "heter_ps_base.h",
    "heter_ps.cu", "cudf/concurrent_unordered_map.cuh.h", "cudf/hash_functions.cuh", "cudf/managed.cuh",
    "cudf/managed_allocator.cuh", "cudf/LICENSE",
]
# translation units of the reference compiled as they are
# (feature_value.cu is compiled through harness/ref_instantiate.cu, which includes it unmodified)
REF_TUS = ["heter_comm_kernel.cu", "hashtable_kernel.cu", "heter_resource.cc", "heter_ps.cu"]
DEFINES = ["-DPADDLE_WITH_HETERPS", "-DPADDLE_WITH_CUDA", "-DNDEBUG"]   # NDEBUG: a Release build, like Paddle's


def sha(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def copy_sources() -> dict:
    man = {}
    for f in FILES:
        src, dst = REF / HPS / f, SRC / HPS / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        if src.exists():
            shutil.copyfile(src, dst)
            assert sha(src) == sha(dst)
        if not dst.exists():
            raise RuntimeError(f"missing reference source {src} (and no earlier copy at {dst})")
        man[f] = sha(dst)
    (OUT / "MANIFEST.json").write_text(json.dumps({"root": str(REF / HPS), "sha256": man}, indent=1))
    return man


def build(verbose: bool = True) -> Path:
    copy_sources()
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    obj = OUT / "obj"
    obj.mkdir(parents=True, exist_ok=True)
    inc = ["-I", str(HERE / "shims"), "-I", str(SRC)]
    flags = ["-O3", "-std=c++17", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC", "-w",
             "--expt-relaxed-constexpr", "--expt-extended-lambda", "-lineinfo"] + DEFINES + inc
    jobs = []
    for f in REF_TUS:
        jobs.append((SRC / HPS / f, obj / (Path(f).stem + ".o"), ["-x", "cu"]))
    jobs.append((HERE / "harness" / "ref_shim.cc", obj / "ref_shim.o", ["-x", "cu"]))
    jobs.append((HERE / "harness" / "ref_harness.cu", obj / "ref_harness.o", []))
    jobs.append((HERE / "harness" / "ref_instantiate.cu", obj / "ref_instantiate.o", []))

    def run(job):
        src, o, extra = job
        if o.exists() and o.stat().st_mtime > max(src.stat().st_mtime, *(p.stat().st_mtime for p in (HERE / "shims").rglob("*.h"))):
            return
        cmd = [nvcc] + flags + extra + ["-c", str(src), "-o", str(o)]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        (obj / (o.stem + ".log")).write_text(" ".join(cmd) + "\n" + r.stdout)
        if r.returncode != 0:
            raise RuntimeError(f"compile failed: {src}\n{r.stdout[-6000:]}")

    with ThreadPoolExecutor(max_workers=6) as ex:
        list(ex.map(run, jobs))
    lib = OUT / "libref_heterps.so"
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100,code=sm_100", "-o", str(lib)] + [str(j[1]) for j in jobs] + \
          ["-lnccl", "-lcurand", "-lpthread"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout)
    if verbose:
        print("built", lib)
    return lib


if __name__ == "__main__":
    build()
