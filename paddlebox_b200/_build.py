"""In-tree build of the native libraries.

* ``libpbx_cuda.so`` — every sm_100a kernel (csrc/cuda/*.cu), compiled with
  ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (cross-compiles without a GPU).
* ``libpbx_host.so`` — the host runtime (csrc/host/*.cc): data feed, dataset, channel,
  key collection, CPU table tier, metrics.

Both land in ``paddlebox_b200/lib/`` so they travel with the source tree (no JIT cache).
Only stale objects are rebuilt (mtime of the source and of every header).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = Path(__file__).resolve().parent / "lib"
BUILD_DIR = ROOT / "build"

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "--expt-relaxed-constexpr",
    "-I", str(INCLUDE),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-Wall", "-Wno-unused-function", "-I", str(INCLUDE)]


def _nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def _newest_header_mtime() -> float:
    m = 0.0
    for pat in ("cuda/*.cuh", "cuda/*.h", "host/*.h"):
        for p in CSRC.glob(pat):
            m = max(m, p.stat().st_mtime)
    for p in INCLUDE.glob("*.h"):
        m = max(m, p.stat().st_mtime)
    return m


def _run(cmd: list[str], log: Path | None = None) -> None:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("build command failed:\n  " + " ".join(cmd) + "\n" + r.stdout)


def _build_lib(name: str, sources: list[Path], compiler: list[str], flags: list[str], link: list[str],
               force: bool, verbose: bool) -> Path:
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    obj_dir = BUILD_DIR / name
    obj_dir.mkdir(parents=True, exist_ok=True)
    out = LIB_DIR / f"lib{name}.so"
    hdr_m = _newest_header_mtime()
    jobs = []
    objs = []
    for src in sources:
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        stale = force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_m)
        if stale:
            cmd = compiler + flags + ["-c", str(src), "-o", str(obj)]
            jobs.append((cmd, obj_dir / (src.stem + ".log")))
    if jobs:
        if verbose:
            print(f"[paddlebox_b200] compiling {len(jobs)} file(s) for lib{name}.so", file=sys.stderr)
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _run(*j), jobs))
    if jobs or not out.exists():
        _run(link + ["-o", str(out)] + [str(o) for o in objs])
    return out


def build_cuda(force: bool = False, verbose: bool = True) -> Path:
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build libpbx_cuda.so")
    sources = sorted((CSRC / "cuda").glob("*.cu"))
    flags = NVCC_FLAGS + ["-Xptxas", "-v"]
    return _build_lib("pbx_cuda", sources, [nvcc], flags, [nvcc, "-shared", "-gencode",
                      "arch=compute_100a,code=sm_100a"], force, verbose)


def build_host(force: bool = False, verbose: bool = True) -> Path:
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    sources = sorted((CSRC / "host").glob("*.cc"))
    return _build_lib("pbx_host", sources, [cxx], CXX_FLAGS, [cxx, "-shared", "-pthread", "-ldl"], force, verbose)


def build_all(force: bool = False, verbose: bool = True) -> dict[str, Path]:
    out = {"host": build_host(force, verbose)}
    if _nvcc() is not None:
        out["cuda"] = build_cuda(force, verbose)
    return out


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv)
    for k, v in libs.items():
        print(k, v)
