#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of the shipped CUDA library (no GPU needed):

    python benchmarks/sass_summary.py > profiles/sass_summary.txt

Proves which hardware paths each kernel uses: UTCHMMA = tcgen05.mma, UTMALDG = TMA load,
LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier, RED/ATOM = reductions/atomics,
MEMBAR/ERRBAR = fences, LDG.E.128 = 16-byte global loads."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "paddlebox_b200", "lib", "libpbx_cuda.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "RED", "ATOM",
         "LDG.E.128", "STG.E.128", "LDS.128", "BAR.SYNC", "MEMBAR", "ERRBAR", "SHFL", "HMMA", "FFMA"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"] + names, capture_output=True, text=True, check=True).stdout.split("\n")
        return [re.sub(r"\(anonymous namespace\)::", "", o) for o in out]
    except Exception:
        return names


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1)
            kernels[cur]["_n"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + ".") or (w == "LDG.E.128" and op.startswith("LDG.E.128")) \
                        or (w == "STG.E.128" and op.startswith("STG.E.128")):
                    kernels[cur][w] += 1
    names = list(kernels)
    print("# SASS mnemonic counts per kernel of paddlebox_b200/lib/libpbx_cuda.so (cuobjdump -sass, sm_100a)")
    print("# UTCHMMA = tcgen05.mma, UTMALDG = TMA load, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier")
    for raw, nice in zip(names, demangle(names)):
        c = kernels[raw]
        short = re.sub(r"\(.*", "", nice)
        print(f"{short}\n    instrs={c['_n']} " + str({k: v for k, v in c.items() if k != '_n'}))


if __name__ == "__main__":
    sys.exit(main())
