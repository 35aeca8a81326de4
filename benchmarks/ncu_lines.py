#!/usr/bin/env python
"""Warp-stall samples of one kernel aggregated by CUDA source line (needs -lineinfo):
    python benchmarks/ncu_lines.py gpurun_out/prof.ncu-rep k_sparse_pull [top_n]
Reads `ncu --page source --print-source cuda,sass --csv`, sums the SASS rows under every source line."""
import csv
import io
import subprocess
import sys
from collections import defaultdict


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                          "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, fname, cur = None, "?", None
    agg = defaultdict(lambda: defaultdict(int))
    text = {}
    seen_kernel = 0
    for r in rows:
        if not r:
            continue
        if r[0] == "Kernel Name":
            seen_kernel += 1
            if seen_kernel > 1:
                break          # first matching launch only
            continue
        if r[0] in ("File Path", "Function Name", "Module Name"):
            continue
        if r[0] == "File Name":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = {n: i for i, n in enumerate(r)}
            stall_cols = [(n, i) for i, n in enumerate(r) if n.startswith("stall_") and "Not Issued" not in n]
            samp_col = r.index("# Samples")
            continue
        if hdr is None:
            continue
        if r[0] != "":
            try:
                cur = (fname, int(r[0]))
            except ValueError:
                continue
            text[cur] = r[1].strip()
            continue
        if cur is None or len(r) <= samp_col:
            continue
        try:
            agg[cur]["samples"] += int(r[samp_col])
        except ValueError:
            continue
        for n, i in stall_cols:
            try:
                agg[cur][n] += int(r[i])
            except ValueError:
                pass
    tot = sum(v["samples"] for v in agg.values()) or 1
    print(f"# {kern}: {tot} warp-stall samples; top {top} source lines")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        st = sorted(((c, n[6:]) for n, c in v.items() if n != "samples" and c), reverse=True)[:3]
        print(f"{100 * v['samples'] / tot:5.1f}%  {key[0]}:{key[1]:<4d} {text.get(key, '')[:100]:<100s} {st}")


if __name__ == "__main__":
    main()
