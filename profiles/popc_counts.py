#!/usr/bin/env python
"""POPC / LOP3 / VIMNMX / IMAD warp-instruction counts of the match kernels from an ncu report's source page
(per-SASS-instruction "Instructions Executed"; needs --import-source on at capture time).

    python profiles/popc_counts.py gpurun_out/r2x_full.ncu-rep > profiles/r2x_popc.txt
"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter


def kernel_counts(rep, kernel):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass",
                          "--kernel-name", f"regex:{kernel}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, counts, seen_hdr = None, Counter(), 0
    for r in rows:
        if not r:
            continue
        if "Instructions Executed" in r and "Source" in r:
            seen_hdr += 1
            if seen_hdr > 1:
                break  # second launch of the same kernel
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        try:
            n = int(r[hdr.index("Instructions Executed")] or 0)
        except ValueError:
            continue
        op = re.sub(r"^@!?U?P[0-9T]+\s+", "", r[hdr.index("Source")].strip()).split(" ")[0].split(".")[0]
        counts[op] += n
    return counts


def main(rep):
    for kernel, label in (("knn2_kernel", "brute-force 2-NN, 128 pairs of 1000 x 1000 (radius 0)"),
                          ("knn2_window_kernel", "windowed 2-NN, 128 pairs, r = 32 px")):
        c = kernel_counts(rep, kernel)
        tot = sum(c.values())
        if not tot:
            print(f"{kernel}: not in the report")
            continue
        print(f"{kernel} — {label}")
        print(f"  warp-instructions executed: {tot}")
        for op, n in c.most_common(12):
            print(f"    {op:10s} {n:12d}  {100.0 * n / tot:5.1f} %")
        popc = c.get("POPC", 0)
        print(f"  POPC lane-operations (x32): {popc * 32}")
        if kernel == "knn2_kernel":
            pairs = 128 * 1000 * 1000
            print(f"  POPC per (query, train) pair: {popc * 32 / pairs:.2f} issued (SURVEY 8d counts 8 per pair algorithmically)")


if __name__ == "__main__":
    main(sys.argv[1])
