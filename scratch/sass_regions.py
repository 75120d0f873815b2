#!/usr/bin/env python
"""SASS-level digest of one kernel in an ncu report: splits the instruction stream at BAR / SYNCS /
backward branches and prints executed warp-instructions, avg active threads and stall samples per region.
    python scratch/sass_regions.py rep kernel_regex [min_share]"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", f"regex:{sys.argv[2]}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = None; ins = []
for r in rows:
    if r and r[0] == "Address":
        if hdr is not None: break
        hdr = r; continue
    if hdr and len(r) > 8 and r[0].startswith("0x"):
        ins.append(r)
iA, iS, iN, iI, iT = (hdr.index(k) for k in ("Address", "Source", "# Samples", "Instructions Executed", "Thread Instructions Executed"))
tot = sum(int(r[iI]) for r in ins); tots = sum(int(r[iN]) for r in ins)
print("total warp-inst", tot, "samples", tots, "sass lines", len(ins))
# per-instruction listing compressed: print runs
cur = []; 
def flush():
    global cur
    if not cur: return
    n = sum(int(r[iI]) for r in cur); s = sum(int(r[iN]) for r in cur); th = sum(int(r[iT]) for r in cur)
    if n / tot >= float(sys.argv[3]) if len(sys.argv) > 3 else 0.0:
        ops = {}
        for r in cur:
            op = r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]
            op = op.split(".")[0]
            ops[op] = ops.get(op, 0) + int(r[iI])
        top = ", ".join(f"{k}:{100*v/n:.0f}%" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:6]) if n else ""
        print(f"[{ins.index(cur[0]):4d}-{ins.index(cur[-1]):4d}] inst {100*n/tot:5.1f}%  smp {100*s/max(1,tots):5.1f}%  lanes {th/max(1,n):4.1f}  n={len(cur):3d}  exec/inst={n/len(cur)/1e3:7.1f}k  {top}")
    cur = []
prev = None
for r in ins:
    c = int(r[iI])
    # new region when the execution count changes by more than 20 %
    if prev is not None and (c > prev * 1.25 or c < prev * 0.8):
        flush()
    cur.append(r); prev = c
flush()
