#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals of one kernel of an ncu report.

    python profiles/hot.py gpurun_out/x.ncu-rep fast_kernel [top_n]

Uses `ncu -i <rep> --page source --csv --print-source cuda,sass` (needs -lineinfo at compile time and
--import-source on at capture time).  Only the first captured launch of the kernel is summarised.
"""
import csv
import io
import subprocess
import sys


def main(rep, kernel, top=40):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", f"regex:{kernel}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, out, files, cur = None, {}, 0, ""
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            if cur in out or any(k[0] == cur for k in out):
                break  # second launch of the same kernel
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or not r[0].isdigit():
            continue
        i_inst, i_smp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        try:
            inst, smp = int(r[i_inst] or 0), int(r[i_smp] or 0)
        except ValueError:
            continue
        key = (cur, int(r[0]))
        a = out.setdefault(key, [0, 0, r[1][:100]])
        a[0] += inst
        a[1] += smp
    ti = sum(v[0] for v in out.values()) or 1
    ts = sum(v[1] for v in out.values()) or 1
    print(f"{kernel}: warp-instructions {ti}, samples {ts}")
    for (f, ln), (inst, smp, src) in sorted(out.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{100*smp/ts:5.1f}% smp {100*inst/ti:5.1f}% inst  {f}:{ln}: {src.strip()}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
