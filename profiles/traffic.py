#!/usr/bin/env python
"""DRAM traffic per step and per launch of every stage, from a per-kernel summary CSV (summarize.py).

    python profiles/traffic.py profiles/r1c_full.csv profiles/traffic.json

bench.py reads profiles/traffic.json for `roofline.traffic` (dram__bytes_read.sum + dram__bytes_write.sum of
the dominant kernel, per launch, from the `ncu --set full` capture of the same bench command at batch 128).
"""
import csv
import json
import sys

STAGE_OF = {"resize_band_kernel": "pyramid", "resize_kernel": "pyramid", "apron_kernel": "pyramid", "fast_kernel": "fast",
            "select_kernel": "select", "blur_kernel": "blur", "describe_kernel": "describe", "knn2_window_kernel": "match",
            "knn2_kernel": "match", "match_compact_kernel": "match"}


def main(src, dst):
    rows = list(csv.reader(open(src)))
    h, units = rows[0], rows[1]
    i_name, i_rd, i_wr, i_t = (h.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
    scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
    seen = {}
    # one step = the launches from the first apron_kernel / resize up to the next one: keep the first occurrence
    # of each (kernel, grid) so that a capture spanning more than one step is not double counted
    i_grid = h.index("launch__grid_size")
    for r in rows[2:]:
        name = r[i_name].replace("void ", "").split("<")[0].strip()
        key = (name, r[i_grid])
        if key in seen:
            continue
        seen[key] = (float(r[i_rd]) * scale[units[i_rd]] + float(r[i_wr]) * scale[units[i_wr]], float(r[i_t]))
    out = {}
    for (name, grid), (b, t) in seen.items():
        st = STAGE_OF.get(name)
        if st is None:
            continue
        e = out.setdefault(st, {"dram_bytes_per_step": 0.0, "launches_per_step": 0, "ncu_us_per_step": 0.0})
        e["dram_bytes_per_step"] += b
        e["launches_per_step"] += 1
        e["ncu_us_per_step"] += t
    for e in out.values():
        e["dram_bytes_per_launch"] = e["dram_bytes_per_step"] / e["launches_per_step"]
    out["_source"] = {"csv": src, "batch": 128, "note": "ncu --set full --clock-control none, one resident step after warm-up"}
    json.dump(out, open(dst, "w"), indent=1)
    for k, v in out.items():
        print(k, v)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
