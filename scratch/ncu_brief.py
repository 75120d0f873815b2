#!/usr/bin/env python
"""Short per-kernel digest of an ncu report: time, instructions, issue/pipe utilisation, stalls."""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
def g(r, k):
    return r[h.index(k)] if k in h else "?"
for r in rows[2:]:
    print(g(r, "Kernel Name").split("(")[0], "grid", g(r, "launch__grid_size"), "regs", g(r, "launch__registers_per_thread"))
    print("  us", g(r, "gpu__time_duration.sum"), "warp-inst", g(r, "smsp__inst_executed.sum"), "lanes/inst", g(r, "smsp__thread_inst_executed_per_inst_executed.ratio"))
    print("  issue%", g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), "warps_active%", g(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
          "alu%", g(r, "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"), "fma%", g(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
          "lsu%", g(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"))
    print("  dram rd MB", g(r, "dram__bytes_read.sum"), "wr", g(r, "dram__bytes_write.sum"), "dram%", g(r, "dram__throughput.avg.pct_of_peak_sustained_elapsed"), "L2hit", g(r, "lts__t_sector_hit_rate.pct"))
    st = {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): float(r[i]) for i, k in enumerate(h) if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and r[i]}
    print("  stalls", ", ".join(f"{k} {v:.2f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:7]))
    print("  smem wavefronts", g(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"), "conflicts", g(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"))
