#!/usr/bin/env python
"""Key figures of one kernel from an .ncu-rep (raw page): time, launch shape, pipe utilisation, stall cycles per issue.
Usage: ncu_brief.py report.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, d = rows[0], rows[2]
g = dict(zip(h, d))
for k in ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
          "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
          "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
          "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
          "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__icc_request_hit_rate.pct",
          "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
          "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum"]:
    if k in g:
        print(f"{k}: {g[k][:120]}")
st = []
for n, v in g.items():
    if n.startswith("smsp__average_warps_issue_stalled_") and "not_issued" not in n:
        try:
            st.append((float(v), n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
        except ValueError:
            pass
print("stall cycles per issue:", ", ".join(f"{n} {v:.2f}" for v, n in sorted(st, reverse=True)[:9]))
