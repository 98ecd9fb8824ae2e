#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the per-kernel table kept under profiles/
(one row per captured launch, the columns DESIGN.md and bench.py's roofline.traffic refer to).

  ncu -i gpurun_out/prof_full.ncu-rep --page raw --csv > gpurun_out/prof_full_raw.csv
  python tools/ncu_summary.py gpurun_out/prof_full_raw.csv > profiles/rNN_ncu_full_summary.csv
"""
import csv
import sys

COLS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_srcunit_ltcfabric.sum",
        "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait",
        "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_not_selected",
        "smsp__pcsamp_warps_issue_stalled_short_scoreboard", "smsp__pcsamp_warps_issue_stalled_lg_throttle",
        "smsp__pcsamp_warps_issue_stalled_branch_resolving", "smsp__pcsamp_warps_issue_stalled_barrier",
        "smsp__pcsamp_warps_issue_stalled_membar"]

with open(sys.argv[1]) as f:
    rows = list(csv.reader(l for l in f if not l.startswith("==")))
header, units, data = rows[0], rows[1], rows[2:]
idx = {n: i for i, n in enumerate(header)}
keep = [c for c in COLS if c in idx]
out = csv.writer(sys.stdout)
out.writerow(["Kernel Name"] + ["%s [%s]" % (c, units[idx[c]]) if units[idx[c]] else c for c in keep])
for r in data:
    if len(r) != len(header):
        continue
    out.writerow([r[idx["Kernel Name"]]] + [r[idx[c]].replace(",", "") for c in keep])
