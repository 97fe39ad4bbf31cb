#!/usr/bin/env python
"""Key figures per kernel from an `ncu --set full` report.  usage: ncu_summary.py report.ncu-rep"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed.avg.per_cycle_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
]
STALL = "smsp__average_warps_issue_stalled_"
TAIL = "_per_issue_active.ratio"
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")], " grid", r[hdr.index("Grid Size")] if "Grid Size" in hdr else "")
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"  {w:62s} {r[i]:>16s} {units[i]}")
    st = [(float(r[i]), hdr[i][len(STALL):-len(TAIL)]) for i in range(len(hdr))
          if hdr[i].startswith(STALL) and hdr[i].endswith(TAIL) and r[i]]
    print("  warp stall cycles per issued instruction:", ", ".join(f"{n} {v:.2f}" for v, n in sorted(st, reverse=True)[:7]))
