#!/usr/bin/env python
"""ncu_summary.py <report.ncu-rep> [kernel-index] -- print the handful of ncu metrics the roofline notes cite."""
import csv
import re
import subprocess
import sys

rep = sys.argv[1]
idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
d, u = dict(zip(hdr, data[idx])), dict(zip(hdr, units))
pats = [r"^Kernel Name$", r"^Grid Size$", r"^Block Size$", r"^gpu__time_duration.sum$", r"^dram__bytes_(read|write).sum$",
        r"^gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed$", r"^lts__throughput.avg.pct_of_peak_sustained_elapsed$",
        r"^l1tex__throughput.avg.pct_of_peak_sustained_elapsed$", r"^sm__throughput.avg.pct_of_peak_sustained_elapsed$",
        r"^sm__warps_active.avg.pct_of_peak_sustained_active$", r"^sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active$",
        r"^sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active$", r"^sm__inst_executed_pipe_(alu|fma|fmaheavy|fp64|xu|uniform).avg.pct_of_peak_sustained_active$",
        r"^smsp__issue_active.avg.pct_of_peak_sustained_active$", r"^smsp__inst_executed.sum$",
        r"^l1tex__data_pipe_lsu_wavefronts_mem_shared.sum$", r"^l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum$",
        r"^l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed$", r"^lts__t_sectors_srcunit_tex_op_(read|red|write).sum$",
        r"^lts__t_sector_hit_rate.pct$", r"^lts__d_atomic_input_cycles_active.avg.pct_of_peak_sustained_elapsed$",
        r"^launch__registers_per_thread$", r"^launch__occupancy_limit_(registers|shared_mem|warps|blocks)$", r"^sm__cycles_elapsed.max$",
        r"^launch__shared_mem_per_block_dynamic$"]
for k in hdr:
    if any(re.search(p, k) for p in pats):
        print(f"{k} = {d[k]} {u[k]}")
st = [(float(d[k].replace(",", "")), k) for k in hdr
      if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio")]
for v, k in sorted(st, reverse=True)[:8]:
    print(f"{v:8.3f} {k}")
