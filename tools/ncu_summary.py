#!/usr/bin/env python
"""Summarise an ncu report for profiles/: key raw metrics + per-function attribution.
usage: ncu_summary.py <report.ncu-rep> <nvdisasm -gi sass> <mangled kernel substring> <title> [kernel-name regex when the report holds several]"""
import csv
import io
import os
import subprocess
import sys

rep, sass, kernel, title = sys.argv[1:5]
ksel = ["-k", "regex:" + sys.argv[5]] if len(sys.argv) > 5 else []
raw = subprocess.run(["ncu", "-i", rep] + ksel + ["--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[-1]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]
print(f"# {title}")
print(f"# ncu --set full --clock-control none --import-source on; report {rep.split('/')[-1]} (not committed); one launch")
for h, u, v in zip(hdr, units, vals):
    if h in want:
        print(f"{h},{v} {u}")
stalls = [(h, float(v)) for h, v in zip(hdr, vals) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and v]
stalls.sort(key=lambda kv: -kv[1])
print("stall cycles per issued instruction: " + ", ".join(f"{h.split('stalled_')[1].split('_per_issue')[0]} {v:.2f}" for h, v in stalls[:8]))
src = subprocess.run(["ncu", "-i", rep] + ksel + ["--page", "source", "--csv"], capture_output=True, text=True).stdout
tmp = "/tmp/_ncu_source.csv"
open(tmp, "w").write(src)
tool = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_by_line.py")
out = subprocess.run([sys.executable, tool, tmp, sass, kernel, "12"], capture_output=True, text=True).stdout
print("# attribution by source line (top 12) and by function (share of warp instructions, of thread instructions, of stall samples, average active threads)")
print(out)
