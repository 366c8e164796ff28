#!/usr/bin/env python
"""Text summary of one kernel of an `ncu --set full` report: duration, DRAM bytes, pipe utilisation, issue and stall
mix, per-source-line hot spots.  usage: python tools/ncu_summary.py REPORT.ncu-rep KERNEL_REGEX > profiles/rNN_x.txt"""
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", f"regex:{pat}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units, v = rows[0], rows[1], rows[2]
col = {n: i for i, n in enumerate(h)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__warps_eligible.avg.per_cycle_active"]
print(f"# {rep}: kernel {v[col['Kernel Name']] if 'Kernel Name' in col else pat} (ncu --set full --clock-control none)")
for n in want:
    if n in col:
        print(f"{n:70s} {v[col[n]]:>16s} {units[col[n]]}")
print("# stall reasons (warps per issue-active cycle)")
st = [(float(v[i]), n) for n, i in col.items() if n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio")]
for x, n in sorted(st, reverse=True)[:8]:
    print(f"{n:90s} {x:8.3f}")
print("# hottest source lines")
sys.stdout.flush()
subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "ncu_lines.py"), rep, pat, "25"])
