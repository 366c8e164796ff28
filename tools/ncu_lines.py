#!/usr/bin/env python
"""Per-CUDA-source-line instruction counts and stall samples of one kernel from an .ncu-rep (needs -lineinfo and
--import-source on).  usage: python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [top_n]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                      f"regex:{pat}"], capture_output=True, text=True).stdout
rows, fname, first_fn, launches = [], "", None, 0
for r in csv.reader(io.StringIO(txt)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Kernel Name":
        launches += 1
        if launches > 1:
            break  # first matching launch only
        continue
    if r[0] == "Line No" or not r[0].isdigit():
        continue
    try:
        rows.append((fname, int(r[0]), r[1].strip(), int(r[7]), int(r[4])))
    except (ValueError, IndexError):
        pass
tot_i = sum(x[3] for x in rows) or 1
tot_s = sum(x[4] for x in rows) or 1
print(f"# {rep} {pat}: {tot_i} warp instructions, {tot_s} stall samples attributed to source lines")
for f, ln, src, ins, smp in sorted(rows, key=lambda x: -x[3])[:top]:
    print(f"{100 * ins / tot_i:5.1f}% inst {100 * smp / tot_s:5.1f}% smp  {f}:{ln:<4d} {src[:110]}")
