#!/usr/bin/env python
"""Instruction and stall-sample shares of the scan kernel by source region (markers are searched in the source).
usage: python tools/ncu_regions.py REPORT.ncu-rep"""
import re
import subprocess
import sys

src = open("tatajuba_b200/csrc/tjb_scan.cuh").read().split("\n")
marks = [("helpers", "__device__ __forceinline__ int dp4a_us"), ("setup", "tjb_scan_kernel (const uint8_t"),
         ("stage", "auto stage_tile = "), ("flush", "auto flush = "), ("phaseB", "auto process_batch = "),
         ("loop", "---- main loop"), ("pass1", "---- pass 1"), ("stage_next", "the derived arrays are complete"),
         ("pass2", "---- pass 2"), ("compact", "---- compaction into"), ("loop_tail", "n_old = tail - head; // what")]
starts = []
for name, m in marks:
    ln = next(i + 1 for i, l in enumerate(src) if m in l)
    starts.append((ln, name))
starts.sort()
out = subprocess.run([sys.executable, "tools/ncu_lines.py", sys.argv[1], "tjb_scan", "1000"], capture_output=True, text=True).stdout
tot, oth = {n: [0.0, 0.0] for _, n in starts}, {}
for l in out.split("\n"):
    m = re.match(r"\s*([\d.]+)% inst\s+([\d.]+)% smp\s+(\S+):(\d+)", l)
    if not m:
        continue
    i, s, f, ln = float(m[1]), float(m[2]), m[3], int(m[4])
    if f == "tjb_scan.cuh":
        name = [n for a, n in starts if a <= ln]
        t = tot[name[-1]] if name else oth.setdefault("head", [0.0, 0.0])
    else:
        t = oth.setdefault(f, [0.0, 0.0])
    t[0] += i
    t[1] += s
print(out.split("\n")[0])
for k, v in list(tot.items()) + list(oth.items()):
    print(f"{k:34s} inst {v[0]:5.1f}%  smp {v[1]:5.1f}%")
