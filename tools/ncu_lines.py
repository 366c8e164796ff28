#!/usr/bin/env python
"""instructions executed per CUDA source line of a kernel in an .ncu-rep captured with --import-source on
(ncu -i <rep> --page source --csv --print-source cuda,sass): top lines and sums per line range"""
import csv
import subprocess
import sys
from collections import defaultdict


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    only = sys.argv[3] if len(sys.argv) > 3 else None  # substring of the kernel name
    per = defaultdict(lambda: [0, 0, ""])  # (file, line) -> [instructions, samples, text]
    fname, hdr, kernel = None, None, ""
    for r in rows:
        if len(r) >= 2 and r[0] in ("Kernel Name", "Function Name"):
            kernel = r[1]
            continue
        if only and only not in kernel:
            continue
        if len(r) >= 2 and r[0] == "File Name":
            fname = r[1].split("/")[-1]
            continue
        if len(r) > 7 and r[0] == "Line No":
            hdr = r
            ie, ns = hdr.index("Instructions Executed"), hdr.index("# Samples")
            continue
        if hdr and len(r) == len(hdr):
            try:
                line = int(r[0])
            except ValueError:
                continue
            key = (fname, line)
            if r[1].strip():
                per[key][2] = r[1].strip()
            try:
                per[key][0] += int(r[ie] or 0)
                per[key][1] += int(r[ns] or 0)
            except ValueError:
                pass
    total = sum(v[0] for v in per.values())
    tsamp = sum(v[1] for v in per.values())
    print(f"total instructions {total}  samples {tsamp}")
    for (f, l), v in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{100 * v[0] / total:5.1f}% inst {100 * v[1] / max(tsamp, 1):5.1f}% smp  {f}:{l:<4d} {v[2][:110]}")


if __name__ == "__main__":
    main()
