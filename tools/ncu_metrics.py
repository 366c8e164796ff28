#!/usr/bin/env python
"""prints the metrics that matter of every kernel in an .ncu-rep (ncu -i <rep> --page raw --csv), one block per launch"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum",
        "sm__inst_executed_pipe_xu.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum", "pcie__read_bytes.sum", "pcie__write_bytes.sum"]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    name_col = hdr.index("Kernel Name")
    for r in rows[2:]:
        print("==", r[name_col][:60])
        for w in WANT:
            if w in hdr:
                print(f"  {w:75s} {r[hdr.index(w)]}")
        stalls = sorted(((float(r[i].replace(',', '')) if r[i] not in ('', 'n/a') else 0.0, h) for i, h in enumerate(hdr)
                         if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")), reverse=True)
        for v, h in stalls[:7]:
            print(f"  stall {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:40s} {v:.2f}")


if __name__ == "__main__":
    main()
