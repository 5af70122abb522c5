#!/usr/bin/env python3
"""Selected raw metrics of every kernel in an ncu report. usage: tools/ncu_raw.py report.ncu-rep [extra_metric ...]"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "smsp__cycles_active.max", "smsp__cycles_active.min", "sm__cycles_elapsed.max",
        "sm__cycles_active.avg", "sm__cycles_active.min", "sm__cycles_active.max",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size"]


def main():
    rep = sys.argv[1]
    want = WANT + sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    kn = h.index("Kernel Name")
    print("kernel", [r[kn][:28] for r in rows[2:]])
    for w in want:
        if w in h:
            i = h.index(w)
            print("%-62s %-10s %s" % (w, rows[1][i], [r[i] for r in rows[2:]]))


if __name__ == "__main__":
    main()
