#!/usr/bin/env python
"""Condenses one kernel of an .ncu-rep (ncu --set full) into the metric,unit,value CSV kept under
profiles/:   python tools/ncu_summary.py report.ncu-rep "header comment" > profiles/xyz.csv"""
import csv
import subprocess
import sys

KEEP = ("Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_static",
        "launch__occupancy_limit_registers", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio")
PREFIX = ("sm__inst_executed_pipe_", "sm__pipe_", "smsp__average_warps_issue_stalled_")


def main():
    rep, note = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    names, units, vals = rows[0], rows[1], rows[2]
    w = csv.writer(sys.stdout, lineterminator="\n")
    print("# " + note)
    w.writerow(["metric", "unit", "value"])
    for n, u, v in zip(names, units, vals):
        if n in KEEP or (n.startswith(PREFIX) and ".avg" in n and "pct_of_peak_sustained_active" in n) \
                or (n.startswith(PREFIX[2]) and n.endswith("per_issue_active.ratio")):
            w.writerow([n, u, v])


if __name__ == "__main__":
    main()
