#!/usr/bin/env python
"""Summary CSV of an .ncu-rep (ncu -i ... --page raw --csv): one row per captured launch with the metrics the roofline
arithmetic uses.  Usage: python tools/ncu_summary.py in.ncu-rep [more.ncu-rep ...] > out.csv"""
import csv
import io
import subprocess
import sys

KEEP = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]


def main():
    w = csv.writer(sys.stdout)
    first = True
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]
        cols = [hdr.index(k) for k in KEEP if k in hdr]
        if first:
            w.writerow(["report"] + [hdr[c] for c in cols])
            first = False
        w.writerow(["units of " + rep.split("/")[-1]] + [units[c] for c in cols])  # ncu picks units per report
        for r in rows[2:]:
            w.writerow([rep.split("/")[-1]] + [r[c] for c in cols])


if __name__ == "__main__":
    main()
