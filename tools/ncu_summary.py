#!/usr/bin/env python3
"""Summarise an ncu report (.ncu-rep) into a small CSV: one row per profiled launch with the metrics the
roofline discussion needs.  Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/rN/x_summary.csv"""
import csv
import subprocess
import sys

KEYS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "launch__registers_per_thread", "lts__t_sector_hit_rate.pct"]


def main():
    raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [(k, hdr.index(k)) for k in KEYS if k in hdr]
    w = csv.writer(sys.stdout)
    w.writerow([k + (f" [{units[i]}]" if units[i] else "") for k, i in idx])
    for r in rows[2:]:
        out = []
        for k, i in idx:
            v = r[i]
            if k == "Kernel Name":
                v = v.split("(")[0].replace("<unnamed>::", "")
            out.append(v)
        w.writerow(out)


if __name__ == "__main__":
    main()
