#!/usr/bin/env python
"""Reduce an `ncu --set full` report to the per-launch columns quoted in DESIGN.md / profiles/*_ncu_summary.csv
(base units: bytes, ns, plain counts - bench.py reads the DRAM byte columns).

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1x_ncu_summary.csv
"""
import csv
import io
import subprocess
import sys

COLS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread",
    "smsp__inst_executed.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers",
    "l1tex__t_sector_hit_rate.pct",
    "lts__t_sector_hit_rate.pct",
]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], check=True, capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    name = head.index("Kernel Name")
    idx = [head.index(c) for c in COLS]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Kernel Name"] + ["%s [%s]" % (head[i], units[i]) for i in idx])
        for r in body:
            w.writerow([r[name].split("(")[0]] + [r[i] for i in idx])
    print("wrote", out, len(body), "launches")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
