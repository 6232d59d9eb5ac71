#!/usr/bin/env python
"""Key metrics of the first kernel in an .ncu-rep, as short text (the summaries under profiles/ were made with this).

    python bench/ncu_key.py gpurun_out/<name>.ncu-rep > profiles/<name>.txt

Reads the report with ``ncu -i <rep> --page raw --csv`` and prints duration, pipe utilisation, issue-slot use,
occupancy, memory throughput, shared-memory bank conflicts, DRAM bytes and the top stall reasons per issue.
"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]


def main():
    if len(sys.argv) != 2:
        sys.exit(__doc__)
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    if len(rows) < 3:
        sys.exit("no kernel in the report")
    d = dict(zip(rows[0], zip(rows[1], rows[2])))
    for k in KEYS:
        if k in d:
            print(k, d[k][1], d[k][0])
    stalls = []
    for name, (_unit, val) in d.items():
        if name.startswith("smsp__average_warps_issue_stalled") and name.endswith("per_issue_active.ratio") and val:
            short = name.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")
            stalls.append((float(val.replace(",", "")), short))
    for v, name in sorted(stalls, reverse=True)[:7]:
        print(round(v, 2), name)


if __name__ == "__main__":
    main()
