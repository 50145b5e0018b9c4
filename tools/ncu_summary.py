#!/usr/bin/env python
"""Markdown table of the metrics that matter from one `ncu --set full` capture (reads `ncu -i X.ncu-rep --page raw --csv`).

    python tools/ncu_summary.py gpurun_out/X.ncu-rep [--stalls]   > profiles/rNN_<kernel>_ncu.md
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.avg.per_second",
    "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
    "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    names, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(names, zip(vals, units)))
        print(f"| metric | value | unit |\n|---|---|---|\n| Kernel Name | {d['Kernel Name'][0]} |  |")
        for m in METRICS:
            if m in d:
                print(f"| {m} | {d[m][0]} | {d[m][1]} |")
        if "--stalls" in sys.argv:
            for n in names:
                if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio"):
                    v = float(d[n][0] or 0)
                    if v >= 0.01:
                        print(f"| {n} | {v:.3f} | warps per issue |")
        print()


if __name__ == "__main__":
    main()
