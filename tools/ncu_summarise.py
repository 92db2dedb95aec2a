#!/usr/bin/env python3
"""Turns `gpurun_out/prof_*.ncu-rep` into the markdown tables of profiles/ncu_single_gpu_kernels.md.

  python tools/ncu_summarise.py gpurun_out/prof_crc32c_kernel.ncu-rep [...]   # prints one section per report
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warp_latency_issue_stalled_lg_throttle.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
]


def main():
    for path in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        header, units, data = rows[0], rows[1], rows[2:]
        for row in data:
            rec = dict(zip(header, row))
            unit = dict(zip(header, units))
            print(f"## {rec.get('Kernel Name', '?')}\n\n| metric | value |\n|---|---|")
            for m in METRICS:
                if m in rec and rec[m] != "":
                    print(f"| {m} | {rec[m]} {unit.get(m, '')} |")
            print()


if __name__ == "__main__":
    main()
