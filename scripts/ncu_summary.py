#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into markdown for profiles/.

usage: scripts/ncu_summary.py gpurun_out/prof.ncu-rep [title] > profiles/xyz.md
"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg",
]


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    print(f"# {title}\n")
    print("Captured with `ncu --set full --clock-control none --import-source on` on one B200 (numbers under the "
          "profiler are for shares and traffic, not bench values).\n")
    for r in rows[2:]:
        print(f"## {r[h.index('Kernel Name')]}\n")
        print("| metric | value | unit |\n|---|---|---|")
        for w in WANT:
            if w in h:
                print(f"| {w} | {r[h.index(w)]} | {units[h.index(w)]} |")
        print()


if __name__ == "__main__":
    main()
