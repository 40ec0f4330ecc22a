#!/usr/bin/env python
"""Prints selected metrics of an .ncu-rep (first captured kernel) as CSV: tools/ncu_summary.py rep [substr ...]"""
import csv
import subprocess
import sys

DEFAULT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
           "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_uniform", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "smsp__inst_executed.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
           "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "lts__t_sectors_srcunit_tex_op_read.sum"]


def main():
    rep, extra = sys.argv[1], sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units, vals = rows[0], rows[1], rows[2]
    print("metric,unit,value")
    for k, u, v in zip(head, units, vals):
        if any(k == d or (d in k and not d.startswith("Kernel")) for d in DEFAULT) or any(e in k for e in extra):
            print(f"{k},{u},{v}")


if __name__ == "__main__":
    main()
