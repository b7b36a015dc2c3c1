#!/usr/bin/env python3
"""Condenses `ncu -i X.ncu-rep --page raw --csv` into the handful of metrics DESIGN.md cites.

usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/NAME.txt
"""
import csv
import subprocess
import sys

KEEP = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct",
    "gpu__dram_throughput.avg.pct", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit", "sm__cycles_elapsed.avg.per_second",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed.sum.per_cycle_elapsed", "sm__inst_executed.sum ", "smsp__inst_executed.sum ",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ",
    "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum ", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum ",
    "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum ", "lts__t_bytes.sum ", "lts__throughput.avg.pct",
    "smsp__average_warps_issue_stalled_", "sm__sass_inst_executed_op_global_st", "smsp__inst_executed_op_global_st",
)


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True,
                         check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print("== kernel: %s" % name)
        for h, u, v in zip(hdr, units, r):
            if any((h + " ").startswith(k) or h.startswith(k.strip()) for k in KEEP):
                print("%-90s %-14s %s" % (h, u, v))


if __name__ == "__main__":
    main()
