#!/usr/bin/env python
"""Markdown summary of one kernel of an ncu report (development tool; feeds profiles/).

  python tools/ncu_summary.py gpurun_out/ncu_transport.ncu-rep [row] > profiles/xxx.md
"""
import csv
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
    "smsp__average_warp_latency_per_inst_issued.ratio",
]


def main():
    rep = sys.argv[1]
    row = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    raw = (open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2 + row]
    d = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
    print(f"kernel: `{d['Kernel Name'][1]}`\n")
    print("| metric | unit | value |\n|---|---|---|")
    for k in KEEP:
        if k in d:
            print(f"| {k} | {d[k][0]} | {d[k][1]} |")
    for k in sorted(d):
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
            try:
                if float(d[k][1]) >= 0.1:
                    print(f"| {k} | {d[k][0]} | {d[k][1]} |")
            except ValueError:
                pass


if __name__ == "__main__":
    main()
