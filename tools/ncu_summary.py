#!/usr/bin/env python3
"""Condense an `ncu --page raw --csv` export into one line per kernel launch (the figures quoted in
profiles/README.md and DESIGN.md). Usage: python tools/ncu_summary.py gpurun_out/<tag>/pipeline_full_raw.csv"""
import csv
import sys

KEYS = [("ms", "gpu__time_duration.sum"), ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        ("fma%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        ("lsu%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        ("l1%", "l1tex__throughput.avg.pct_of_peak_sustained_active"),
        ("l2%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("dram%", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("rdMB", "dram__bytes_read.sum"), ("wrMB", "dram__bytes_write.sum"),
        ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("lanes", "smsp__thread_inst_executed_per_inst_executed.ratio"), ("regs", "launch__registers_per_thread"),
        ("inst", "smsp__inst_executed.sum")]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0]
        parts = []
        for k, m in KEYS:
            if m in col:
                v = r[col[m]].replace(",", "")
                u = units[col[m]]
                try:
                    x = float(v)
                    if k in ("rdMB", "wrMB"):
                        x *= {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6}.get(u, 1.0)
                    if k == "ms":
                        x *= {"s": 1e3, "ms": 1.0, "us": 1e-3, "ns": 1e-6}.get(u, 1.0)
                    v = f"{x:.3g}" if k != "inst" else f"{x:.4g}"
                except ValueError:
                    pass
                parts.append(f"{k}={v}")
        print(f"{name:22s} " + " ".join(parts))


if __name__ == "__main__":
    main(sys.argv[1])
