#!/usr/bin/env python3
"""Key metrics per kernel from an .ncu-rep (raw page). usage: python tools/ncu_summary.py rep [out.csv]"""
import csv
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
out = [[w for w, _ in idx], [units[i] for _, i in idx]] + [[r[i] for _, i in idx] for r in rows[2:]]
if len(sys.argv) > 2:
    csv.writer(open(sys.argv[2], "w")).writerows(out)
short = {"Kernel Name": "kernel", "launch__grid_size": "grid", "launch__block_size": "blk", "launch__registers_per_thread": "regs",
         "launch__occupancy_limit_shared_mem": "occ_smem", "launch__occupancy_limit_registers": "occ_reg",
         "gpu__time_duration.sum": "time", "dram__bytes_read.sum": "dram_rd", "dram__bytes_write.sum": "dram_wr",
         "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram%", "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm%",
         "sm__warps_active.avg.pct_of_peak_sustained_active": "warps%", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64%",
         "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor%", "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%",
         "smsp__inst_executed.sum": "inst", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wf",
         "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "bank_conf", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": "l1tex%",
         "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2%", "sm__cycles_elapsed.max": "cycles"}
for r in out[2:]:
    print("  ".join(f"{short[w]}={v[:38]}{u if short[w] in ('time','dram_rd','dram_wr') else ''}" for (w, _), v, u in zip(idx, r, out[1])))
