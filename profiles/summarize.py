#!/usr/bin/env python
"""Extract the counters DESIGN.md quotes from an .ncu-rep (ncu --set full) into a text summary.

usage: python profiles/summarize.py gpurun_out/prof.ncu-rep > profiles/<name>.txt
"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak",
        "smsp__issue_active.avg.pct", "sm__inst_issued.avg.pct", "sm__inst_executed_pipe_alu.avg.pct",
        "sm__inst_executed_pipe_fma", "sm__inst_executed_pipe_lsu", "sm__inst_executed_pipe_uniform",
        "sm__pipe_alu_cycles_active.avg.pct", "sm__pipe_fma_cycles_active.avg.pct", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__throughput.avg.pct", "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__warp_issue_stalled",
        "smsp__average_warp", "sm__sass_thread_inst_executed_op_integer"]

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    print(f"== {name[:110]}  grid={r[hdr.index('Grid Size')]} block={r[hdr.index('Block Size')]}")
    for h, u, v in zip(hdr, units, r):
        if any(h.startswith(w) for w in WANT) and ".min" not in h and ".max.pct" not in h and ".sum.p" not in h:
            print(f"  {h} = {v} {u}")
