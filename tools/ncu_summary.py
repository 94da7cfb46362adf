#!/usr/bin/env python
"""Prints the metrics the roofline discussion uses from an .ncu-rep (ncu --set full): `python tools/ncu_summary.py REP [substr ...]`."""
import csv
import subprocess
import sys

DEFAULT = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__cluster", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe", "sm__inst_executed_pipe_tensor", "sm__pipe_fp64", "sm__inst_executed_pipe_fp64",
    "sm__pipe_fma_cycles_active.avg.pct", "sm__pipe_alu_cycles_active.avg.pct", "smsp__issue_active.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "l1tex__data_bank_reads.avg.pct", "l1tex__data_bank_writes.avg.pct",
    "smsp__average_warps_issue_stalled", "smsp__average_warp_latency_issue_stalled",
]
rep = sys.argv[1]
want = sys.argv[2:] or DEFAULT
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
for r in rows[2:]:
    name = r[h.index("Kernel Name")] if "Kernel Name" in h else "?"
    print(f"== {name[:100]}")
    for i, n in enumerate(h):
        if any(n.startswith(w) or (w.startswith("*") and w[1:] in n) for w in want):
            print(f"{n:90s} {u[i]:>10s} {r[i]}")
