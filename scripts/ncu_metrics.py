#!/usr/bin/env python
"""Key metrics of every kernel in an ncu report as one CSV (metric, unit, one column per kernel).
usage: scripts/ncu_metrics.py gpurun_out/x.ncu-rep > profiles/rN_ncu_x_metrics.csv"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
KEEP = ("Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma", "sm__inst_executed_pipe_alu.avg.pct",
        "sm__inst_executed_pipe_lsu.avg.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "smsp__average_warps_issue_stalled",
        "smsp__average_warp_latency_per_inst_issued", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__sass_average_data_bytes_per_sector_mem_global", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "sm__cycles_elapsed.max")
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [r[hdr.index("Kernel Name")][:48] for r in data])
for i, h in enumerate(hdr):
    if any(h.startswith(k) for k in KEEP) and ".max" not in h and ".min" not in h and ".per_second" not in h and ".pct_of_peak_sustained_elapsed" not in h[len("dram__bytes"):] if h.startswith("dram__bytes") else any(h.startswith(k) for k in KEEP) and ".min" not in h and not h.endswith(".max.pct_of_peak_sustained_active"):
        w.writerow([h, units[i]] + [r[i] for r in data])
