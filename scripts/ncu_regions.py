#!/usr/bin/env python
"""Summarise an ncu report of solve_kernel: headline metrics, opcode mix, and stall samples per execution-count
region (instructions executed equally often belong to the same straight-line region of the warp loop).
usage: scripts/ncu_regions.py gpurun_out/solve_x.ncu-rep"""
import collections, csv, io, re, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.per_cycle_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread"]
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h:75s} {vals[i]} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; data = rows[2:]
c = {h: i for i, h in enumerate(hdr)}
iS, iE, iN = c["Source"], c["Instructions Executed"], c["# Samples"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
regions = collections.OrderedDict()
for r in data:
    regions.setdefault(int(r[iE]), []).append(r)
total = sum(int(r[iN]) for r in data)
maxe = max(regions)
print(f"total samples {total}; region = instructions sharing an execution count (passes ~ {maxe})")
for e, rs in sorted(regions.items(), key=lambda kv: -sum(int(r[iN]) for r in kv[1]))[:8]:
    n = sum(int(r[iN]) for r in rs)
    ops = collections.Counter()
    for r in rs:
        m = re.match(r"\s*(@!?U?P\w+\s+)?([A-Z0-9_]+)", r[iS]); ops[m.group(2) if m else "?"] += 1
    fp64 = sum(v for k, v in ops.items() if k in ("DADD", "DMUL", "DFMA", "DSETP"))
    st = collections.Counter()
    for r in rs:
        for h in stalls: st[h] += int(r[c[h]])
    top = ", ".join(f"{k[6:]} {100*v/max(n,1):.0f}%" for k, v in st.most_common(5))
    print(f"exec {e:9d} ({e/maxe:5.2f}/pass) instrs {len(rs):5d} fp64 {fp64:4d} samples {100*n/total:5.1f}%  samples/fp64 {n/max(fp64,1):6.0f} | {top}")
    print("      ops:", dict(ops.most_common(10)))
