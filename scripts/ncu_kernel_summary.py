#!/usr/bin/env python
"""Summary of the first kernel of an ncu report (development aid): duration, registers, instructions per launch, issue
activity, the busiest units, warp-state stalls per issued instruction and the executed-opcode histogram.
usage: ncu_kernel_summary.py <report.ncu-rep> [units of work per CTA]   (second argument: e.g. tents per CTA)"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
per_cta = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
m, u = dict(zip(hdr, vals)), dict(zip(hdr, units))
print("kernel:", m.get("Kernel Name"))
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
for k in keys:
    if k in m:
        print(f"{k:95s} {m[k]} {u[k]}")
grid = float(m["launch__grid_size"])
inst = float(m["smsp__inst_executed.sum"])
print(f"warp instructions per unit of work ({per_cta:g} per CTA): {inst / (grid * per_cta):.1f}")
print("stall cycles per issued instruction (> 0.15):")
for k in hdr:
    if "issue_stalled" in k and "per_issue_active" in k and float(m[k] or 0) > 0.15:
        print(f"  {k.split('stalled_')[1].split('_per')[0]:20s} {float(m[k]):.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, body = rows[1], rows[2:]
S, IE, SRC = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
ti, ts = sum(int(r[IE]) for r in body), sum(int(r[S]) for r in body)
ops = {}
for r in body:
    t = r[SRC].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    o = ops.setdefault(op, [0, 0])
    o[0] += int(r[IE])
    o[1] += int(r[S])
print("opcode: % of executed instructions / % of warp samples")
for op, (i, s) in sorted(ops.items(), key=lambda x: -x[1][0])[:16]:
    print(f"  {op:8s} {i / ti * 100:5.1f} / {s / ts * 100:5.1f}")
