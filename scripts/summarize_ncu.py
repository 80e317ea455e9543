#!/usr/bin/env python
"""Turns the ncu outputs a gpurun visit leaves in gpurun_out/ into small text summaries under profiles/.

usage: summarize_ncu.py <tag>   (reads gpurun_out/<tag>_launches.csv and gpurun_out/<tag>_prof_tent.ncu-rep)
"""
import csv
import io
import json
import os
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

KEYS = ["gpu__time_duration.sum", "sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.sum", "sm__sass_thread_inst_executed_op_dfma_pred_on.sum",
        "sm__sass_thread_inst_executed_op_dmul_pred_on.sum", "sm__sass_thread_inst_executed_op_dadd_pred_on.sum",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]

lc = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
if os.path.exists(lc):
    rows = [r for r in csv.reader(l for l in open(lc) if l.startswith('"'))]
    hdr = rows[0]
    kn, mv, gs = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    agg = defaultdict(lambda: [0, 0.0])
    lines = []
    for r in rows[1:]:
        ns = float(r[mv].replace(",", ""))
        agg[r[kn]][0] += 1
        agg[r[kn]][1] += ns
        lines.append(f"{r[0]:>4} {r[kn][:70]:70s} grid={r[gs]:>12} {ns/1e3:12.1f} us")
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(out, f"{tag}_launches.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised): share per kernel\n")
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{100*ns/tot:6.2f}%  launches={n:4d}  total={ns/1e6:10.3f} ms  {k}\n")
        f.write("\n# per launch\n" + "\n".join(lines) + "\n")
    print(open(os.path.join(out, f"{tag}_launches.txt")).read()[:1500])

rep = os.path.join(ROOT, "gpurun_out", f"{tag}_prof_tent.ncu-rep")
if os.path.exists(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(os.path.join(out, f"{tag}_tent_kernel_full.txt"), "w") as f:
        f.write("# ncu --set full --clock-control none, selected metrics per captured launch\n")
        f.write("kernel: " + data[0][hdr.index("Kernel Name")] + "\n")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                f.write(f"{k:90s} [{units[i]}] " + "  ".join(r[i] for r in data) + "\n")
    i_r, i_w = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tr = [float(r[i_r].replace(",", "")) * scale[units[i_r]] + float(r[i_w].replace(",", "")) * scale[units[i_w]] for r in data]
    grids = [int(r[hdr.index("launch__grid_size")].replace(",", "")) for r in data]
    with open(os.path.join(out, "traffic.json"), "w") as f:
        json.dump({"tag": tag, "dram_bytes_per_launch": sum(tr) / len(tr), "dram_bytes_per_tent": sum(tr) / sum(grids),
                   "captured_launches": len(tr), "grid_sizes": grids}, f)
        f.write("\n")
    print(open(os.path.join(out, f"{tag}_tent_kernel_full.txt")).read())

    # ---- executed FP64 work per tent-element (roofline numerator of bench.py): the captured launches are levels of the
    # benched slab (28^3 cells x 6 tets, order 4, dt = 0.05); a level is identified by its tent count = grid size
    def col(name):
        i = hdr.index(name)
        return [float(r[i].replace(",", "")) for r in data]

    try:
        sys.path.insert(0, ROOT)
        import numpy as np
        import ngstrefftz_b200 as ng
        cells = int(os.environ.get("NGST_PROFILE_CELLS", "28"))
        mesh = ng.StructuredMesh([cells] * 3, [0.0] * 3, [1.0] * 3)
        ts = ng.TentSlab(mesh, "edge")
        ts.SetMaxWavespeed(1.0)
        assert ts.PitchTents(0.05, True, 2 / 3)
        e = ts.export()
        nels = np.diff(e["el_off"])
        per_level = {}
        for l in range(int(e["level"].max()) + 1):
            m = e["level"] == l
            per_level.setdefault(int(m.sum()), []).append(int(nels[m].sum()))
        te = []
        for g in grids:
            assert g in per_level and len(per_level[g]) == 1, f"grid size {g} does not identify one level"
            te.append(per_level[g][0])
        dmma = col("sm__inst_executed_pipe_tensor_subpipe_dmma.sum")
        dfma = col("sm__sass_thread_inst_executed_op_dfma_pred_on.sum")
        dmul = col("sm__sass_thread_inst_executed_op_dmul_pred_on.sum")
        dadd = col("sm__sass_thread_inst_executed_op_dadd_pred_on.sum")
        kname = data[0][hdr.index("Kernel Name")]
        rec = {"tag": tag, "kernel": "tent_kernel_fast" if "tent_kernel_fast" in kname else kname, "kernel_instance": kname,
               "D": 3, "order": 4, "workload": f"bench.py --cells {cells} --level-launches, mid-slab levels",
               "formula": "DMMA warp instructions x 512 + 2 x DFMA + DMUL + DADD thread instructions (pred on), per tent-element",
               "captured_grid_sizes": grids, "captured_tent_elements": te,
               "dmma_flops_per_tent_element": 512.0 * sum(dmma) / sum(te),
               "dfma_flops_per_tent_element": (2.0 * sum(dfma) + sum(dmul) + sum(dadd)) / sum(te),
               "dram_bytes_per_tent_element": sum(tr) / sum(te),
               "pipe_shared_pct": col("sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_elapsed"),
               "dmma_pipe_pct": col("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed"),
               "fp64_pipe_pct": col("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"),
               "duration_ms_under_ncu": col("gpu__time_duration.sum")}
        rec["flops_per_tent_element"] = rec["dmma_flops_per_tent_element"] + rec["dfma_flops_per_tent_element"]
        name = os.environ.get("NGST_EXECUTED_JSON", "executed_flops.json")
        with open(os.path.join(out, name), "w") as f:
            json.dump(rec, f, indent=1)
            f.write("\n")
        print(json.dumps(rec, indent=1))
    except Exception as ex:  # the text summaries above stand on their own
        print("executed_flops.json not written:", repr(ex))
