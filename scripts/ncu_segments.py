#!/usr/bin/env python
"""Per-barrier-segment breakdown of warp-state samples of the first kernel in an ncu report (development aid).
usage: ncu_segments.py <report.ncu-rep> [min_share_percent]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
minshare = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
print(rows[starts[0]][1])
hdr = rows[starts[0] + 1]
body = rows[starts[0] + 2:starts[1]] if len(starts) > 1 else rows[starts[0] + 2:]
S, SRC, IE = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
names = [hdr[c] for c in stall_cols]


def new(start):
    return {"n": 0, "samples": 0, "inst": 0, "dmma": 0, "st": [0] * len(stall_cols), "start": start, "ops": {}}


seg, cur = [], new(0)
for idx, r in enumerate(body):
    src, s, ie = r[SRC], int(r[S]), int(r[IE])
    cur["n"] += 1
    cur["samples"] += s
    cur["inst"] += ie
    toks = src.split()
    op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
    cur["ops"][op] = cur["ops"].get(op, 0) + s
    if "DMMA" in src:
        cur["dmma"] += ie
    for k, c in enumerate(stall_cols):
        cur["st"][k] += int(r[c])
    if "BAR.SYNC" in src:
        seg.append(cur)
        cur = new(idx + 1)
seg.append(cur)
tot = sum(x["samples"] for x in seg)
print("total samples", tot)
for x in seg:
    if x["samples"] < tot * minshare / 100:
        continue
    top = sorted(zip(x["st"], names), reverse=True)[:5]
    ops = sorted(x["ops"].items(), key=lambda kv: -kv[1])[:6]
    print(f"seg@{x['start']:6d} n={x['n']:5d} samples={100*x['samples']/tot:5.1f}% inst={x['inst']:9d} dmma={x['dmma']:8d} | "
          + ", ".join(f"{n[6:]}={100*v/max(x['samples'],1):.0f}%" for v, n in top) + " | "
          + ", ".join(f"{o}:{100*v/max(x['samples'],1):.0f}%" for o, v in ops))
