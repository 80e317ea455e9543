"""Host-side model of the level-synchronous launch scheme at bench.py's workload: list scheduling of the tents of each level
(most elements first, as Solver orders them) on 444 CTA slots with cost ~ elements + 1.5, against perfect packing.
Output quoted in DESIGN.md section 9."""
import sys, heapq
sys.path.insert(0, "/root/repo")
import numpy as np
import ngstrefftz_b200 as ng
n = 28
mesh = ng.StructuredMesh([n]*3, [0.0]*3, [1.0]*3)
ts = ng.TentSlab(mesh, "edge"); ts.SetMaxWavespeed(1.0)
assert ts.PitchTents(0.05, True, 2/3)
e = ts.export()
nels = np.diff(e["el_off"]); lev = e["level"]
nl = lev.max()+1
slots = 444
tot_ideal = 0.0; tot_real = 0.0
for l in range(nl):
    w = np.sort(nels[lev == l])[::-1].astype(float) + 1.5   # per-tent cost ~ elements + fixed part (solve etc.)
    h = [0.0]*slots
    for x in w:
        t = heapq.heappop(h); heapq.heappush(h, t + x)
    real = max(h); ideal = w.sum()/slots
    tot_ideal += ideal; tot_real += real
    if l < 3 or l > nl-8: print(l, len(w), round(ideal,1), round(real,1))
print("levels", nl, "ideal", round(tot_ideal), "level-synchronous", round(tot_real), "loss %", round(100*(tot_real/tot_ideal-1),1))
