#!/usr/bin/env python
"""Development aid: where does a streamed step (ngst_twave_set_initial_streamed + propagate [+ output mirror]) spend its time."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import ngstrefftz_b200 as ng
from ngstrefftz_b200._capi import C, c_dp, check, lib
from _cases import dirichlet, standing

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 28
f = standing(3)
mesh = ng.StructuredMesh([cells] * 3, [0.0] * 3, [1.0] * 3)
ts = ng.TentSlab(mesh, "edge")
ts.SetMaxWavespeed(1.0)
assert ts.PitchTents(0.05, True, 2 / 3)
tw = ng.TWave(4, ts, 1.0)
tw.SetBoundaryCF(dirichlet(f, 3))
wf0 = tw.MakeWavefront(f, 0.0)
pin = torch.empty(wf0.size, dtype=torch.float64, pin_memory=True)
pin.numpy()[:] = wf0.reshape(-1)
out = torch.zeros(wf0.size, dtype=torch.float64, pin_memory=True)
ptr = C.cast(pin.data_ptr(), c_dp)
optr = C.cast(out.data_ptr(), c_dp)
bv = tw.sample_boundary(0.0)
tw.SetInitial(wf0)
tw.propagate_with_values(bv)   # builds the factor cache
tw.reset_time()
tw.SetInitial(wf0)
tw.propagate_with_values(bv)
print(f"plain: propagate device {tw.last_ms():.2f} ms")


def step(streamed, mirror, n=4):
    check(lib.ngst_twave_set_output_mirror(tw._h, optr if mirror else None))
    res = []
    for _ in range(n):
        tw.reset_time()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if streamed:
            check(lib.ngst_twave_set_initial_streamed(tw._h, ptr, 5))
        else:
            check(lib.ngst_twave_set_initial(tw._h, ptr, 5))
        t1 = time.perf_counter()
        check(lib.ngst_twave_propagate(tw._h))
        t2 = time.perf_counter()
        res.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, tw.last_ms()))
    check(lib.ngst_twave_set_output_mirror(tw._h, None))
    r = np.array(res[1:]).mean(axis=0)
    print(f"streamed={streamed} mirror={mirror}: set_initial {r[0]:.2f} ms, propagate wall {r[1]:.2f} ms (device {r[2]:.2f} ms), "
          f"sum {r[0] + r[1]:.2f} ms")


for streamed in (False, True):
    for mirror in (False, True):
        step(streamed, mirror)
tw.set_kernel_mode(3)
print("single launch forced:")
step(False, False)
