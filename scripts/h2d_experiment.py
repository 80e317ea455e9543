#!/usr/bin/env python
"""How fast do the rows of a pinned host buffer reach the device: blocking ngst_twave_set_initial (copy engine) against
ngst_twave_set_initial_streamed (SM loads from mapped memory) with different numbers of fetch CTAs. Development aid."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import ngstrefftz_b200 as ng
from ngstrefftz_b200._capi import C, c_dp, check, lib
from _cases import standing

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 28
mesh = ng.StructuredMesh([cells] * 3, [0.0] * 3, [1.0] * 3)
ts = ng.TentSlab(mesh, "edge")
ts.SetMaxWavespeed(1.0)
assert ts.PitchTents(0.05, True, 2 / 3)
tw = ng.TWave(4, ts, 1.0)
wf0 = tw.MakeWavefront(standing(3), 0.0)
pin = torch.empty(wf0.size, dtype=torch.float64, pin_memory=True)
pin.numpy()[:] = wf0.reshape(-1)
ptr = C.cast(pin.data_ptr(), c_dp)
tw.SetInitial(wf0)
tw.Propagate()
mb = wf0.nbytes / 1e6


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def plain():
    check(lib.ngst_twave_set_initial(tw._h, ptr, 5))
    tw.Energy()


print(f"{mb:.0f} MB; set_initial + Energy: {timed(plain):.2f} ms")
for ctas in (16, 32, 64, 128, 256, 444):
    check(lib.ngst_twave_set_fetch_ctas(tw._h, ctas))

    def streamed():
        check(lib.ngst_twave_set_initial_streamed(tw._h, ptr, 5))
        tw.Energy()

    print(f"set_initial_streamed ({ctas} CTAs) + Energy: {timed(streamed):.2f} ms")
