#!/usr/bin/env python
"""Small cases that drive every kernel variant once, for compute-sanitizer (scripts/gpu_sanitize.sh): assemble / invert /
apply modes of the fast kernel, dependency-driven and per-level launches, spatial sweep with output mirror, generic kernel
with its factor cache, both quasi-Trefftz kernels, pipelined steps."""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import ngstrefftz_b200 as ng
from _cases import dirichlet, gaussian_1d, standing
from test_oracle_qt import qt_exact


def slab(mesh, dt, c=1.0):
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    return ts


f = standing(3)
ts = slab(ng.UnitCubeMesh(4), 0.15)
for mode in (0, 2, 3):
    tw = ng.TWave(4, ts, 1.0)
    tw.set_kernel_mode(mode)
    tw.SetInitial(f)
    tw.SetBoundaryCF(dirichlet(f, 3))
    out = np.zeros((tw.nel, tw.ncomp * tw.nip))
    tw.SetOutputMirror(out)
    for _ in range(3):
        tw.Propagate()
    assert np.array_equal(out, tw.GetWavefront())
    tw.SetOutputMirror(None)
    wf0 = tw.MakeWavefront(f, 0.0)
    tw.reset_time()
    tw.SetInitialStreamed(wf0)
    tw.Propagate()
    print("fast kernel mode", mode, "energy", tw.Energy())
tw = ng.TWave(4, ts, 1.0)
tw.set_compact_state(False)
tw.SetInitial(f)
tw.SetBoundaryCF(dirichlet(f, 3))
tw.Propagate()
print("public rows", tw.Energy())
c1 = lambda x: np.where(x[:, 0] < 0.5, 1.0, 2.0)  # noqa: E731
t1 = slab(ng.SegMesh(32, 0, 1), 0.2, c1)
tw = ng.TWave(6, t1, c1)
tw.SetInitial(gaussian_1d)
tw.SetBoundaryCF(lambda x, t: 0 * t)
for _ in range(3):
    tw.Propagate()
print("generic kernel + cache", tw.factor_cache(), tw.Energy())
cq = ng.PolynomialCF.affine([1.0, 1.0], 1.0)
fq = qt_exact(2)
for order in (4, 2):
    tq = slab(ng.UnitSquareMesh(4), 0.1, cq)
    tw = ng.TWave(order, tq, cq, BBcf=1)
    tw.SetInitial(fq)
    tw.SetBoundaryCF(lambda x, t: fq(x, t)[:, 3])
    tw.Propagate()
    print("quasi-Trefftz", tw.kernel_name(), tw.Energy())
print("SANITIZE_CASES_OK")
