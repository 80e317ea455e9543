"""GPU vs CPU oracle AT THE BENCHED SIZES: the first tent levels of the slab bench.py times (BASELINE configs[4]'s per-GPU
share: 28^3 cells x 6 tets, order 4, dt = 0.05) and of configs[1..3] at their full sizes, compared row by row.

The oracle needs minutes for a whole slab of this size, but the tents of the first levels only depend on the initial
data: ngst_twave_propagate_levels runs them on the device, the oracle runs the same tents (`ntents_limit`, level order),
and every state row they touch must agree to 1e-10 relative (BASELINE.json north_star). Where the oracle's literal
pitcher (linear scans) finishes in seconds the tent tables are also compared bitwise at this size; otherwise the oracle
runs on the product's tents (oracle_import_tents; the pitchers are compared bitwise in tests/test_pitcher.py and in
tests/test_gpu_parity.py at smaller sizes).
"""
import math
import os
import sys

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, gaussian_1d, standing
from _oracle import Oracle

pytestmark = pytest.mark.gpu
TOL = 1e-10


def first_levels(mesh, order, cw, dt, init, bnd, nlev, oracle_tents=None, oracle_pitch=True, qt=False, expect_kernel=None):
    """GPU: first `nlev` levels. Oracle: the same tents, or only the first `oracle_tents` of level 0 (then nlev == 1).
    Returns (max relative row difference, number of compared rows, number of oracle tents)."""
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(cw)
    assert ts.PitchTents(dt, True, 2 / 3)
    e = ts.export()
    o = Oracle(mesh)
    if oracle_pitch:
        cmax = cw(o.centroids) if callable(cw) else cw
        assert o.pitch(dt, True, 2 / 3, cmax)
        e2 = o.export()
        for k in e2:
            assert np.array_equal(e[k], e2[k]), f"tent table {k} differs from the oracle's at full size"
    else:
        o.import_tents(e, dt)
    level_off = np.concatenate([[0], np.cumsum(np.bincount(e["level"]))])
    assert nlev < len(level_off) - 1
    ntents = int(level_off[nlev])
    if oracle_tents is not None:
        assert nlev == 1 and oracle_tents <= ntents
        ntents = oracle_tents
    tw = ng.TWave(order, ts, cw, BBcf=1) if qt else ng.TWave(order, ts, cw)
    if expect_kernel:
        assert tw.kernel_name() == expect_kernel
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    if qt:
        o.setup_qt(order, cw)
    else:
        o.setup(order, cw)
    o.set_initial(init)
    o.set_boundary_cf(bnd)
    ref0 = o.wavefront()
    wf_init = tw.GetWavefront()
    tw.propagate_levels(nlev)
    o.propagate(ntents_limit=ntents)
    assert tw.timeshift() == 0.0  # a partial slab does not advance time
    wf, ref = tw.GetWavefront(), o.wavefront()
    # tents in level order, ascending id inside a level: what the oracle ran
    ran = np.argsort(e["level"], kind="stable")[:ntents]
    touched = np.unique(np.concatenate([e["els"][e["el_off"][t]:e["el_off"][t + 1]] for t in ran]))
    assert len(touched) > 0 and not np.array_equal(ref[touched], ref0[touched])
    rel = np.abs(wf[touched] - ref[touched]).max() / np.abs(ref).max()
    if oracle_tents is None:
        # rows no tent of these levels touches still hold the initial data, bit for bit, on both sides
        rest = np.setdiff1d(np.arange(mesh.ne), touched)
        assert np.array_equal(wf[rest], wf_init[rest]) and np.array_equal(ref[rest], ref0[rest])
    # 1e-10 (north_star) unless the systems are so ill-conditioned that two backward-stable solvers (the reference's LAPACK
    # inverse, the elimination on the device) legitimately differ by more: cond * eps of the diagonally scaled tent systems
    cond = o.max_cond()
    tol = max(TOL, cond * np.finfo(float).eps)
    print(f"  max scaled cond of the tent systems {cond:.3e} -> cond*eps {cond * np.finfo(float).eps:.2e}, tolerance {tol:.1e}")
    return rel, len(touched), ntents, tol


def test_bench_slab_first_levels_match_oracle(built):
    """The slab bench.py times (131 712 tets, 114 072 tents, 42 levels): tent tables bitwise equal to the oracle's literal
    pitcher, levels 0-1 (7 412 tents, 0.13 M tent-elements) row by row."""
    f = standing(3, 1.0)
    mesh = ng.StructuredMesh([28] * 3, [0.0] * 3, [1.0] * 3)
    rel, nrows, nt, tol = first_levels(mesh, 4, 1.0, 0.05, f, dirichlet(f, 3), 2, expect_kernel="tent_kernel_fast")
    print(f"bench slab: {nt} tents, {nrows} rows compared, max rel diff {rel:.3e}")
    assert nt == 7412 and rel <= TOL, rel


def test_config2_first_levels_match_oracle(built):
    """configs[2] at full size: 26^3 x 6 = 105 456 tets, order 3."""
    f = standing(3, 1.0)
    rel, nrows, nt, tol = first_levels(ng.UnitCubeMesh(26), 3, 1.0, 0.05, f, dirichlet(f, 3), 2, expect_kernel="tent_kernel_fast")
    print(f"config 2: {nt} tents, {nrows} rows compared, max rel diff {rel:.3e}")
    assert rel <= TOL, rel


def test_config1_first_levels_match_oracle(built):
    """configs[1] at full size: 1+1D, order 6, 100 000 elements, piecewise-constant wavespeed (material-interface tents on
    the macro-element path, src/twavetents.cpp:436-539), > 1e6 tents per slab; the first two levels on the product's tents."""
    n = 100000
    cw_el = np.where((np.arange(n) // 1000) % 2 == 0, 1.0, 2.0)
    mesh = ng.SegMesh(n, 0.0, 1.0)
    rel, nrows, nt, tol = first_levels(mesh, 6, cw_el, 8e-5, gaussian_1d, lambda x, t: 0 * t, 2, oracle_pitch=False)
    print(f"config 1: {nt} tents, {nrows} rows compared, max rel diff {rel:.3e}")
    # order-6 1-D tent systems have a scaled condition number of 2e7: cond * eps = 5e-9 is the forced bound here
    assert nt > 50000 and rel <= tol <= 1e-8, (rel, tol)


def test_config3_first_tents_match_oracle(built):
    """configs[3] at full size: 2+1D quasi-Trefftz, c = x + y + 1, order 5, 708^2 x 2 = 1 002 528 triangles. The device runs
    level 0; the oracle the first 2 000 tents of it (tents of one level are independent)."""
    from test_oracle_qt import qt_exact
    c = ng.PolynomialCF.affine([1.0, 1.0], 1.0)
    f = qt_exact(2)
    rel, nrows, nt, tol = first_levels(ng.UnitSquareMesh(708), 5, c, 0.0005, f, lambda x, t: f(x, t)[:, 3], 1,
                                       oracle_tents=2000, oracle_pitch=False, qt=True, expect_kernel="tent_kernel_qt_fast")
    print(f"config 3: {nt} tents, {nrows} rows compared, max rel diff {rel:.3e}")
    assert rel <= tol <= 1e-9, (rel, tol)
