"""GPU parity (K6): libngst_b200.so through its C ABI vs the CPU oracle on the same tents and data.

Bar: tent tables bitwise identical; wavefront and Energy within 1e-10 relative in FP64 (BASELINE.json north_star).
"""
import math

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, gaussian_1d, neumann, poly_solution, standing
from _oracle import Oracle

pytestmark = pytest.mark.gpu
TOL = 1e-10


def both(mesh, order, c, dt, init, bnd, nslabs=1, neu=False):
    if neu:
        mesh.SetBoundaryNames("neumann")
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, c)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, c(o.centroids) if callable(c) else c)
    e1, e2 = ts.export(), o.export()
    for k in e2:
        assert np.array_equal(e1[k], e2[k]), f"tent table {k} differs"
    o.setup(order, c)
    o.set_initial(init)
    o.set_boundary_cf(bnd)
    for _ in range(nslabs):
        tw.Propagate()
        o.propagate()
    assert tw.last_launches() > 0
    return tw, o


def forced_tol(o, what=""):
    """1e-10 (north_star), or cond * eps of the (diagonally scaled) tent systems where that is larger: two backward-stable
    solvers -- the reference's LAPACK inverse-then-multiply, the elimination on the device -- may legitimately differ by
    that much. The condition number is measured by the oracle on the systems it actually inverted and printed."""
    eps = np.finfo(float).eps
    cond = o.max_cond()
    tol = max(TOL, cond * eps)
    print(f"{what}: max scaled cond of the tent systems {cond:.3e}, cond*eps {cond * eps:.2e} -> tolerance {tol:.1e}")
    return tol


def check(tw, o, tol=TOL):
    wf, ref = tw.GetWavefront(), o.wavefront()
    rel = np.abs(wf - ref).max() / np.abs(ref).max()
    assert rel <= tol, rel
    if not tw.fosystem:
        eg, eo = tw.Energy(wf), o.energy(ref)
        assert abs(eg - eo) <= tol * abs(eo), (eg, eo)
        assert abs(tw.Energy() - eg) <= 1e-14 * abs(eg)  # resident state == downloaded state
    assert abs(tw.timeshift() - o_timeshift(o)) == 0
    return rel


def o_timeshift(o):
    from _oracle import lib
    return lib.oracle_timeshift(o.h)


CASES = {
    # name: (mesh factory, order, c, dt, init, bnd, nslabs, neumann)
    "cfg0_2d_p4_h025": (lambda: ng.UnitSquareMesh(4), 4, 1.0, 0.5, standing(2), dirichlet(standing(2), 2), 1, False),
    "1d_p4_dirichlet": (lambda: ng.SegMesh(16, 0, 1), 4, 1.0, 0.5, standing(1), dirichlet(standing(1), 1), 1, False),
    "1d_p4_neumann": (lambda: ng.SegMesh(16, 0, 1), 4, 1.0, 0.5, standing(1), neumann(standing(1), 1), 1, True),
    "1d_p6_first_order": (lambda: ng.SegMesh(16, 0, 1), 6, 1.0, 0.5, first_order(standing(1)), dirichlet(standing(1), 1), 1, False),
    "1d_segpi": (lambda: ng.SegMesh(4, 0, math.pi), 4, 1.0, 0.5, standing(1), dirichlet(standing(1), 1), 1, False),
    "2d_p4_neumann": (lambda: ng.UnitSquareMesh(4), 4, 1.0, 0.5, standing(2), neumann(standing(2), 2), 1, True),
    "2d_p4_fo_2slabs": (lambda: ng.UnitSquareMesh(6), 4, 1.0, 0.3, first_order(standing(2)), dirichlet(standing(2), 2), 2, False),
    "2d_p5_poly": (lambda: ng.UnitSquareMesh(3), 5, 1.0, 0.3, poly_solution(2), dirichlet(poly_solution(2), 2), 1, False),
    "2d_p2_c3": (lambda: ng.UnitSquareMesh(5), 2, 3.0, 0.1, standing(2, 3.0), dirichlet(standing(2, 3.0), 2), 3, False),
    "3d_p3": (lambda: ng.UnitCubeMesh(3), 3, 1.0, 0.2, standing(3), dirichlet(standing(3), 3), 1, False),
    "3d_p4": (lambda: ng.UnitCubeMesh(4), 4, 1.0, 0.2, standing(3), dirichlet(standing(3), 3), 1, False),
    "3d_p4_neumann": (lambda: ng.UnitCubeMesh(3), 4, 1.0, 0.2, standing(3), neumann(standing(3), 3), 1, True),
    "3d_p4_fo": (lambda: ng.UnitCubeMesh(3), 4, 1.0, 0.2, first_order(standing(3)), dirichlet(standing(3), 3), 2, False),
    "3d_p1": (lambda: ng.UnitCubeMesh(2), 1, 1.0, 0.2, standing(3), dirichlet(standing(3), 3), 1, False),
    "3d_box_p3": (lambda: ng.StructuredMesh([4, 2, 3], [0, 0, 0], [1, 0.6, 1.3]), 3, 1.3, 0.15, standing(3, 1.3),
                  dirichlet(standing(3, 1.3), 3), 2, False),
}


@pytest.mark.parametrize("name", list(CASES))
def test_parity_constant_wavespeed(built, name):
    mk, order, c, dt, init, bnd, nslabs, neu = CASES[name]
    tw, o = both(mk(), order, c, dt, init, bnd, nslabs, neu)
    check(tw, o)


def pc1(x):
    return np.where((np.floor(x[:, 0] * 4).astype(int) % 2) == 0, 1.0, 2.0)


MACRO = {
    "1d_p6_piecewise": (lambda: ng.SegMesh(32, 0, 1), 6, pc1, 0.3, gaussian_1d, lambda x, t: 0 * t, 2),
    "2d_p3_piecewise": (lambda: ng.UnitSquareMesh(4), 3, lambda x: np.where(x[:, 0] < 0.5, 1.0, 1.5), 0.2, standing(2),
                        dirichlet(standing(2), 2), 1),
    "3d_p2_piecewise": (lambda: ng.UnitCubeMesh(2), 2, lambda x: np.where(x[:, 2] < 0.5, 1.0, 2.0), 0.2, standing(3),
                        dirichlet(standing(3), 3), 1),
}


@pytest.mark.parametrize("name", list(MACRO))
def test_parity_material_interfaces(built, name):
    """CalcTentMacroEl (twavetents.cpp:436-539). 1e-10, except where cond * eps of the tent systems (printed) is larger: the
    order-6 1-D systems have a scaled condition number of 2e7, so the oracle's inverse-then-multiply
    (twavetents.cpp:64-70) and a pivoted elimination legitimately differ by up to 5e-9."""
    mk, order, c, dt, init, bnd, nslabs = MACRO[name]
    tw, o = both(mk(), order, c, dt, init, bnd, nslabs)
    tol = forced_tol(o, name)
    assert tol <= (1e-8 if name.startswith("1d") else TOL)
    check(tw, o, tol=tol)


def test_known_answer_on_gpu(built):
    """K1 on the device: a cubic wave polynomial is reproduced to round-off, Error() agrees with the oracle's."""
    f = poly_solution(3, 1.2)
    tw, o = both(ng.UnitCubeMesh(2), 4, 1.2, 0.25, f, dirichlet(f, 3), nslabs=2)
    wf = tw.GetWavefront()
    exact = tw.MakeWavefront(f, 0.5)
    assert np.abs(wf - exact).max() <= 1e-10 * np.abs(exact).max()
    assert tw.Error(wf, exact) <= 1e-10 * math.sqrt(2 * tw.Energy(exact))
    ref = o.wavefront()
    assert abs(tw.L2Error(wf, 0 * wf) - o.l2error(ref, 0 * ref)) <= 1e-10 * o.l2error(ref, 0 * ref)
    assert abs(tw.MaxAdiam() - o.max_adiam()) <= 1e-15 * o.max_adiam()


def test_size_independent_properties_large(built):
    """At a size the oracle does not finish in seconds: energy decay with homogeneous data and determinism
    (two solvers on the same data give bitwise identical states)."""
    f = standing(3)
    mesh = ng.UnitCubeMesh(12)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.05, True, 2 / 3)
    out = []
    for _ in range(2):
        tw = ng.TWave(4, ts, 1.0)
        tw.SetInitial(f)
        tw.SetBoundaryCF(lambda x, t: 0 * t)
        en = [tw.Energy()]
        for _ in range(2):
            tw.Propagate()
            en.append(tw.Energy())
        assert en[2] <= en[1] * (1 + 1e-12) <= en[0] * (1 + 1e-12) ** 2 and en[2] > 0.9 * en[0]
        out.append(tw.GetWavefront())
    assert np.array_equal(out[0], out[1])
    exact = tw.MakeWavefront(f, 0.1)
    assert tw.Error(out[0], exact) < 2e-4


def test_dependency_driven_launch_matches_level_launches(built):
    """One GPU, every tent on the fast kernel: Propagate is ONE launch per slab whose CTAs draw tents in level order and wait
    on per-tent hazard counters (replacing RunParallelDependency, paralleldepend.hpp:20-131). Mode 2 keeps one launch per
    level. Same tents, same arithmetic per tent: the states must be bitwise identical."""
    f = standing(3)
    mesh = ng.UnitCubeMesh(10)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.08, True, 2 / 3)
    out, launches = [], []
    for mode in (3, 2):
        tw = ng.TWave(4, ts, 1.0)
        tw.set_kernel_mode(mode)
        tw.SetInitial(f)
        tw.SetBoundaryCF(dirichlet(f, 3))
        for _ in range(3):
            tw.Propagate()
        out.append(tw.GetWavefront())
        launches.append(tw.last_launches())
    assert launches[0] == 1 and launches[1] == ts.GetNLayers()
    assert np.array_equal(out[0], out[1])


@pytest.mark.parametrize("D,fo", [(1, False), (2, False), (3, False), (2, True), (3, True)])
def test_get_wave_projection(built, D, fo):
    """GetWave (src/__init__.py:7-43): the element-wise L2 projection reproduces a polynomial state exactly (degree p for
    u, p-1 for the first-order components), its basis is orthonormal in the rule, and it equals the numpy least-squares
    projection for a non-polynomial state."""
    order = 3
    mesh = {1: lambda: ng.SegMesh(6, 0, 1), 2: lambda: ng.UnitSquareMesh(3), 3: lambda: ng.UnitCubeMesh(2)}[D]()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.1, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    pf = poly_solution(D, 1.0)  # cubic u, quadratic derivatives
    init = first_order(pf) if fo else pf
    tw.SetInitial(init)
    w = tw.GetWave()
    assert w.order == (order - 1 if fo else order)
    assert w.coef.shape[:2] == (mesh.ne, D + 1 if fo else 1)
    wf = tw.GetWavefront().reshape(mesh.ne, -1, tw.nip)
    got = w.at_quadrature()
    want = wf[:, :got.shape[1], :]
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
    rule_w = tw.quadrature()[1][0] / tw.quadrature()[1][0].sum() / math.factorial(D)
    gram = w.psi_at_points.T @ (rule_w[:, None] * w.psi_at_points)
    assert np.abs(gram - np.eye(gram.shape[0])).max() < 1e-12
    # monomial form of the basis agrees with its point values (evaluation off the rule points goes through it)
    xq = np.zeros((tw.nip, D))
    x, _ = tw.quadrature()
    P = mesh.points[mesh.elements[0]]
    lam = np.linalg.lstsq(np.vstack([P.T, np.ones(D + 1)]), np.vstack([x[0].T, np.ones(tw.nip)]), rcond=None)[0]
    xq = lam[:D].T
    assert np.abs(w(0, xq) - got[0]).max() <= 1e-10 * np.abs(got[0]).max()
    # non-polynomial state: compare with a weighted least-squares projection done in numpy
    f = standing(D)
    tw.SetInitial(first_order(f) if fo else f)
    w2 = tw.GetWave()
    wf2 = tw.GetWavefront().reshape(mesh.ne, -1, tw.nip)[:, :w2.coef.shape[1], :]
    sw = np.sqrt(rule_w)
    ref = np.linalg.lstsq(sw[:, None] * w2.psi_at_points, (sw * wf2).reshape(-1, tw.nip).T, rcond=None)[0]
    assert np.abs(w2.coef.reshape(-1, ref.shape[0]).T - ref).max() <= 1e-12 * np.abs(ref).max()


def test_import_reference_wavefront_drops_simd_padding(built):
    """The reference's wavefront has snip = nip rounded up to the SIMD width columns per component (twavetents.cpp:192, 741)."""
    mesh = ng.UnitCubeMesh(2)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.1, True, 2 / 3)
    tw = ng.TWave(4, ts, 1.0)
    tw.SetInitial(standing(3))
    wf = tw.GetWavefront()
    assert tw.nip == 125
    padded = np.zeros((mesh.ne, 5, 128))
    padded[:, :, :125] = wf.reshape(mesh.ne, 5, 125)
    padded[:, :, 125:] = padded[:, :, 124:125]  # pad = last point repeated
    assert np.array_equal(tw.import_reference_wavefront(padded.reshape(mesh.ne, -1), simd_width=4), wf)
    with pytest.raises(ValueError):
        tw.import_reference_wavefront(np.zeros((mesh.ne, 21)))


def test_errors_on_gpu(built):
    from ngstrefftz_b200._capi import NgstError
    mesh = ng.UnitSquareMesh(2)
    ts = ng.TentSlab(mesh)
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.1, True, 2 / 3)
    tw = ng.TWave(2, ts, 1.0)
    with pytest.raises(NgstError) as ei:
        tw.Propagate()  # no initial data
    assert ei.value.code == -2
    with pytest.raises(NgstError):
        tw.SetInitial(np.zeros((mesh.ne, 7 * tw.nip)))  # neither D+1 nor D+2 components
    with pytest.raises(NgstError) as ei:
        ng.TWave(2, ts, 1.0, device=99)
    assert ei.value.code == -3


# ---------------------------------------------------------------------------------------------- quasi-Trefftz (QTWaveTents)
def qt_both(mesh, order, dt, nslabs=1, cpoly=None, bpoly=None, generic=False):
    from test_oracle_qt import qt_exact
    D = mesh.dim
    c = cpoly or ng.PolynomialCF.affine([1.0] * D, 1.0)
    f = qt_exact(D)
    bnd = lambda x, t: f(x, t)[:, D + 1]  # noqa: E731
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, c, BBcf=bpoly if bpoly is not None else 1)
    if generic:
        tw.set_kernel_mode(1)
    tw.SetInitial(f)
    tw.SetBoundaryCF(bnd)
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, c(o.centroids))
    e1, e2 = ts.export(), o.export()
    for k in e2:
        assert np.array_equal(e1[k], e2[k]), f"tent table {k} differs"
    o.setup_qt(order, c, bpoly)
    o.set_initial(f)
    o.set_boundary_cf(bnd)
    for _ in range(nslabs):
        tw.Propagate()
        o.propagate()
    # D = 2, orders 3-5 run on the monomial-space DMMA kernel, everything else on the generic quasi-Trefftz kernel
    want = "tent_kernel_qt_fast" if (D == 2 and 3 <= order <= 5 and not generic) else "tent_kernel_qt"
    assert tw.kernel_name() == want and tw.last_launches() > 0
    return tw, o, f


QT_CASES = {
    "1d_p4_h8": (lambda: ng.SegMesh(8, 0, math.pi), 4, 1.0, 1, None),
    "1d_p3_2slabs": (lambda: ng.SegMesh(16, 0, 1), 3, 0.2, 2, None),
    "2d_p4": (lambda: ng.UnitSquareMesh(4), 4, 0.25, 1, None),
    "2d_p5": (lambda: ng.UnitSquareMesh(3), 5, 0.2, 1, None),
    "2d_p3_varB": (lambda: ng.UnitSquareMesh(4), 3, 0.2, 2, "B"),
    "2d_p5_varB_3slabs": (lambda: ng.UnitSquareMesh(5), 5, 0.15, 3, "B"),
    "2d_p2": (lambda: ng.UnitSquareMesh(4), 2, 0.2, 1, None),
    "2d_p4_generic": (lambda: ng.UnitSquareMesh(4), 4, 0.25, 1, "generic"),
    "2d_p5_generic": (lambda: ng.UnitSquareMesh(3), 5, 0.2, 1, "generic"),
}


@pytest.mark.parametrize("name", list(QT_CASES))
def test_parity_quasi_trefftz(built, name):
    """QTWaveTents (twavetents.cpp:967-1144): per-tent basis recursion, per-point wavespeed, cone volume terms."""
    mk, order, dt, nslabs, var = QT_CASES[name]
    mesh = mk()
    bpoly = ng.PolynomialCF(mesh.dim, {(0, 0): 1.0, (1, 0): 0.2, (0, 1): 0.1}) if var == "B" else None
    tw, o, f = qt_both(mesh, order, dt, nslabs, bpoly=bpoly, generic=(var == "generic"))
    wf, ref = tw.GetWavefront(), o.wavefront()
    rel = np.abs(wf - ref).max() / np.abs(ref).max()
    tol = forced_tol(o, name)
    print(f"{name}: max rel diff {rel:.3e}")
    assert tol <= 1e-9 and rel <= tol, (rel, tol)
    exact = tw.MakeWavefront(f, nslabs * dt)
    eg, eo = tw.Error(wf, exact), o.error(ref, o.make_wavefront(f, nslabs * dt))
    # Error is the norm of a DIFFERENCE of nearly equal states (the scheme reproduces this solution to ~1e-6 .. 1e-9 of
    # its size): a state difference of tol * |state| moves it by up to tol * |state| in absolute terms
    scale = math.sqrt(2 * o.energy(ref))
    print(f"{name}: Error {eg:.6e} vs oracle {eo:.6e}, |state| {scale:.3e}")
    assert abs(eg - eo) <= 10 * tol * scale + 1e-10 * eo, (eg, eo)
    assert abs(tw.Energy(wf) - o.energy(ref)) <= 2 * tol * o.energy(ref)


def test_reference_doctest_digit_qt_on_gpu(built):
    """test/tents.py:133-141, second entry (h = 8): 0.001..."""
    tw, o, f = qt_both(ng.SegMesh(8, 0, math.pi), 4, 1.0)
    err = tw.Error(tw.GetWavefront(), tw.MakeWavefront(f, 1.0))
    assert repr(err).startswith("0.001"), err


def test_pointwise_wavespeed_in_error_plain_trefftz(built):
    """Plain Trefftz with a varying wavespeed (TestQTrefftz(..., None), test/tents.py:150-157): Error evaluates c per point."""
    from test_oracle_qt import qt_exact
    c = ng.PolynomialCF.affine([1.0], 1.0)
    f = qt_exact(1)
    tw, o = both(ng.SegMesh(16, 0, math.pi), 4, c, 1.0, f, lambda x, t: f(x, t)[:, 2])
    tw.set_point_wavespeed(c)
    o.set_point_wavespeed(c)
    wf = tw.GetWavefront()
    err = tw.Error(wf, tw.MakeWavefront(f, 1.0))
    assert repr(err).startswith("0.01"), err
    # Error is the norm of a difference of nearly equal states: it moves by up to tol * |state| in absolute terms
    ref = o.wavefront()
    scale = math.sqrt(2 * o.energy(ref))
    assert abs(err - o.error(ref, o.make_wavefront(f, 1.0))) <= 10 * forced_tol(o, "plain Trefftz, c = x + 1") * scale


@pytest.mark.parametrize("D,mk,order,dt", [(2, lambda: ng.UnitSquareMesh(5), 4, 0.3), (3, lambda: ng.UnitCubeMesh(4), 4, 0.2)])
def test_parity_on_level_sync_tents(built, D, mk, order, dt):
    """The solver on the tents of the level-synchronous pitcher (ngst_slab_pitch_level_sync, pick rule "minimal (level, vertex
    number)") against the oracle on ITS serial loop with the same rule: tables bitwise, wavefront to 1e-10."""
    f = standing(D)
    mesh = mk()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3, level_sync=True)
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, 1.0, rule=1)
    e1, e2 = ts.export(), o.export()
    for k in e2:
        assert np.array_equal(e1[k], e2[k]), f"tent table {k} differs"
    tw = ng.TWave(order, ts, 1.0)
    tw.SetInitial(f)
    tw.SetBoundaryCF(dirichlet(f, D))
    o.setup(order, 1.0)
    o.set_initial(f)
    o.set_boundary_cf(dirichlet(f, D))
    for _ in range(2):
        tw.Propagate()
        o.propagate()
    check(tw, o)
