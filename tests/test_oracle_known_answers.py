"""Pins of the CPU oracle (oracle/): the reference's own doctest digits for the reproducible 1-D cases
(test/tents.py:34-75) and mesh-/rule-independent known answers K1, K3, K5 of SURVEY.md section 8.3."""
import math

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, neumann, poly_solution, standing
from _oracle import Oracle


def solve(mesh, order, c, dt, f, bnd, nslabs=1, neu=False, init=None):
    if neu:
        mesh.SetBoundaryNames("neumann")
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, c(o.centroids) if callable(c) else c)
    o.setup(order, c)
    o.set_initial(init if init is not None else f)
    o.set_boundary_cf(bnd)
    for _ in range(nslabs):
        o.propagate()
    return o


# ---- reference doctest pins, test/tents.py:46-48 and :67-69 (ELLIPSIS patterns "0.12...", "...e-05", "...e-06")
@pytest.mark.parametrize("neu", [False, True])
def test_reference_doctest_digits_1d(built, neu):
    got = []
    for mesh in (ng.SegMesh(4, 0, math.pi), ng.SegMesh(8, 0, 1), ng.SegMesh(16, 0, 1)):
        f = standing(1, 1.0)
        o = solve(mesh, 4, 1.0, 0.5, f, neumann(f, 1) if neu else dirichlet(f, 1), neu=neu)
        got.append(o.error(o.wavefront(), o.make_wavefront(f, 0.5)))
    assert repr(got[0]).startswith("0.12")
    assert repr(got[1]).endswith("e-05")
    assert repr(got[2]).endswith("e-06")


# ---- K1: polynomial solutions of degree <= p are reproduced to round-off
@pytest.mark.parametrize("D,n", [(1, 6), (2, 3), (3, 2)])
@pytest.mark.parametrize("neu", [False, True])
@pytest.mark.parametrize("c", [1.0, 1.7])
def test_trefftz_exactness(built, D, n, neu, c):
    mesh = {1: ng.SegMesh, 2: lambda k, a, b: ng.UnitSquareMesh(k), 3: lambda k, a, b: ng.UnitCubeMesh(k)}[D](n, 0, 1)
    f = poly_solution(D, c)
    dt = 0.3
    o = solve(mesh, 4, c, dt, f, neumann(f, D) if neu else dirichlet(f, D), nslabs=2, neu=neu)
    exact = o.make_wavefront(f, 2 * dt)
    wf = o.wavefront()
    assert np.abs(wf - exact).max() <= 1e-10 * np.abs(exact).max()
    assert o.error(wf, exact) <= 1e-10 * math.sqrt(2 * o.energy(exact))
    # K3: energy equals the analytic (quadrature-exact) value
    assert abs(o.energy(wf) - o.energy(exact)) <= 1e-10 * o.energy(exact)


def test_first_order_input_matches_second_order(built):
    """SetInitial with D+1 components switches to the first-order system with nbasis-1 (twavetents.hpp:198-203);
    both discretise the same wave, so the gradient parts agree to discretisation accuracy."""
    f = standing(2, 1.0)
    o2 = solve(ng.UnitSquareMesh(4), 4, 1.0, 0.3, f, dirichlet(f, 2))
    o1 = solve(ng.UnitSquareMesh(4), 4, 1.0, 0.3, f, dirichlet(f, 2), init=first_order(f))
    assert o1.nbasis() == o2.nbasis() - 1 == 24
    a, b = o2.wavefront()[:, o2.nip:], o1.wavefront()
    assert np.abs(a - b).max() < 2e-4
    ex = o1.make_wavefront(first_order(f), 0.3)
    assert np.abs(b - ex).max() < 5e-2


# ---- K3: energy is non-increasing slab to slab (dissipative upwind DG), homogeneous Dirichlet data (dt u = 0)
def test_energy_non_increasing(built):
    f = standing(2, 1.0)
    mesh = ng.UnitSquareMesh(4)
    o = Oracle(mesh)
    assert o.pitch(0.2, True, 2 / 3, 1.0)
    o.setup(3, 1.0)
    o.set_initial(f)
    o.set_boundary_cf(lambda x, t: 0 * t)
    en = [o.energy(o.make_wavefront(f, 0.0))]
    for _ in range(4):
        o.propagate()
        en.append(o.energy(o.wavefront()))
    assert all(en[i + 1] <= en[i] * (1 + 1e-12) for i in range(4)), en
    assert en[-1] > 0.5 * en[0]


# ---- K5: convergence on the standing wave
@pytest.mark.parametrize("D", [1, 2])
def test_convergence_rate(built, D):
    f = standing(D, 1.0)
    errs = []
    for n in (4, 8, 16):
        mesh = ng.SegMesh(n, 0, 1) if D == 1 else ng.UnitSquareMesh(n)
        o = solve(mesh, 3, 1.0, 0.25, f, dirichlet(f, D))
        errs.append(o.error(o.wavefront(), o.make_wavefront(f, 0.25)))
    rates = [math.log2(errs[i] / errs[i + 1]) for i in range(2)]
    assert min(rates) > 2.2 and rates[-1] > 2.8, (errs, rates)  # order p in the energy norm


def test_piecewise_wavespeed_macro_elements(built):
    """Interface tents carry ndomains stacked copies (CalcTentMacroEl, twavetents.cpp:436-539): a wave that is a
    polynomial solution in each medium with matching traces is reproduced; here: u = t (dt u = 1, grad u = 0)."""
    c = lambda x: np.where(x[:, 0] < 0.5, 1.0, 2.0)  # noqa: E731
    f = lambda x, t: np.stack([np.broadcast_to(np.asarray(t, float), (len(x),)), 0 * x[:, 0], 0 * x[:, 0], 1 + 0 * x[:, 0]], axis=1)  # noqa: E731
    o = solve(ng.UnitSquareMesh(4), 3, c, 0.2, f, lambda x, t: 1 + 0 * t, nslabs=2)
    wf, exact = o.wavefront(), o.make_wavefront(f, 0.4)
    assert np.abs(wf - exact).max() < 1e-10
