"""Quasi-Trefftz oracle (oracle/oracle_qt.cpp restating QTWaveTents / QTWaveBasis / CalcDDWaveOperator) against the
reference's own doctest digits (test/tents.py:133-157, stored in tests/golden/reference_doctest_pins.json) and against
method-independent properties. No GPU."""
import json
import math
import os

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from ngstrefftz_b200._poly import PolynomialCF, monomial_exponents
from _oracle import Oracle

HERE = os.path.dirname(os.path.abspath(__file__))
CA = 2.5


def qt_exact(D):
    """(u, grad u, dt u) of test/tents.py:166-191: u = (x [+ y] + 1)^ca exp(-sqrt(D ca (ca-1)) t), c = x [+ y] + 1."""
    k = math.sqrt(D * CA * (CA - 1))

    def f(x, t):
        t = np.broadcast_to(np.asarray(t, float), (len(x),))
        s = x.sum(axis=1) + 1
        e = np.exp(-k * t)
        return np.stack([s ** CA * e] + [CA * s ** (CA - 1) * e] * D + [-k * s ** CA * e], axis=1)

    return f


def match(pattern, value):
    """doctest ELLIPSIS match of repr(value)"""
    r = repr(float(value))
    head, _, tail = pattern.partition("...")
    return r.startswith(head) and r.endswith(tail)


def run(mesh, order, dt, qt):
    D = mesh.dim
    c = PolynomialCF.affine([1.0] * D, 1.0)
    f = qt_exact(D)
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, c(o.centroids))
    if qt:
        o.setup_qt(order, c)
    else:
        o.setup(order, c)
        o.set_point_wavespeed(c)
    o.set_initial(f)
    o.set_boundary_cf(lambda x, t: f(x, t)[:, D + 1])
    o.propagate()
    return o.error(o.wavefront(), o.make_wavefront(f, dt))


@pytest.mark.parametrize("name,qt", [("TestQTrefftz_1d_qt", True), ("TestQTrefftz_1d_plain", False)])
def test_reference_doctest_digits_qtrefftz(built, name, qt):
    pins = json.load(open(os.path.join(HERE, "golden", "reference_doctest_pins.json")))[name]
    for (n, a, b), pattern, stored in zip(pins["meshes"], pins["pattern"], pins["oracle"]):
        err = run(ng.SegMesh(n, a, b), pins["order"], pins["dt"], qt)
        assert match(pattern, err), (pattern, err)
        assert abs(err - stored) <= 1e-9 * stored  # the committed oracle value (same arithmetic, any host)


def test_taylor_tables_against_finite_differences():
    c = PolynomialCF(2, {(0, 0): 1.3, (1, 0): 0.7, (0, 1): -0.2, (1, 1): 0.1, (2, 0): 0.05})
    x0 = np.array([[0.3, 0.4], [1.0, -0.5]])
    tab = c.taylor_inverse_square(x0, 3)
    expo = monomial_exponents(2, 3)
    g = lambda x: 1.0 / c(x) ** 2  # noqa: E731
    h = 1e-3
    for j, (a, b) in enumerate(expo):
        if a + b > 2:
            continue
        # central differences of order a in x and b in y
        def dx(fun, n, d):
            if n == 0:
                return fun
            e = np.zeros(2)
            e[d] = h
            return dx(lambda x: (fun(x + e) - fun(x - e)) / (2 * h), n - 1, d)
        val = dx(dx(g, a, 0), b, 1)(x0) / (math.factorial(a) * math.factorial(b))
        assert np.allclose(tab[:, j], val, rtol=1e-4, atol=1e-6)
    # B = c itself: exact shift of a polynomial
    assert np.allclose(c.taylor(x0, 2)[:, 0], c(x0))


def test_qt_basis_reduces_to_trefftz_for_constant_coefficients(built):
    """With 1/c^2 = const and B = 1 the quasi-Trefftz recursion (trefftzfespace.cpp:1109-1185) yields exact wave polynomials:
    every basis function satisfies A u_tt = Laplace u."""
    from _oracle import lib, _dp, _ip
    D, p, A0 = 2, 4, 0.49
    nA, nB = len(monomial_exponents(D, p - 2)), len(monomial_exponents(D, p - 1))
    aa, bb = np.zeros(nA), np.zeros(nB)
    aa[0], bb[0] = A0, 1.0
    N = math.comb(D + p, p) + math.comb(D + p - 1, p - 1)
    nnz = lib.oracle_qt_basis(D, p, _dp(aa), _dp(bb), None, None, None)
    rowptr, col, val = np.zeros(N + 1, np.int32), np.zeros(nnz, np.int32), np.zeros(nnz)
    lib.oracle_qt_basis(D, p, _dp(aa), _dp(bb), _ip(rowptr), _ip(col), _dp(val))
    expo = monomial_exponents(D + 1, p)
    idx = {e: i for i, e in enumerate(expo)}
    for i in range(N):
        row = {expo[col[j]]: val[j] for j in range(rowptr[i], rowptr[i + 1])}
        res = {}
        for e, v in row.items():
            for d in range(D):
                if e[d] >= 2:
                    k = tuple(e[m] - 2 * (m == d) for m in range(D + 1))
                    res[k] = res.get(k, 0.0) + v * e[d] * (e[d] - 1)
            if e[D] >= 2:
                k = tuple(e[m] - 2 * (m == D) for m in range(D + 1))
                res[k] = res.get(k, 0.0) - A0 * v * e[D] * (e[D] - 1)
        assert max((abs(v) for v in res.values()), default=0.0) < 1e-12
    assert len(idx) == math.comb(D + 1 + p, p)


def test_qt_convergence_2d(built):
    """2-D quasi-Trefftz on structured meshes, c = x + y + 1 (test/tents.py:184-191 on netgen meshes: 0.004.., 0.001.., ..e-05)."""
    errs = [run(ng.UnitSquareMesh(n), 4, 0.25, True) for n in (2, 4, 8)]
    rates = [math.log2(errs[i] / errs[i + 1]) for i in range(2)]
    assert errs[0] < 2e-2 and min(rates) > 2.5, (errs, rates)
    plain = run(ng.UnitSquareMesh(8), 4, 0.25, False)
    assert errs[2] < 0.1 * plain  # quasi-Trefftz beats the piecewise-constant-speed Trefftz basis by far
