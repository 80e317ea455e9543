"""Trefftz basis (src/trefftzfespace.cpp:611-682) and simplex rules: product vs oracle vs known answers (K2)."""
import math
from itertools import product

import numpy as np
import pytest

from _oracle import oracle_basis, reference_rule


def product_basis(D, order, fo):
    from ngstrefftz_b200._capi import lib, iptr, dptr, check
    sizes = np.zeros(3, np.int32)
    check(lib.ngst_basis(D, order, fo, iptr(sizes), None, None, None, None))
    N, npoly, nnz = (int(x) for x in sizes)
    rowptr, col, val = np.zeros(N + 1, np.int32), np.zeros(nnz, np.int32), np.zeros(nnz)
    expo = np.zeros((npoly, D + 1), np.int32)
    check(lib.ngst_basis(D, order, fo, iptr(sizes), iptr(rowptr), iptr(col), dptr(val), iptr(expo)))
    return N, npoly, rowptr, col, val, expo


# (D, p) -> (N, npoly, nnz)
SIZES = {(2, 4): (25, 35, 46), (1, 6): (13, 28, 28), (3, 3): (30, 35, 45), (2, 5): (36, 56, None), (3, 4): (55, 70, 103)}


@pytest.mark.parametrize("D,p", list(SIZES))
def test_sizes(built, D, p):
    N, npoly, rowptr, col, val, expo = product_basis(D, p, 0)
    eN, enp, ennz = SIZES[(D, p)]
    assert (N, npoly) == (eN, enp)
    assert N == math.comb(D + p, p) + math.comb(D + p - 1, p - 1)  # twavetents.hpp:160-161
    if ennz is not None:
        assert len(val) == ennz
    assert product_basis(D, p, 1)[0] == N - 1  # first-order system drops the constant (trefftzfespace.cpp:680)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("p", [1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("fo", [0, 1])
def test_product_equals_oracle(built, D, p, fo):
    N, npoly, rowptr, col, val, expo = product_basis(D, p, fo)
    orow, ocol, oval = oracle_basis(D, p, fo)
    assert np.array_equal(rowptr, orow) and np.array_equal(col, ocol) and np.array_equal(val, oval)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("p", [2, 4, 6])
def test_rows_are_wave_polynomials(built, D, p):
    """K2: every basis polynomial satisfies u_tt = Laplace(u) exactly (c = 1), in exact rational arithmetic on exponents."""
    N, npoly, rowptr, col, val, expo = product_basis(D, p, 0)
    # monomial order: first coordinate outermost (scalarmappedfe.cpp:526-531)
    want = [e for e in product(range(p + 1), repeat=D + 1) if sum(e) <= p]
    assert [tuple(r) for r in expo] == want
    idx = {tuple(e): i for i, e in enumerate(expo)}
    dense = np.zeros((N, npoly))
    for i in range(N):
        dense[i, col[rowptr[i]:rowptr[i + 1]]] = val[rowptr[i]:rowptr[i + 1]]
    assert np.linalg.matrix_rank(dense) == N
    for i in range(N):
        box = {}
        for j in range(npoly):
            a = dense[i, j]
            if a == 0:
                continue
            e = list(expo[j])
            for d in range(D + 1):
                if e[d] >= 2:
                    f = e.copy()
                    f[d] -= 2
                    sgn = -1.0 if d == D else 1.0
                    box[tuple(f)] = box.get(tuple(f), 0.0) + sgn * a * e[d] * (e[d] - 1)
        assert all(abs(v) < 1e-12 for v in box.values()), (i, box)


@pytest.mark.parametrize("dim", [1, 2, 3])
@pytest.mark.parametrize("order", [2, 6, 8, 10, 12])
def test_rule_matches_scipy_and_integrates_exactly(built, dim, order):
    from ngstrefftz_b200._capi import lib, iptr, dptr, check
    import ctypes as C
    nip = C.c_int(0)
    check(lib.ngst_rule(dim, order, C.byref(nip), None, None))
    xi, w = np.zeros((nip.value, dim)), np.zeros(nip.value)
    check(lib.ngst_rule(dim, order, C.byref(nip), dptr(xi), dptr(w)))
    rx, rw = reference_rule(dim, order)
    assert nip.value == (order // 2 + 1) ** dim == len(rw)
    assert np.allclose(xi, rx, rtol=0, atol=2e-15) and np.allclose(w, rw, rtol=1e-13, atol=1e-17)
    assert abs(w.sum() - 1 / math.factorial(dim)) < 1e-15
    # exactness on monomials of total degree <= order: int x^a = prod a_i! / (dim + |a|)!
    for e in product(range(order + 1), repeat=dim):
        if sum(e) > order:
            continue
        exact = math.prod(math.factorial(k) for k in e) / math.factorial(dim + sum(e))
        got = (w * np.prod(xi ** np.array(e), axis=1)).sum()
        assert abs(got - exact) < 1e-14, (e, got, exact)


@pytest.mark.parametrize("dim,degree,npts", [(2, 4, 6), (2, 6, 12), (2, 8, 16), (2, 10, 25), (3, 2, 4), (3, 4, 14), (3, 6, 24), (3, 8, 46)])
def test_compact_rules_are_exact_positive_and_interior(built, dim, degree, npts):
    """Compact symmetric rules of the fast kernel's top-face Gram (scripts/gen_compact_rules.py): exact for every monomial
    of total degree <= degree (integral of lambda^alpha = alpha! / (|alpha| + dim)!), positive weights, interior points."""
    import itertools
    import math
    from ngstrefftz_b200._capi import C, check, dptr, lib
    n = C.c_int(0)
    check(lib.ngst_compact_rule(dim, degree, C.byref(n), None, None))
    assert n.value == npts
    xi, w = np.zeros((npts, dim)), np.zeros(npts)
    check(lib.ngst_compact_rule(dim, degree, C.byref(n), dptr(xi), dptr(w)))
    lam = np.hstack([xi, 1 - xi.sum(axis=1, keepdims=True)])
    assert w.min() > 0 and lam.min() > 1e-3
    for alpha in itertools.product(range(degree + 1), repeat=dim + 1):
        if sum(alpha) <= degree:
            exact = math.prod(math.factorial(a) for a in alpha) / math.factorial(sum(alpha) + dim)
            assert abs(np.sum(w * np.prod(lam ** np.array(alpha), axis=1)) - exact) <= 1e-13 * exact
    # no table: falls back to the Duffy rule of that order
    check(lib.ngst_compact_rule(3, 10, C.byref(n), None, None))
    assert n.value == 216
