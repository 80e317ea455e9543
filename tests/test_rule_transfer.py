"""ngst_rule_transfer: the matrix that carries state-rule point values to the compact rule through the degree-p
polynomials. Properties the fast kernel's planned compact right-hand-side pass rests on:
(1) weak identity for ARBITRARY point data g:  sum_q w_q phi(x_q) g_q = sum_r v_r phi(y_r) (M g)_r  for deg(phi) <= p;
(2) M reproduces polynomials of degree <= p at the compact points; (3) a second transfer changes nothing (projection)."""
import ctypes as C
from itertools import product

import numpy as np
import pytest

from ngstrefftz_b200._capi import check, dptr, iptr, lib


def rule(fn, dim, k):
    nip = C.c_int(0)
    check(fn(dim, k, C.byref(nip), None, None))
    xi, w = np.zeros((nip.value, dim)), np.zeros(nip.value)
    check(fn(dim, k, C.byref(nip), dptr(xi), dptr(w)))
    return xi, w


def transfer(dim, p):
    sizes = np.zeros(2, dtype=np.int32)
    check(lib.ngst_rule_transfer(dim, 2 * p, 2 * p, p, iptr(sizes), None))
    M = np.zeros((sizes[0], sizes[1]))
    check(lib.ngst_rule_transfer(dim, 2 * p, 2 * p, p, iptr(sizes), dptr(M)))
    return M


# (dim, p) for which a compact rule of degree 2p is tabulated (host/quadrature.hpp compact_rule_points)
@pytest.mark.parametrize("dim,p", [(2, 2), (2, 3), (2, 4), (2, 5), (3, 1), (3, 2), (3, 3), (3, 4)])
def test_transfer_identities(built, dim, p):
    xa, wa = rule(lib.ngst_rule, dim, 2 * p)
    xb, wb = rule(lib.ngst_compact_rule, dim, 2 * p)
    M = transfer(dim, p)
    assert M.shape == (len(wb), len(wa))
    rng = np.random.default_rng(100 * dim + p)
    g = rng.standard_normal(len(wa))          # arbitrary point data (e.g. sampled initial data)
    Mg = M @ g
    expo = [e for e in product(range(p + 1), repeat=dim) if sum(e) <= p]
    scale = np.abs(wa).sum() * np.abs(g).max()
    for e in expo:
        pa, pb = np.prod(xa ** np.array(e), axis=1), np.prod(xb ** np.array(e), axis=1)
        assert abs((wa * pa * g).sum() - (wb * pb * Mg).sum()) <= 1e-13 * scale, e
    # polynomial data: reproduced at the compact points
    coef = rng.standard_normal(len(expo))
    ga = sum(c * np.prod(xa ** np.array(e), axis=1) for c, e in zip(coef, expo))
    gb = sum(c * np.prod(xb ** np.array(e), axis=1) for c, e in zip(coef, expo))
    assert np.abs(M @ ga - gb).max() <= 1e-12 * np.abs(gb).max()
    # Energy-type quantities of polynomial states agree on both rules (integrand of degree 2p)
    assert abs((wa * ga * ga).sum() - (wb * gb * gb).sum()) <= 1e-13 * (wa * ga * ga).sum()


def test_transfer_size_query_and_errors(built):
    sizes = np.zeros(2, dtype=np.int32)
    check(lib.ngst_rule_transfer(3, 8, 8, 4, iptr(sizes), None))
    assert list(sizes) == [46, 125]
    from ngstrefftz_b200 import NgstError
    with pytest.raises(NgstError):
        check(lib.ngst_rule_transfer(5, 8, 8, 4, iptr(sizes), None))
    with pytest.raises(NgstError):   # 220 polynomials of degree <= 9 cannot be orthonormalised on the 8-point rule
        check(lib.ngst_rule_transfer(3, 2, 2, 9, iptr(sizes), np.zeros(4 * 8).ctypes.data_as(C.POINTER(C.c_double))))


def test_rows_written_by_tents_are_polynomials_of_degree_p(built):
    """Fact (ii) of DESIGN.md section 9: after a slab every state row is a polynomial of degree <= p on its element (one tent
    writes all points of a row), so projecting it onto P_p in the rule's inner product changes nothing -- checked on the
    oracle's Propagate; the sampled initial rows (a standing wave) are NOT polynomials."""
    import ngstrefftz_b200 as ng
    from _cases import dirichlet, standing
    from _oracle import Oracle
    for D, mesh, p, dt in ((2, ng.UnitSquareMesh(4), 4, 0.3), (3, ng.UnitCubeMesh(2), 3, 0.2)):
        o = Oracle(mesh)
        assert o.pitch(dt, True, 2 / 3, 1.0)
        o.setup(p, 1.0)
        o.set_initial(standing(D))
        o.set_boundary_cf(dirichlet(standing(D), D))
        xa, wa = rule(lib.ngst_rule, D, 2 * p)
        expo = [e for e in product(range(p + 1), repeat=D) if sum(e) <= p]
        V = np.stack([np.prod(xa ** np.array(e), axis=1) for e in expo], axis=1)
        # W-orthogonal projector onto P_p at the rule points
        G = V.T @ (wa[:, None] * V)
        P = V @ np.linalg.solve(G, V.T * wa[None, :])
        init = o.wavefront().reshape(o.nel, -1, len(wa))
        assert np.abs(init - init @ P.T).max() > 1e-6 * np.abs(init).max()
        o.propagate()
        rows = o.wavefront().reshape(o.nel, -1, len(wa))
        assert np.abs(rows - rows @ P.T).max() <= 1e-11 * np.abs(rows).max()
