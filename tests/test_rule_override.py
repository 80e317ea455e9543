"""Rule tables from a real NGSolve install (ngst_rule_override): host-side behaviour.
A valid table (here: the built-in rule with its points permuted, which is how NGSolve's ordering could differ) replaces
the rule in force; tables with the wrong point count, non-positive weights, outside points or insufficient exactness are
rejected; removing the override restores the built-in rule."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng


@pytest.fixture
def clean():
    yield
    for dim in (1, 2, 3):
        for order in (4, 8):
            ng.SetIntegrationRule(dim, order)


@pytest.mark.parametrize("dim,order", [(1, 8), (2, 8), (3, 8), (3, 4)])
def test_override_roundtrip(built, clean, dim, order):
    xi, w = ng.GetIntegrationRule(dim, order)
    perm = np.random.default_rng(dim * 10 + order).permutation(len(w))
    ng.SetIntegrationRule(dim, order, xi[perm], w[perm])
    x2, w2 = ng.GetIntegrationRule(dim, order)
    assert np.array_equal(x2, xi[perm]) and np.array_equal(w2, w[perm])
    # other (dim, order) pairs are untouched
    other = ng.GetIntegrationRule(dim, order + 2)
    assert len(other[1]) == ((order + 2) // 2 + 1) ** dim
    ng.SetIntegrationRule(dim, order)
    x3, w3 = ng.GetIntegrationRule(dim, order)
    assert np.array_equal(x3, xi) and np.array_equal(w3, w)


def test_invalid_tables_are_rejected(built, clean):
    xi, w = ng.GetIntegrationRule(2, 8)
    with pytest.raises(ng.NgstError):          # wrong number of points
        ng.SetIntegrationRule(2, 8, xi[:-1], w[:-1])
    bad = w.copy()
    bad[3] = -bad[3]
    with pytest.raises(ng.NgstError):          # non-positive weight
        ng.SetIntegrationRule(2, 8, xi, bad)
    out = xi.copy()
    out[0] = [0.9, 0.9]
    with pytest.raises(ng.NgstError):          # point outside the simplex
        ng.SetIntegrationRule(2, 8, out, w)
    moved = xi.copy()
    moved[5, 0] += 1e-6
    with pytest.raises(ng.NgstError):          # no longer exact
        ng.SetIntegrationRule(2, 8, moved, w)
    with pytest.raises(ValueError):            # bad dimension
        ng.SetIntegrationRule(4, 8, xi, w)
    from ngstrefftz_b200._capi import check, dptr, lib
    with pytest.raises(ng.NgstError):          # the same through the C ABI
        check(lib.ngst_rule_override(4, 8, len(w), dptr(xi), dptr(w)))
    with pytest.raises(ng.NgstError):          # negative order
        check(lib.ngst_rule_override(2, -2, len(w), dptr(xi), dptr(w)))
    with pytest.raises(ValueError):
        ng.SetIntegrationRule(2, 8, xi, w[:-1])
    x2, w2 = ng.GetIntegrationRule(2, 8)       # nothing was installed
    assert np.array_equal(x2, xi) and np.array_equal(w2, w)
