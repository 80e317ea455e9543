"""Device side of the rule override (ngst_rule_override): with a permuted rule installed, the solver agrees with the
oracle run on the same permuted rule, and its state is the default-rule state with the columns permuted.
(First run on a B200 in round 2, profiles/r04a_pytest_gpu.txt; the host side is covered by tests/test_rule_override.py.)"""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, standing
from _oracle import Oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D,mk,order,dt", [(1, lambda: ng.SegMesh(16, 0, 1), 4, 0.3), (2, lambda: ng.UnitSquareMesh(4), 4, 0.3),
                                            (3, lambda: ng.UnitCubeMesh(3), 4, 0.2), (3, lambda: ng.UnitCubeMesh(3), 3, 0.2)])
def test_permuted_rule(D, mk, order, dt):
    f, bnd = standing(D), dirichlet(standing(D), D)

    def run():
        ts = ng.TentSlab(mk(), "edge")
        ts.SetMaxWavespeed(1.0)
        assert ts.PitchTents(dt, True, 2 / 3)
        tw = ng.TWave(order, ts, 1.0)
        tw.SetInitial(f)
        tw.SetBoundaryCF(bnd)
        tw.Propagate()
        return tw, tw.GetWavefront()

    _, base = run()
    xi, w = ng.GetIntegrationRule(D, 2 * order)
    perm = np.random.default_rng(7).permutation(len(w))
    try:
        ng.SetIntegrationRule(D, 2 * order, xi[perm], w[perm])
        tw, wf = run()
        o = Oracle(mk())
        assert o.pitch(dt, True, 2 / 3, 1.0)
        o.setup(order, 1.0, rule=(xi[perm], w[perm]))
        o.set_initial(f)
        o.set_boundary_cf(bnd)
        o.propagate()
        ref = o.wavefront()
        assert np.abs(wf - ref).max() <= 1e-10 * np.abs(ref).max()
        nip = len(w)
        cols = (np.arange(tw.ncomp)[:, None] * nip + perm[None, :]).reshape(-1)
        assert np.abs(wf - base[:, cols]).max() <= 1e-11 * np.abs(base).max()
        assert abs(tw.Energy() - tw.Energy(wf)) <= 1e-14 * tw.Energy(wf)
    finally:
        ng.SetIntegrationRule(D, 2 * order)
