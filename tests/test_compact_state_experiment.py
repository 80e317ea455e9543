"""CPU proof of the fast kernel's planned compact state: the oracle's Propagate with the bottom-face
terms integrated on the compact rule, state rows transferred with ngst_rule_transfer's matrix, equals the restated
reference algorithm up to round-off -- for sampled (non-polynomial) initial rows, rows written by earlier tents,
Dirichlet and Neumann boundaries, first- and second-order form, 1 to 3 slabs."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, neumann, standing
from _oracle import Oracle
from test_rule_transfer import rule, transfer
from ngstrefftz_b200._capi import lib

CASES = {
    "2d_p4": (lambda: ng.UnitSquareMesh(4), 2, 4, 0.3, False, False, 2),
    "2d_p5_neumann": (lambda: ng.UnitSquareMesh(3), 2, 5, 0.3, True, False, 2),
    "2d_p3_first_order": (lambda: ng.UnitSquareMesh(4), 2, 3, 0.3, False, True, 3),
    "3d_p3": (lambda: ng.UnitCubeMesh(2), 3, 3, 0.2, False, False, 2),
    "3d_p4": (lambda: ng.UnitCubeMesh(2), 3, 4, 0.2, False, False, 1),
    "3d_p2_neumann_fo": (lambda: ng.UnitCubeMesh(2), 3, 2, 0.2, True, True, 2),
}


@pytest.mark.parametrize("name", list(CASES))
def test_compact_bottom_face_terms_reproduce_the_reference(built, name):
    mk, D, p, dt, neu, fo, nslabs = CASES[name]
    f = standing(D)

    def run(compact):
        mesh = mk()
        if neu:
            mesh.SetBoundaryNames("neumann")
        o = Oracle(mesh)
        assert o.pitch(dt, True, 2 / 3, 1.0)
        o.setup(p, 1.0)
        o.set_initial(first_order(f) if fo else f)
        o.set_boundary_cf(neumann(f, D) if neu else dirichlet(f, D))
        if compact:
            xb, wb = rule(lib.ngst_compact_rule, D, 2 * p)
            assert len(wb) < o.nip          # the compact rule really is smaller
            o.set_compact_rhs(xb, wb, transfer(D, p))
        for _ in range(nslabs):
            o.propagate(nthreads=2)
        return o.wavefront()

    ref, got = run(False), run(True)
    assert np.abs(got - ref).max() <= 1e-10 * np.abs(ref).max()
    assert not np.array_equal(got, ref)     # (a different summation, not the same code path)
