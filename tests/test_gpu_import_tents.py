"""The solver on imported tents (ngst_slab_import_tents): same tents -> bitwise the same wavefront as on the product's own
slab; the same tents in another valid serial numbering -> still bitwise the same (a tent's arithmetic does not depend on
its id); and a slab that is re-pitched or re-imported under a live solver is refused instead of run on stale tables."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, standing
from _oracle import Oracle
from ngstrefftz_b200._capi import NgstError
from test_import_tents import reorder_by_level

pytestmark = pytest.mark.gpu


def solve(ts, order, f, D, nslabs=2):
    tw = ng.TWave(order, ts, 1.0)
    tw.SetInitial(f)
    tw.SetBoundaryCF(dirichlet(f, D))
    for _ in range(nslabs):
        tw.Propagate()
    return tw, tw.GetWavefront()


@pytest.mark.parametrize("D,mk,order,dt", [(1, lambda: ng.SegMesh(16, 0, 1), 4, 0.3), (2, lambda: ng.UnitSquareMesh(5), 4, 0.3),
                                            (3, lambda: ng.UnitCubeMesh(4), 4, 0.2)])
def test_solver_on_imported_tents(built, D, mk, order, dt):
    f = standing(D)
    mesh = mk()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    _, want = solve(ts, order, f, D)
    e = ts.export()
    same = ng.TentSlab.from_reference(mesh, e, dt)
    _, got = solve(same, order, f, D)
    assert np.array_equal(got, want)
    if D > 1:
        other, _ = reorder_by_level(e)
        _, got2 = solve(ng.TentSlab.from_reference(mesh, other, dt), order, f, D)
        assert np.array_equal(got2, want)
    # the oracle on the same imported tents (its own import path) agrees to 1e-10
    o = Oracle(mesh)
    o.import_tents(same.export(), dt)
    o.setup(order, 1.0)
    o.set_initial(f)
    o.set_boundary_cf(dirichlet(f, D))
    o.propagate()
    o.propagate()
    ref = o.wavefront()
    assert np.abs(got - ref).max() <= 1e-10 * np.abs(ref).max()


def test_repitch_under_live_solver_is_refused(built):
    """the device tables are staged once; the reference re-reads its slab on every Propagate."""
    f = standing(2)
    mesh = ng.UnitSquareMesh(4)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.3, True, 2 / 3)
    tw = ng.TWave(3, ts, 1.0)
    tw.SetInitial(f)
    tw.SetBoundaryCF(dirichlet(f, 2))
    tw.Propagate()
    assert ts.PitchTents(0.1, True, 2 / 3)  # more, smaller tents
    with pytest.raises(NgstError) as ei:
        tw.Propagate()
    assert ei.value.code == -2
    tw2 = ng.TWave(3, ts, 1.0)  # a new solver for the new tents works
    tw2.SetInitial(f)
    tw2.SetBoundaryCF(dirichlet(f, 2))
    tw2.Propagate()
    assert np.isfinite(tw2.Energy())


def test_initial_data_switch_resizes_scratch(built):
    """SetInitial(first order) -> Error -> SetInitial(second order) -> Error must not overflow the scratch rows."""
    from _cases import first_order
    f = standing(2)
    mesh = ng.UnitSquareMesh(4)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.2, True, 2 / 3)
    tw = ng.TWave(3, ts, 1.0)
    tw.SetInitial(first_order(f))
    wf = tw.GetWavefront()
    assert tw.Error(wf, wf) == 0.0 and tw.Energy(wf) > 0
    tw.SetInitial(f)
    wf2 = tw.GetWavefront()
    assert wf2.shape[1] > wf.shape[1]
    assert tw.Error(wf2, wf2) == 0.0 and abs(tw.Energy(wf2) - tw.Energy()) <= 1e-14 * tw.Energy()
    tw.Propagate()
    assert np.isfinite(tw.Energy())
