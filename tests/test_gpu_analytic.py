"""Device-side sampling of analytic data (cuda/sample.cu through ngst_twave_*_analytic):
the samples agree with the host evaluation of the same formulas, and a Propagate fed from device-sampled initial and
boundary data agrees with the oracle fed from host-evaluated data.

Tolerances: samples 1e-13 relative to the largest sample (device sin/cos/exp differ from libm by a few ulp; the
polynomial family is exact up to summation order); propagated state 1e-10 relative (BASELINE.json north_star)."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from ngstrefftz_b200._analytic import poly_solution_terms
from _cases import dirichlet, neumann
from _oracle import Oracle

pytestmark = pytest.mark.gpu
SAMPLE_TOL = 1e-13
TOL = 1e-10


def solver(mesh, order, c, dt, neu=False):
    if neu:
        mesh.SetBoundaryNames("neumann")
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    return ts, ng.TWave(order, ts, c)


DATA = {
    "standing1": lambda: (ng.SegMesh(16, 0, 1), ng.StandingWave(1, 1.0), 1.0),
    "standing2": lambda: (ng.UnitSquareMesh(4), ng.StandingWave(2, 1.0), 1.0),
    "standing3": lambda: (ng.UnitCubeMesh(2), ng.StandingWave(3, 1.3), 1.3),
    "pulse2": lambda: (ng.UnitSquareMesh(5), ng.PlanePulse(2, 1.0, 20.0, 0.3, [1.0, 0.5]), 1.0),
    "pulse3": lambda: (ng.UnitCubeMesh(2), ng.PlanePulse(3, 0.8, 10.0, 0.2, [1.0, -0.5, 0.25]), 0.8),
    "poly1": lambda: (ng.SegMesh(8, 0, 1), ng.PolynomialWave(1, poly_solution_terms(1, 1.0)), 1.0),
    "poly2": lambda: (ng.UnitSquareMesh(3), ng.PolynomialWave(2, poly_solution_terms(2, 1.0)), 1.0),
    "poly3": lambda: (ng.UnitCubeMesh(2), ng.PolynomialWave(3, poly_solution_terms(3, 1.0)), 1.0),
}


@pytest.mark.parametrize("name", list(DATA))
@pytest.mark.parametrize("fo", [False, True])
def test_make_wavefront_matches_host(name, fo):
    mesh, data, c = DATA[name]()
    _, tw = solver(mesh, 3, c, 0.2)
    d = data.FirstOrder() if fo else data
    for time in (0.0, 0.37):
        dev = tw.MakeWavefront(d, time)              # device
        host = tw.MakeWavefront(lambda x, t: d(x, t), time)  # host evaluation of the same formulas
        assert dev.shape == host.shape
        assert np.abs(dev - host).max() <= SAMPLE_TOL * max(np.abs(host).max(), 1.0)


@pytest.mark.parametrize("name", list(DATA))
@pytest.mark.parametrize("neu", [False, True])
@pytest.mark.parametrize("fo", [False, True])
def test_propagate_device_data_vs_oracle(name, neu, fo):
    mesh, data, c = DATA[name]()
    D = data.D
    if neu and D == 1 and name != "standing1":
        pytest.skip("one Neumann case in 1-D is enough")
    order, dt, nslabs = (4, 0.2, 2)
    ts, tw = solver(mesh, order, c, dt, neu)
    init = data.FirstOrder() if fo else data
    tw.SetInitial(init)
    tw.SetBoundaryCF(data)
    assert tw.fosystem == int(fo)
    o = Oracle(mesh)
    assert o.pitch(dt, True, 2 / 3, c)
    o.setup(order, c)
    o.set_initial(lambda x, t: init(x, t))
    o.set_boundary_cf(neumann(data, D) if neu else dirichlet(data, D))
    for _ in range(nslabs):
        tw.Propagate()
        o.propagate()
    assert tw.last_launches() > 0
    wf, ref = tw.GetWavefront(), o.wavefront()
    rel = np.abs(wf - ref).max() / np.abs(ref).max()
    assert rel <= TOL, rel
    # the device-sampled slab equals the slab fed with host-sampled values of the same datum
    tw2 = ng.TWave(order, ts, c)
    tw2.SetInitial(lambda x, t: init(x, t))
    tw2.SetBoundaryCF(neumann(data, D) if neu else dirichlet(data, D))
    for _ in range(nslabs):
        tw2.Propagate()
    wf2 = tw2.GetWavefront()
    # (device and host evaluate sin / cos / powers with different last bits; both runs then go through the same kernels)
    assert np.abs(wf - wf2).max() <= 1e-11 * np.abs(wf2).max()


def test_polynomial_is_reproduced_exactly():
    """K1 (Trefftz exactness) with everything sampled on the device: a degree-3 polynomial solution stays exact."""
    mesh, data, c = DATA["poly2"]()
    _, tw = solver(mesh, 4, c, 0.3)
    tw.SetInitial(data)
    tw.SetBoundaryCF(data)
    tw.Propagate()
    tw.Propagate()
    exact = tw.MakeWavefront(data, tw.timeshift())
    wf = tw.GetWavefront()
    assert tw.Error(wf, exact) <= 1e-10 * np.sqrt(2 * tw.Energy(exact))


@pytest.mark.parametrize("name", ["standing1", "standing2", "pulse3", "poly2"])
@pytest.mark.parametrize("fo", [False, True])
def test_error_against_device_sampled_exact_solution(name, fo):
    """Error / L2Error of the resident state vs analytic data, all on the device, equal the host-array route."""
    mesh, data, c = DATA[name]()
    _, tw = solver(mesh, 4, c, 0.2)
    tw.SetInitial(data.FirstOrder() if fo else data)
    tw.SetBoundaryCF(data)
    tw.Propagate()
    d = data.FirstOrder() if fo else data
    wf = tw.GetWavefront()
    exact = tw.MakeWavefront(lambda x, t: d(x, t), tw.timeshift())
    want = tw.Error(wf, exact)
    got = tw.Error(None, data)
    scale = np.sqrt(np.sum(exact ** 2)) + 1.0
    assert abs(got - want) <= 1e-12 * scale, (got, want)
    assert abs(tw.Error(wf, data) - want) <= 1e-12 * scale
    if not fo:
        assert abs(tw.L2Error(None, data) - tw.L2Error(wf, exact)) <= 1e-12 * scale
    else:
        with pytest.raises(ng.NgstError):
            tw.L2Error(None, data)
    # at another time the error is the distance of the two exact samples (triangle inequality, loose check)
    assert tw.Error(None, data, time=0.0) > 0 or name.startswith("poly")


def test_values_then_analytic_again():
    """propagate_with_values replaces the device-side datum for one slab; the next Propagate samples it again."""
    mesh, data, c = DATA["standing2"]()
    ts, tw = solver(mesh, 4, c, 0.25)
    ref = ng.TWave(4, ts, c)
    for s in (tw, ref):
        s.SetInitial(data)
        s.SetBoundaryCF(data)
    ref.Propagate()
    ref.Propagate()
    host = ng.TWave(4, ts, c)
    host.SetBoundaryCF(dirichlet(data, 2))
    vals = host.sample_boundary(0.0)
    tw.propagate_with_values(vals)
    tw.Propagate()
    a, b = tw.GetWavefront(), ref.GetWavefront()
    assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max()


def test_argument_errors():
    mesh, data, c = DATA["standing2"]()
    _, tw = solver(mesh, 3, c, 0.25)
    with pytest.raises(ValueError):
        tw.SetInitial(ng.StandingWave(3))
    from ngstrefftz_b200._capi import check, dptr, lib
    bad = np.array([1.0, 2.0])
    with pytest.raises(ng.NgstError):
        check(lib.ngst_twave_set_initial_analytic(tw._h, 0, dptr(bad), 2, 0))
    with pytest.raises(ng.NgstError):
        check(lib.ngst_twave_set_boundary_analytic(tw._h, 7, dptr(bad), 2))
