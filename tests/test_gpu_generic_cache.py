"""Factor cache of the generic kernel (tent_generic.cu gmode 1 / 2): solvers whose tents all run on the generic kernel with
systems of up to 32 unknowns -- 1-D meshes, incl. the material-interface (macro-element) tents of BASELINE configs[1] --
keep the inverse of every tent matrix after the first Propagate and afterwards only build right-hand sides, apply the
stored inverse and evaluate."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, gaussian_1d, neumann, standing
from _oracle import Oracle
from test_gpu_parity import forced_tol

pytestmark = pytest.mark.gpu


def pc(x):
    return np.where((np.floor(x[:, 0] * 4).astype(int) % 2) == 0, 1.0, 2.0)


CASES = {
    "1d_p4": (lambda: ng.SegMesh(32, 0, 1), 4, 1.0, 0.3, standing(1), dirichlet(standing(1), 1), False),
    "1d_p6_piecewise": (lambda: ng.SegMesh(64, 0, 1), 6, pc, 0.2, gaussian_1d, lambda x, t: 0 * t, False),
    "1d_p4_neumann": (lambda: ng.SegMesh(24, 0, 1), 4, 1.0, 0.3, standing(1), neumann(standing(1), 1), True),
    "1d_p5_fo": (lambda: ng.SegMesh(24, 0, 1), 5, 1.0, 0.3, first_order(standing(1)), dirichlet(standing(1), 1), False),
}


@pytest.mark.parametrize("name", list(CASES))
def test_generic_factor_cache(built, name):
    mk, order, c, dt, init, bnd, neu = CASES[name]

    def mesh():
        m = mk()
        if neu:
            m.SetBoundaryNames("neumann")
        return m

    nslabs = 3
    m0 = mesh()
    ts = ng.TentSlab(m0, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    out = {}
    for cache in (True, False):
        tw = ng.TWave(order, ts, c)
        tw.set_factor_cache(cache)
        tw.SetInitial(init)
        tw.SetBoundaryCF(bnd)
        assert tw.kernel_name() == "tent_kernel_generic"
        states = [tw.factor_cache()[0]]
        for _ in range(nslabs):
            tw.Propagate()
            states.append(tw.factor_cache()[0])
        out[cache] = tw.GetWavefront()
        assert states == ([1, 2, 2, 2] if cache else [0, 0, 0, 0])
        if cache:
            assert tw.factor_cache()[1] > 0
            tw.reset_time()
            tw.SetInitial(init)
            tw.Propagate()           # applies the stored inverses
            again = tw.GetWavefront()
            tw2 = ng.TWave(order, ts, c)
            tw2.SetInitial(init)
            tw2.SetBoundaryCF(bnd)
            tw2.Propagate()          # builds them
            assert np.array_equal(again, tw2.GetWavefront())
    o = Oracle(mesh())
    assert o.pitch(dt, True, 2 / 3, c(o.centroids) if callable(c) else c)
    o.setup(order, c)
    o.set_initial(init)
    o.set_boundary_cf(bnd)
    for _ in range(nslabs):
        o.propagate()
    ref = o.wavefront()
    tol = forced_tol(o, name)
    d = np.abs(out[True] - out[False]).max() / np.abs(out[False]).max()
    rel = np.abs(out[True] - ref).max() / np.abs(ref).max()
    print(f"{name}: stored inverse vs LU every step {d:.3e}, vs oracle {rel:.3e} (tolerance {tol:.1e})")
    assert d <= tol and rel <= tol


WARP_CASES = dict(CASES)
# levels of several hundred tents (several CTAs of four warps, a partly filled last one), material-interface tents among them
WARP_CASES["1d_p6_piecewise_1000"] = (lambda: ng.SegMesh(1000, 0, 1), 6, lambda x: np.where((np.floor(x[:, 0] * 10).astype(int) % 2) == 0, 1.0, 2.0),
                                      0.01, gaussian_1d, lambda x, t: 0 * t, False)


# outside the 1-D kernel's range the default stays the CTA variant; the CTA code on a warp (5) is bitwise the same there too
WARP_CASES["2d_p2"] = (lambda: ng.UnitSquareMesh(5), 2, 1.0, 0.2, standing(2), dirichlet(standing(2), 2), False)
WARP_CASES["3d_p1"] = (lambda: ng.UnitCubeMesh(3), 1, 1.0, 0.2, standing(3), dirichlet(standing(3), 3), False)
WARP_CASES["1d_p8"] = (lambda: ng.SegMesh(24, 0, 1), 8, 1.0, 0.3, standing(1), dirichlet(standing(1), 1), False)


@pytest.mark.parametrize("name", list(WARP_CASES))
def test_warp_per_tent_apply_is_bitwise_the_cta_one(built, name):
    """The stored inverses of the generic kernel applied by one WARP per tent -- set_factor_cache(3), the default: in 1-D at
    orders 4-6 the lanes-as-functions kernel tent_kernel_small1d (a CTA per tent elsewhere); 5: tent_kernel_generic<D, true>, the
    statements of the CTA variant executed by 32 lanes with warp-level barriers. Same arithmetic in the same order: same bits as
    the CTA variant (4)."""
    mk, order, c, dt, init, bnd, neu = WARP_CASES[name]
    m = mk()
    if neu:
        m.SetBoundaryNames("neumann")
    ts = ng.TentSlab(m, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    out, names = [], []
    for variant in (4, 3, 5):
        tw = ng.TWave(order, ts, c)
        tw.set_factor_cache(variant)
        tw.SetInitial(init)
        tw.SetBoundaryCF(bnd)
        for _ in range(3):
            tw.Propagate()
        names.append(tw.kernel_name())
        out.append(tw.GetWavefront())
    auto = "tent_kernel_small1d" if (m.dim == 1 and 4 <= order <= 6) else "tent_kernel_generic"
    assert names == ["tent_kernel_generic", auto, "tent_kernel_generic_warp"], names
    dflt = ng.TWave(order, ts, c)   # no selection: the automatic choice
    dflt.SetInitial(init)
    dflt.SetBoundaryCF(bnd)
    for _ in range(3):
        dflt.Propagate()
    assert dflt.kernel_name() == auto and np.array_equal(dflt.GetWavefront(), out[0])
    assert np.isfinite(out[0]).all() and np.array_equal(out[0], out[1]) and np.array_equal(out[0], out[2])

