"""Compact state of the fast kernel (tent_fast.cu, CS = 1): the tents read and write an internal
copy of the state on the compact rule exact to degree 2p, the public order-2p rows are written by the last tent that
contains an element. It must reproduce the restated reference (the oracle, which integrates on the order-2p points) to
1e-10 and the public-row variant of the same kernel to round-off, in every launch scheme and over several slabs."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, neumann, standing
from _oracle import Oracle

pytestmark = pytest.mark.gpu


def run(mesh, order, dt, init, bnd, nslabs, compact, mode=0):
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    tw.set_compact_state(compact)
    tw.set_kernel_mode(mode)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    assert tw.compact_state() == compact
    en = [tw.Energy()]
    for _ in range(nslabs):
        tw.Propagate()
        en.append(tw.Energy())
    return tw, tw.GetWavefront(), en


CASES = {
    "3d_p4": (lambda: ng.UnitCubeMesh(4), 4, 0.2, lambda D: standing(D), lambda D: dirichlet(standing(D), D), False),
    "3d_p4_neumann": (lambda: ng.UnitCubeMesh(3), 4, 0.2, lambda D: standing(D), lambda D: neumann(standing(D), D), True),
    "3d_p4_fo": (lambda: ng.UnitCubeMesh(3), 4, 0.2, lambda D: first_order(standing(D)), lambda D: dirichlet(standing(D), D), False),
    "3d_p3": (lambda: ng.UnitCubeMesh(4), 3, 0.2, lambda D: standing(D), lambda D: dirichlet(standing(D), D), False),
    "2d_p4": (lambda: ng.UnitSquareMesh(6), 4, 0.3, lambda D: standing(D), lambda D: dirichlet(standing(D), D), False),
    "2d_p5_fo": (lambda: ng.UnitSquareMesh(5), 5, 0.3, lambda D: first_order(standing(D)), lambda D: dirichlet(standing(D), D), False),
}


@pytest.mark.parametrize("name", list(CASES))
def test_compact_state_matches_public_rows_and_oracle(built, name):
    mk, order, dt, init, bnd, neu = CASES[name]
    nslabs = 3

    def mesh():
        m = mk()
        if neu:
            m.SetBoundaryNames("neumann")
        return m

    D = mesh().dim
    tw1, wf1, en1 = run(mesh(), order, dt, init(D), bnd(D), nslabs, True)
    tw0, wf0, en0 = run(mesh(), order, dt, init(D), bnd(D), nslabs, False)
    scale = np.abs(wf0).max()
    d = np.abs(wf1 - wf0).max() / scale
    print(f"{name}: compact vs public rows {d:.3e}")
    assert d <= 1e-10, d   # measured 2e-14 .. 3e-11 (2-D order 5: scaled cond of the tent systems 1e6)
    for a, b in zip(en1, en0):
        assert abs(a - b) <= 1e-10 * abs(b)
    m = mesh()
    o = Oracle(m)
    assert o.pitch(dt, True, 2 / 3, 1.0)
    o.setup(order, 1.0)
    o.set_initial(init(D))
    o.set_boundary_cf(bnd(D))
    for _ in range(nslabs):
        o.propagate()
    ref = o.wavefront()
    rel = np.abs(wf1 - ref).max() / np.abs(ref).max()
    print(f"{name}: compact vs oracle {rel:.3e}")
    assert rel <= 1e-10, rel
    # the resident public state is what GetWavefront copies (rows of elements no tent touches included)
    assert abs(tw1.Energy() - tw1.Energy(wf1)) <= 1e-14 * abs(tw1.Energy(wf1))


def test_compact_state_launch_schemes_bitwise(built):
    """Per-level launches and the dependency-driven single launch give the same bits with the compact state."""
    f = standing(3)
    out = []
    for mode in (2, 3):
        tw, wf, _ = run(ng.UnitCubeMesh(8), 4, 0.08, f, dirichlet(f, 3), 2, True, mode)
        out.append(wf)
    assert np.array_equal(out[0], out[1])


def test_switching_modes_keeps_the_state_consistent(built):
    """A slab on the generic kernel (public rows) between compact slabs: the compact copy is rebuilt from the public rows."""
    f = standing(3)
    mesh = ng.UnitCubeMesh(4)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.15, True, 2 / 3)
    tw = ng.TWave(4, ts, 1.0)
    tw.SetInitial(f)
    tw.SetBoundaryCF(dirichlet(f, 3))
    tw.Propagate()
    tw.set_kernel_mode(1)   # generic DFMA kernel
    tw.Propagate()
    tw.set_kernel_mode(0)
    tw.Propagate()
    ref, wf_ref, _ = run(ng.UnitCubeMesh(4), 4, 0.15, f, dirichlet(f, 3), 3, False)
    wf = tw.GetWavefront()
    assert np.abs(wf - wf_ref).max() <= 1e-10 * np.abs(wf_ref).max()


def test_material_interfaces_switch_the_compact_state_off(built):
    """Slabs with tents on the generic kernel (more than one wavespeed region) keep the public rows only."""
    mesh = ng.UnitSquareMesh(4)
    c = lambda x: np.where(x[:, 0] < 0.5, 1.0, 1.5)  # noqa: E731
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(0.2, True, 2 / 3)
    tw = ng.TWave(3, ts, c)
    tw.SetInitial(standing(2))
    assert not tw.compact_state()


# ---------------------------------------------------------------------------------------------- factor cache (MODE 1 / 2)
@pytest.mark.parametrize("name", ["3d_p4", "3d_p4_neumann", "3d_p4_fo", "2d_p4"])
def test_factor_cache_matches_assembly_every_step_and_oracle(built, name):
    """The first Propagate stores the inverse of every tent matrix, later ones apply it (tent_fast.cu MODE 1 / 2): same
    result as assembling and solving in every Propagate up to cond * eps, the oracle's (inverse-then-multiply, as the
    reference) to 1e-10, and a slab repeated from the same data is bitwise the same whether it built or applied the cache."""
    mk, order, dt, init, bnd, neu = CASES[name]

    def mesh():
        m = mk()
        if neu:
            m.SetBoundaryNames("neumann")
        return m

    D = mesh().dim
    nslabs = 3
    ts = ng.TentSlab(mesh(), "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    out = {}
    for cache in (True, False):
        tw = ng.TWave(order, ts, 1.0)
        tw.set_factor_cache(cache)
        tw.SetInitial(init(D))
        tw.SetBoundaryCF(bnd(D))
        states = [tw.factor_cache()[0]]
        for _ in range(nslabs):
            tw.Propagate()
            states.append(tw.factor_cache()[0])
        out[cache] = tw.GetWavefront()
        if cache:
            assert states == [1, 2, 2, 2] and tw.factor_cache()[1] > 0
            # the same first slab again, now applying the stored inverses: bitwise what the building pass produced
            tw.reset_time()
            tw.SetInitial(init(D))
            tw.Propagate()
            again = tw.GetWavefront()
            tw2 = ng.TWave(order, ts, 1.0)
            tw2.SetInitial(init(D))
            tw2.SetBoundaryCF(bnd(D))
            tw2.Propagate()
            assert np.array_equal(again, tw2.GetWavefront())
        else:
            assert states == [0, 0, 0, 0]
    d = np.abs(out[True] - out[False]).max() / np.abs(out[False]).max()
    print(f"{name}: cached inverse vs solve every step {d:.3e}")
    assert d <= 5e-11, d
    o = Oracle(mesh())
    assert o.pitch(dt, True, 2 / 3, 1.0)
    o.setup(order, 1.0)
    o.set_initial(init(D))
    o.set_boundary_cf(bnd(D))
    for _ in range(nslabs):
        o.propagate()
    ref = o.wavefront()
    rel = np.abs(out[True] - ref).max() / np.abs(ref).max()
    print(f"{name}: cached inverse vs oracle {rel:.3e}")
    assert rel <= 1e-10, rel


def test_factor_cache_follows_the_system(built):
    """First- and second-order input have different bases: each keeps its own stored inverses."""
    f = standing(3)
    mesh = ng.UnitCubeMesh(3)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.2, True, 2 / 3)
    tw = ng.TWave(4, ts, 1.0)
    tw.SetBoundaryCF(dirichlet(f, 3))
    tw.SetInitial(f)
    tw.Propagate()
    assert tw.factor_cache()[0] == 2
    tw.SetInitial(first_order(f))
    assert tw.factor_cache()[0] == 1
    tw.Propagate()
    tw.SetInitial(f)
    assert tw.factor_cache()[0] == 2
    tw.reset_time()
    tw.Propagate()
    ref = ng.TWave(4, ts, 1.0)
    ref.set_factor_cache(False)
    ref.SetInitial(f)
    ref.SetBoundaryCF(dirichlet(f, 3))
    ref.Propagate()
    a, b = tw.GetWavefront(), ref.GetWavefront()
    assert np.abs(a - b).max() <= 5e-11 * np.abs(b).max()


def test_apply_variant_with_four_ctas_is_bitwise_the_default_one(built):
    """set_factor_cache(2): the apply kernel compiled for 128 registers / 55 KB (four CTAs per SM), reading the stored inverse
    from L2 instead of staging it in shared memory -- same arithmetic, same bits."""
    f = standing(3)
    mesh = ng.UnitCubeMesh(5)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.1, True, 2 / 3)
    out = []
    for variant in (True, 2):
        tw = ng.TWave(4, ts, 1.0)
        tw.set_factor_cache(variant)
        tw.SetInitial(f)
        tw.SetBoundaryCF(dirichlet(f, 3))
        for _ in range(3):
            tw.Propagate()
        out.append(tw.GetWavefront())
    assert np.array_equal(out[0], out[1])
