"""Host-side sampling of the boundary datum (TWaveTents._boundary_values): the Python logic around the three C-ABI calls
it makes, run here against a stand-in for the library (no GPU): the datum gets ONE read-only array of boundary points for
all slabs, times are the points' local times plus the slab's timeshift, Dirichlet / Neumann / mixed faces are laid out
as ngst_twave_set_boundary_values expects them (Dirichlet: Q values per face; Neumann: D x Q)."""
import ctypes as C

import numpy as np

import ngstrefftz_b200 as ng
from _cases import dirichlet, neumann, standing


class FakeLib:
    def __init__(self, xt, kind, timeshift=0.0):
        self.xt, self.kind, self.shift, self.point_calls = xt, kind, timeshift, 0

    def ngst_twave_boundary_npoints(self, h):
        return len(self.kind)

    def ngst_twave_boundary_points(self, h, xt_ptr, kind_ptr):
        self.point_calls += 1
        C.memmove(xt_ptr, self.xt.ctypes.data, self.xt.nbytes)
        C.memmove(kind_ptr, self.kind.ctypes.data, self.kind.nbytes)
        return 0

    def ngst_twave_timeshift(self, h):
        return self.shift


def solver(D, Q, datum):
    tw = object.__new__(ng.TWaveTents)
    tw._h, tw._bddatum, tw._bnd_pts, tw.D, tw.nip = None, datum, None, D, Q
    return tw


def points(D, nfaces, Q, kinds, seed=1):
    rng = np.random.default_rng(seed)
    xt = np.ascontiguousarray(rng.random((nfaces * Q, D + 1)))
    kind = np.repeat(np.asarray(kinds, dtype=np.int32), Q)
    return xt, kind


def test_dirichlet_values_and_one_read_only_point_array(monkeypatch):
    D, Q, nf = 2, 4, 3
    xt, kind = points(D, nf, Q, [0, 0, 0])
    fake = FakeLib(xt, kind, timeshift=0.5)
    monkeypatch.setattr(ng, "lib", fake)
    seen = []
    f = standing(D)

    def datum(x, t):
        seen.append(x)
        return dirichlet(f, D)(x, t)

    tw = solver(D, Q, datum)
    v0 = tw._boundary_values()          # the solver's current timeshift
    v1 = tw._boundary_values(1.25)      # an explicit one (bench.py pre-samples the slabs of a run)
    assert fake.point_calls == 1 and seen[0] is seen[1] and not seen[0].flags.writeable
    assert np.array_equal(seen[0], xt[:, :D])
    assert np.array_equal(v0, f(xt[:, :D], xt[:, D] + 0.5)[:, D + 1])
    assert np.array_equal(v1, f(xt[:, :D], xt[:, D] + 1.25)[:, D + 1])


def test_neumann_and_mixed_layout(monkeypatch):
    D, Q, nf = 2, 3, 4
    f = standing(D)
    xt, kind = points(D, nf, Q, [1, 1, 1, 1], seed=2)
    monkeypatch.setattr(ng, "lib", FakeLib(xt, kind))
    v = solver(D, Q, neumann(f, D))._boundary_values()
    full = f(xt[:, :D], xt[:, D])
    want = np.concatenate([full[fc * Q:(fc + 1) * Q, 1:D + 1].T.reshape(-1) for fc in range(nf)])
    assert np.array_equal(v, want)
    # mixed faces with a dict datum: face by face, Dirichlet Q values, Neumann D x Q
    xt, kind = points(D, nf, Q, [0, 1, 0, 1], seed=3)
    monkeypatch.setattr(ng, "lib", FakeLib(xt, kind))
    v = solver(D, Q, {"dirichlet": dirichlet(f, D), "neumann": neumann(f, D)})._boundary_values()
    full = f(xt[:, :D], xt[:, D])
    parts = []
    for fc, k in enumerate([0, 1, 0, 1]):
        rows = full[fc * Q:(fc + 1) * Q]
        parts.append(rows[:, D + 1] if k == 0 else rows[:, 1:D + 1].T.reshape(-1))
    assert np.array_equal(v, np.concatenate(parts))


def test_no_datum_or_no_boundary(monkeypatch):
    xt, kind = points(1, 2, 2, [0, 0])
    monkeypatch.setattr(ng, "lib", FakeLib(xt, kind))
    assert solver(1, 2, None)._boundary_values() is None
    monkeypatch.setattr(ng, "lib", FakeLib(xt[:0], kind[:0]))
    assert solver(1, 2, dirichlet(standing(1), 1))._boundary_values() is None
