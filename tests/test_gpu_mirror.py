"""Output mirror (cuda/mirror.cu, ngst_twave_set_output_mirror): the host buffer filled row by row during Propagate is
bitwise what GetWavefront copies afterwards, for every launch scheme (one launch per level, dependency-driven single
launch, generic kernel, quasi-Trefftz) and for pageable (page-locked by the call) as well as pinned buffers."""
import math

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, standing

pytestmark = pytest.mark.gpu


def make(mesh, order, dt, init, D, mode=0):
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    tw.set_kernel_mode(mode)
    tw.SetInitial(init)
    tw.SetBoundaryCF(dirichlet(standing(D), D))
    return tw


CASES = {
    "1d_p6": (lambda: ng.SegMesh(64, 0, 1), 6, 0.3, standing(1), 1),
    "2d_p4": (lambda: ng.UnitSquareMesh(8), 4, 0.3, standing(2), 2),
    "2d_p4_first_order": (lambda: ng.UnitSquareMesh(6), 4, 0.3, first_order(standing(2)), 2),
    "3d_p4": (lambda: ng.UnitCubeMesh(4), 4, 0.2, standing(3), 3),
    "3d_p3": (lambda: ng.UnitCubeMesh(5), 3, 0.2, standing(3), 3),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_mirror_equals_get_wavefront(name, mode):
    mk, order, dt, init, D = CASES[name]
    tw = make(mk(), order, dt, init, D, mode)
    buf = np.full((tw.nel, tw.ncomp * tw.nip), np.nan)
    tw.SetOutputMirror(buf)
    for _ in range(3):
        tw.Propagate()
        assert tw.last_launches() > 0
        assert np.array_equal(buf, tw.GetWavefront())
    # switched off: the buffer is left alone and the solver carries on
    keep = buf.copy()
    tw.SetOutputMirror(None)
    tw.Propagate()
    assert np.array_equal(buf, keep)
    assert not np.array_equal(tw.GetWavefront(), keep)


def test_mirror_same_result_as_without():
    mk, order, dt, init, D = CASES["3d_p4"]
    a, b = make(mk(), order, dt, init, D), make(mk(), order, dt, init, D)
    buf = np.zeros((a.nel, a.ncomp * a.nip))
    a.SetOutputMirror(buf)
    for _ in range(2):
        a.Propagate()
        b.Propagate()
    assert np.array_equal(buf, b.GetWavefront())


def test_mirror_pinned_buffer():
    torch = pytest.importorskip("torch")
    mk, order, dt, init, D = CASES["2d_p4"]
    tw = make(mk(), order, dt, init, D)
    pin = torch.empty(tw.nel * tw.ncomp * tw.nip, dtype=torch.float64, pin_memory=True)
    buf = pin.numpy().reshape(tw.nel, -1)
    buf[:] = np.nan
    tw.SetOutputMirror(buf)
    tw.Propagate()
    assert np.array_equal(buf, tw.GetWavefront())
    tw.SetOutputMirror(None)


def test_mirror_quasi_trefftz():
    from test_oracle_qt import qt_exact
    mesh = ng.UnitSquareMesh(6)
    c = ng.PolynomialCF.affine([1.0, 1.0], 1.0)
    f = qt_exact(2)
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(0.1, True, 2 / 3)
    for generic in (False, True):
        tw = ng.TWave(4, ts, c, BBcf=1)
        if generic:
            tw.set_kernel_mode(1)
        tw.SetInitial(f)
        tw.SetBoundaryCF(lambda x, t: f(x, t)[:, 3])
        buf = np.zeros((tw.nel, tw.ncomp * tw.nip))
        tw.SetOutputMirror(buf)
        tw.Propagate()
        tw.Propagate()
        assert np.array_equal(buf, tw.GetWavefront())


def test_mirror_follows_set_initial_and_argument_errors():
    mk, order, dt, init, D = CASES["2d_p4"]
    ts = ng.TentSlab(mk(), "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    with pytest.raises(ng.NgstError):  # row length unknown before SetInitial
        from ngstrefftz_b200._capi import check, dptr, lib
        check(lib.ngst_twave_set_output_mirror(tw._h, dptr(np.zeros(8))))
    tw.SetInitial(init)
    tw.SetBoundaryCF(dirichlet(standing(D), D))
    with pytest.raises(ValueError):
        tw.SetOutputMirror(np.zeros(7))
    buf = np.zeros((tw.nel, tw.ncomp * tw.nip))
    tw.SetOutputMirror(buf)
    tw.SetInitial(init)          # same row length: the mirror stays
    tw.Propagate()
    assert np.array_equal(buf, tw.GetWavefront())
    keep = buf.copy()
    tw.SetInitial(first_order(init))   # shorter rows: the mirror is dropped
    tw.Propagate()
    assert np.array_equal(buf, keep)
    assert math.isfinite(tw.GetWavefront().sum())
