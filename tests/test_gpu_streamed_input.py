"""Streamed input (ngst_twave_set_initial_streamed): the copy engine delivers the caller's host buffer in chunks of
consecutive elements while Propagate runs as a spatial sweep -- one dependency-driven launch, tents ticketed in the order
in which their dependency cones are complete, compact rows of first-touched elements built on the fly, finished chunks
copied out when an output mirror is set. Whatever the launch scheme (the sweep, or the fallbacks that wait for the whole
input), the result must be bitwise what SetInitial + Propagate gives."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, standing

pytestmark = pytest.mark.gpu


def make(D, n, order, dt, c=1.0):
    mesh = {1: lambda: ng.SegMesh(n, 0, 1), 2: lambda: ng.UnitSquareMesh(n), 3: lambda: ng.UnitCubeMesh(n)}[D]()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    return mesh, ts


@pytest.mark.parametrize("D,n,order,dt,fo,mode,compact", [
    (3, 6, 4, 0.1, False, 0, True), (3, 6, 4, 0.1, False, 2, True), (3, 6, 4, 0.1, False, 3, True), (3, 5, 4, 0.1, True, 0, True),
    (3, 5, 3, 0.1, False, 0, False), (2, 10, 4, 0.15, False, 0, True), (1, 40, 4, 0.2, False, 0, True), (3, 4, 4, 0.1, False, 1, True)])
def test_streamed_input_is_bitwise_the_plain_path(built, D, n, order, dt, fo, mode, compact):
    f = standing(D)
    init = first_order(f) if fo else f
    mesh, ts = make(D, n, order, dt)
    out = []
    for streamed in (False, True):
        tw = ng.TWave(order, ts, 1.0)
        tw.set_kernel_mode(mode)
        tw.set_compact_state(compact)
        tw.SetBoundaryCF(dirichlet(f, D))
        wf0 = tw.MakeWavefront(init, 0.0)
        for slab in range(3):   # slab 0 builds the factor cache, later ones apply it; every slab restarts from wf0
            tw.reset_time()
            if streamed:
                tw.SetInitialStreamed(wf0)
                if slab == 1:
                    assert np.array_equal(tw.GetWavefront(), wf0)  # reading the state waits for the whole input
            else:
                tw.SetInitial(wf0)
            tw.Propagate()
        out.append((tw.GetWavefront(), tw.Energy()))
    assert np.array_equal(out[0][0], out[1][0])
    assert out[0][1] == out[1][1]


def test_streamed_input_with_output_mirror_and_pinned_buffers(built):
    import torch
    f = standing(3)
    mesh, ts = make(3, 8, 4, 0.08)
    tw = ng.TWave(4, ts, 1.0)
    tw.SetBoundaryCF(dirichlet(f, 3))
    wf0 = tw.MakeWavefront(f, 0.0)
    tw.SetInitial(wf0)
    tw.Propagate()
    want = tw.GetWavefront()
    pin_in = torch.empty(wf0.shape, dtype=torch.float64, pin_memory=True)
    pin_in.numpy()[:] = wf0
    pin_out = torch.zeros(wf0.shape, dtype=torch.float64, pin_memory=True)
    tw.reset_time()
    tw.SetInitialStreamed(pin_in.numpy())
    tw.SetOutputMirror(pin_out.numpy())
    tw.Propagate()
    assert np.array_equal(pin_out.numpy(), want)
    # the same buffer again (no re-registration), other contents: the new rows are what the tents see
    pin_in.numpy()[:] = 0.5 * wf0
    pin_out.zero_()
    tw.reset_time()
    tw.SetInitialStreamed(pin_in.numpy())
    tw.Propagate()
    tw.SetOutputMirror(None)
    tw2 = ng.TWave(4, ts, 1.0)
    tw2.SetBoundaryCF(dirichlet(f, 3))
    tw2.SetInitial(0.5 * wf0)
    tw2.Propagate()
    assert np.array_equal(pin_out.numpy(), tw2.GetWavefront())
