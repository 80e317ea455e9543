"""Pipelining of independent steps (ngst_twave_prefetch_input / use_prefetched / get_wavefront_async / wait_output): the next
step's input is copied in while the current step computes and the current result is copied out while the next step runs.
Each step must give bitwise what the blocking calls give for the same input."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, first_order, standing
from ngstrefftz_b200._capi import check, lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D,n,order,dt,fo", [(3, 6, 4, 0.1, False), (3, 5, 4, 0.1, True), (2, 10, 4, 0.15, False), (1, 40, 6, 0.2, False)])
def test_pipelined_steps_match_blocking_calls(built, D, n, order, dt, fo):
    import torch
    f = standing(D)
    init = first_order(f) if fo else f
    mesh = {1: lambda: ng.SegMesh(n, 0, 1), 2: lambda: ng.UnitSquareMesh(n), 3: lambda: ng.UnitCubeMesh(n)}[D]()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    tw.SetBoundaryCF(dirichlet(f, D))
    wf0 = tw.MakeWavefront(init, 0.0)
    scales = [1.0, 0.5, -2.0, 0.25]
    bvals = tw.sample_boundary(0.0)
    # blocking reference: every step restarts from its own input at time 0
    want = []
    for s in scales:
        tw.reset_time()
        tw.SetInitial(s * wf0)
        tw.propagate_with_values(None if bvals is None else s * bvals)
        want.append(tw.GetWavefront())
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()  # noqa: E731
    ins = [pin(s * wf0) for s in scales]
    bnds = [None if bvals is None else pin(s * bvals) for s in scales]
    outs = [torch.zeros(wf0.shape, dtype=torch.float64).pin_memory() for _ in scales]
    tw.PrefetchInput(ins[0].numpy(), None if bnds[0] is None else bnds[0].numpy())
    for k in range(len(scales)):
        tw.reset_time()
        tw.UsePrefetched()
        if k + 1 < len(scales):
            tw.PrefetchInput(ins[k + 1].numpy(), None if bnds[k + 1] is None else bnds[k + 1].numpy())
        check(lib.ngst_twave_propagate(tw._h))  # (TWaveTents.Propagate would re-sample and upload the boundary datum)
        tw.GetWavefrontAsync(outs[k].numpy())
    tw.WaitOutput()
    for k in range(len(scales)):
        assert np.array_equal(outs[k].numpy(), want[k]), f"step {k}"
