"""ngst_twave_energy_resident_noalloc (include/ngst_b200.h): Energy of the resident state for callers that must not allocate
(no counterpart in the reference). Scratch and result slot are a file-scope
`static __device__` array in csrc/cuda/reduce.cu; the call does two launches, one 8-byte copy and a synchronisation.

Checked against the existing entry point ngst_twave_energy (scratch from cudaMalloc, partial sums added on the host) on two
calls in a row and across solvers of different size that share the static array (stale entries behind the shorter
solver's range must not be read). Tolerance 1e-14 relative, as the other Energy comparisons in tests/; the summation order
is the same on both sides, so the bits are expected to agree as well (printed).

Named test_zz_* so that pytest collects it last: a failure here hides no other test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from ngstrefftz_b200._capi import NgstError, check, lib
from _cases import dirichlet, first_order, standing

pytestmark = pytest.mark.gpu
TOL = 1e-14


def energy_existing(tw):
    out = C.c_double(0)
    check(lib.ngst_twave_energy(tw._h, None, C.byref(out)))
    return out.value


def energy_noalloc(tw):
    out = C.c_double(float("nan"))
    check(lib.ngst_twave_energy_resident_noalloc(tw._h, C.byref(out)))
    return out.value


def make(mesh, order, dt, init, bnd):
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, 1.0)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    return tw


def compare(tw, what):
    ref = energy_existing(tw)
    got = [energy_noalloc(tw), energy_noalloc(tw)]   # two calls in a row on the shared array
    for k, g in enumerate(got):
        print(f"{what}: call {k}: no-alloc {g!r} existing {ref!r} bitwise {g == ref}")
        assert np.isfinite(g) and ref > 0
        assert abs(g - ref) <= TOL * abs(ref), (what, k, g, ref)
    assert got[0] == got[1]   # the same state twice: deterministic
    assert energy_existing(tw) == ref   # the existing path is untouched by the calls in between


def test_noalloc_energy_matches_the_existing_entry_point(built):
    f3, f2 = standing(3), standing(2)
    big = make(ng.UnitCubeMesh(6), 4, 0.1, f3, dirichlet(f3, 3))               # 1 296 tets, second-order state
    small = make(ng.UnitSquareMesh(5), 3, 0.2, first_order(f2), dirichlet(f2, 2))   # 50 triangles, first-order state
    assert lib.ngst_twave_energy_noalloc_max_elements() >= 131712   # the bench slab fits
    compare(big, "3-D initial")
    compare(small, "2-D initial (behind 1 296 stale entries)")
    for step in range(2):
        big.Propagate()
        small.Propagate()
        compare(big, f"3-D after slab {step + 1}")
        compare(small, f"2-D after slab {step + 1}")
    # free device memory does not move across calls (informational: the code path has no allocation to begin with)
    try:
        import torch
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(10):
            energy_noalloc(big)
        print("free device memory before / after 10 calls:", free0, torch.cuda.mem_get_info()[0])
    except Exception as e:  # noqa: BLE001
        print("mem_get_info unavailable:", e)


def test_noalloc_energy_rejects_what_it_cannot_serve(built):
    limit = lib.ngst_twave_energy_noalloc_max_elements()
    n = int(np.ceil(np.sqrt((limit + 1) / 2)))
    mesh = ng.UnitSquareMesh(n)   # 2 n^2 triangles > limit
    assert mesh.ne > limit
    f = standing(2)
    tw = make(mesh, 2, 0.2 / n, f, dirichlet(f, 2))
    out = C.c_double(0)
    rc = lib.ngst_twave_energy_resident_noalloc(tw._h, C.byref(out))
    assert rc == -1   # NGST_ERR_ARG
    msg = lib.ngst_last_error().decode()
    print(msg)
    assert "exceed" in msg and str(limit) in msg
    assert energy_existing(tw) > 0   # the allocating entry point serves it
    # call order: no initial data -> NGST_ERR_STATE, like ngst_twave_energy
    ts = ng.TentSlab(ng.UnitSquareMesh(3), "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.2, True, 2 / 3)
    tw0 = ng.TWave(2, ts, 1.0)
    assert lib.ngst_twave_energy_resident_noalloc(tw0._h, C.byref(out)) == -2
    assert lib.ngst_twave_energy_resident_noalloc(tw0._h, None) == -1
    with pytest.raises(NgstError):
        tw0.EnergyNoAlloc()
