"""ngst_slab_import_tents / TentSlab.from_reference: a slab built from the Tent fields a real ngstents session exposes
(python_tents.cpp:119-128) instead of the product's own pitcher. Host side (no GPU)."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from ngstrefftz_b200._capi import NgstError

MESHES = {1: (lambda: ng.SegMesh(16, 0, 1), 0.5), 2: (lambda: ng.UnitSquareMesh(5), 0.3), 3: (lambda: ng.UnitCubeMesh(3), 0.2)}


def pitched(D):
    mk, dt = MESHES[D]
    mesh = mk()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    return mesh, ts, dt


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("drop", [(), ("level",), ("level", "facet_off", "facets")])
def test_round_trip_is_bitwise(built, D, drop, tmp_path):
    """save_tents -> from_reference reproduces every table, the derived ones (dag, maxslope, level sets) included; levels and
    internal facets can be left out and are then replayed / taken from the mesh."""
    mesh, ts, dt = pitched(D)
    e = ts.export()
    path = str(tmp_path / "tents.npz")
    np.savez(path, **{k: v for k, v in e.items() if k not in drop})
    t2 = ng.TentSlab.from_reference(mesh, path, dt)
    e2 = t2.export()
    assert [k for k in e if not np.array_equal(e[k], e2[k])] == []
    assert t2.GetNTents() == ts.GetNTents() and t2.GetNLayers() == ts.GetNLayers()
    assert t2.GetSlabHeight() == dt and t2.MaxSlope() == ts.MaxSlope()
    assert t2.compare_tents(path) == []


def reorder_by_level(e):
    """The same tents renumbered level by level (a different but equally valid serial order)."""
    perm = np.argsort(e["level"], kind="stable")
    out = {k: e[k][perm] for k in ("vertex", "level", "tbot", "ttop")}
    for off, keys in (("nb_off", ("nbv", "nbtime")), ("el_off", ("els",)), ("facet_off", ("facets",))):
        lens = np.diff(e[off])[perm]
        out[off] = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        for k in keys:
            out[k] = np.concatenate([e[k][e[off][t]:e[off][t + 1]] for t in perm])
    return out, perm


@pytest.mark.parametrize("D", [2, 3])
def test_foreign_tent_numbering_is_accepted(built, D):
    mesh, ts, dt = pitched(D)
    e = ts.export()
    r, perm = reorder_by_level(e)
    t2 = ng.TentSlab.from_reference(mesh, r, dt)
    e2 = t2.export()
    assert np.array_equal(e2["vertex"], e["vertex"][perm]) and np.array_equal(e2["level"], e["level"][perm])
    assert np.array_equal(np.bincount(e2["level"]), np.bincount(e["level"]))
    assert np.array_equal(e2["maxslope"], e["maxslope"][perm])
    # dependencies: every successor has a higher level
    for t in range(len(e2["vertex"])):
        succ = e2["dag"][e2["dag_off"][t]:e2["dag_off"][t + 1]]
        assert np.all(e2["level"][succ] > e2["level"][t])


def test_malformed_lists_are_refused(built):
    mesh, ts, dt = pitched(2)
    e = {k: v.copy() for k, v in ts.export().items()}

    def expect_error(change):
        d = {k: v.copy() for k, v in e.items()}
        change(d)
        with pytest.raises((NgstError, ValueError)):
            ng.TentSlab.from_reference(mesh, d, dt)

    expect_error(lambda d: d["vertex"].__setitem__(0, mesh.nv))            # vertex out of range
    expect_error(lambda d: d["els"].__setitem__(0, mesh.ne))               # element out of range
    expect_error(lambda d: d["ttop"].__setitem__(3, d["tbot"][3]))         # empty tent
    expect_error(lambda d: d["level"].__setitem__(slice(None), 0))         # tents sharing an element on one level
    expect_error(lambda d: d.__setitem__("nbv", d["nbv"][:-1]))            # lengths do not fit
    far = int(np.setdiff1d(np.arange(mesh.ne), e["els"][e["el_off"][0]:e["el_off"][1]])[-1])
    expect_error(lambda d: d["els"].__setitem__(0, far))                   # element without the tent vertex
    # a failed import leaves the slab usable: the original tents are still there
    assert ts.GetNTents() == len(e["vertex"])


def test_levels_are_replayed_like_the_pitcher(built):
    mesh, ts, dt = pitched(3)
    e = ts.export()
    d = {k: v for k, v in e.items() if k != "level"}
    assert np.array_equal(ng.TentSlab.from_reference(mesh, d, dt).export()["level"], e["level"])
