"""Tent pitching: the product's O(log n) pitcher is bit-exact with the literal restatement of tents.cpp:78-359 in
oracle/oracle_pitch.cpp, and satisfies the reference's structural tests
(ngstents/tests/test_causal_tents.py, test_tentlayers.py, test_tent_height_2D.py)  [K4]."""
import math

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _oracle import Oracle


def wavespeed_cases(D):
    return {
        "const": 1.0,
        "const3": 3.0,
        "piecewise": lambda x: np.where(x[:, 0] < 0.5, 1.0, 2.0),
        "smooth": lambda x: 1.0 + x.sum(axis=1),
    }


def meshes():
    return {
        "seg16": lambda: ng.SegMesh(16, 0, 1), "segpi4": lambda: ng.SegMesh(4, 0, math.pi),
        "sq4": lambda: ng.UnitSquareMesh(4), "sq7": lambda: ng.UnitSquareMesh(7),
        "rect": lambda: ng.StructuredMesh([5, 3], [0, 0], [2.0, 0.7]),
        "cube3": lambda: ng.UnitCubeMesh(3), "box": lambda: ng.StructuredMesh([4, 2, 3], [0, 0, 0], [1, 0.6, 1.3]),
    }


def pitch_both(mesh, c, dt, local_ct, global_ct, method="edge"):
    ts = ng.TentSlab(mesh, method)
    ts.SetMaxWavespeed(c)
    ok = ts.PitchTents(dt, local_ct, global_ct)
    o = Oracle(mesh)
    cm = c(o.centroids) if callable(c) else c
    ok2 = o.pitch(dt, local_ct, global_ct, cm, method)
    return ts, o, ok, ok2


@pytest.mark.parametrize("mname", list(meshes()))
@pytest.mark.parametrize("cname", ["const", "piecewise", "smooth"])
@pytest.mark.parametrize("local_ct,global_ct", [(True, 2 / 3), (False, 1.0), (False, 0.5)])
def test_bit_exact_schedule(built, mname, cname, local_ct, global_ct):
    mesh = meshes()[mname]()
    c = wavespeed_cases(mesh.dim)[cname]
    ts, o, ok, ok2 = pitch_both(mesh, c, 0.3, local_ct, global_ct)
    assert ok == ok2
    e1, e2 = ts.export(), o.export()
    assert ts.GetNTents() == o.ntents() and ts.GetNLayers() == o.nlayers()
    for k in e2:
        assert np.array_equal(e1[k], e2[k]), k  # ints identical, doubles bitwise identical


@pytest.mark.parametrize("mname", ["seg16", "sq4", "cube3"])
def test_vol_method_bit_exact(built, mname):
    mesh = meshes()[mname]()
    ts, o, ok, ok2 = pitch_both(mesh, 1.0, 0.2, False, 1.0, method="vol")
    assert ok == ok2
    e1, e2 = ts.export(), o.export()
    for k in e2:
        assert np.array_equal(e1[k], e2[k]), k


@pytest.mark.parametrize("mname", list(meshes()))
@pytest.mark.parametrize("c", [1.0, 3.0])
def test_structural_properties(built, mname, c):
    mesh = meshes()[mname]()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(0.25, True, 2 / 3)
    e = ts.export()
    nt = ts.GetNTents()
    # test_causal_tents.py: slope <= 1/c ; test_tent_height_2D.py: per tent, per edge slope
    assert ts.MaxSlope() <= 1.0 / c + 1e-12
    pts = mesh.points
    for i in range(nt):
        t = ts.GetTent(i)
        for nb, tnb in zip(t.nbv, t.nbtime):
            L = np.linalg.norm(pts[nb] - pts[t.vertex])
            assert abs(t.ttop - tnb) / L <= 1.0 / c + 1e-12
        assert t.ttop > t.tbot
    # test_tentlayers.py: number of distinct levels == GetNLayers()
    assert len(set(e["level"].tolist())) == ts.GetNLayers()
    assert len(e["level_off"]) - 1 == ts.GetNLayers()
    # levels form a valid BSP schedule: disjoint element sets inside a level, DAG edges go strictly upwards
    for l in range(ts.GetNLayers()):
        seen = set()
        for tid in e["level_tents"][e["level_off"][l]:e["level_off"][l + 1]]:
            assert e["level"][tid] == l
            els = e["els"][e["el_off"][tid]:e["el_off"][tid + 1]].tolist()
            assert seen.isdisjoint(els)
            seen.update(els)
    for i in range(nt):
        for j in e["dag"][e["dag_off"][i]:e["dag_off"][i + 1]]:
            assert e["level"][j] > e["level"][i]
    # the front reaches dt everywhere: last tent over each vertex tops out at dt
    top = np.zeros(mesh.nv)
    for i in range(nt):
        top[e["vertex"][i]] = e["ttop"][i]
    assert np.all(top == 0.25)
    assert ts.GetSlabHeight() == 0.25


def test_tent_fields_and_tentdata1d(built):
    mesh = ng.SegMesh(8, 0, 1)
    ts = ng.TentSlab(mesh)
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.2, True, 2 / 3)
    td = ts.TentData1D()
    assert len(td) == ts.GetNTents()
    t0 = ts.GetTent(0)
    assert td[0][0] == (t0.vertex, t0.ttop, t0.tbot, t0.level)
    assert len(t0.nbv) == len(t0.nbtime) and len(t0.els) in (1, 2)
    with pytest.raises(IndexError):
        ts.GetTent(ts.GetNTents())


# ---------------------------------------------------------------------------------------------- level-synchronous pitcher
@pytest.mark.parametrize("D,mk,dt", [(1, lambda: ng.SegMesh(24, 0, 1), 0.5), (2, lambda: ng.UnitSquareMesh(7), 0.3),
                                      (3, lambda: ng.UnitCubeMesh(4), 0.2), (3, lambda: ng.StructuredMesh([4, 2, 3], [0, 0, 0], [1, 0.6, 1.3]), 0.15)])
@pytest.mark.parametrize("local_ct", [True, False])
def test_level_sync_pitcher_equals_serial_loop_with_the_same_rule(built, D, mk, dt, local_ct):
    """ngst_slab_pitch_level_sync picks the ready vertex of minimal (level, vertex number) and handles the ready vertices of
    a level as one parallel phase; the oracle runs the SERIAL greedy loop with that pick rule. Every table must be bitwise
    equal; the slab must be complete and causal; the level sets must be conflict-free."""
    mesh = mk()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, local_ct, 2 / 3, level_sync=True)
    o = Oracle(mesh)
    assert o.pitch(dt, local_ct, 2 / 3, 1.0, rule=1)
    e, eo = ts.export(), o.export()
    assert [k for k in eo if not np.array_equal(e[k], eo[k])] == []
    assert ts.GetNLayers() == o.nlayers()
    # complete: the last tent of every vertex reaches the slab height; causal: slopes below 1 / c
    top = np.zeros(mesh.nv)
    np.maximum.at(top, e["vertex"], e["ttop"])
    assert np.all(top == dt)
    if local_ct:  # (without the local constant the edge algorithm does not guarantee causality in 3-D, whatever the pick rule)
        assert ts.MaxSlope() <= 1.0 + 1e-12
    # tents of one level share no element
    for l in range(int(e["level"].max()) + 1):
        els = np.concatenate([e["els"][e["el_off"][t]:e["el_off"][t + 1]] for t in np.flatnonzero(e["level"] == l)])
        assert len(els) == len(np.unique(els))
    # other tents than the reference's pick rule (that is the point), not more of them on these meshes
    ref = ng.TentSlab(mesh, "edge")
    ref.SetMaxWavespeed(1.0)
    assert ref.PitchTents(dt, local_ct, 2 / 3)
    assert ts.GetNLayers() <= ref.GetNLayers()
