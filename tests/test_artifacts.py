"""Headless exporters and the exchange formats for real-reference artefacts. No GPU."""
import os
import types

import numpy as np

import ngstrefftz_b200 as ng


def pitched(n=3, dt=0.3):
    m = ng.UnitSquareMesh(n)
    ts = ng.TentSlab(m)
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    return m, ts


def test_vtk_export_matches_reference_layout(built, tmp_path):
    """DrawPitchedTentsVTK (ngstents tents.cpp:982-1058): 2 pole points + 2 per element, one tetrahedron per tent-element."""
    m, ts = pitched()
    npts, ncells = ts.DrawPitchedTentsVTK(str(tmp_path / "tents"))
    e = ts.export()
    assert ncells == len(e["els"]) and npts == 2 * ts.GetNTents() + 2 * len(e["els"])
    lines = open(tmp_path / "tents.vtk").read().splitlines()
    assert lines[:4] == ["# vtk DataFile Version 3.0", "vtk output", "ASCII", "DATASET UNSTRUCTURED_GRID"]
    assert lines[4] == f"POINTS {npts} float"
    assert lines[5 + npts] == f"CELLS {ncells} {5 * ncells}"
    first = [float(x) for x in lines[5].split()]
    v = e["vertex"][0]
    assert np.allclose(first, [m.points[v, 0], m.points[v, 1], e["tbot"][0]])
    cell0 = [int(x) for x in lines[6 + npts].split()]
    assert cell0[0] == 4 and cell0[1:3] == [0, 1] and cell0[3:] == [2, 3]
    assert f"tentlevel 1 {npts} float" in lines and f"tentnumber 1 {npts} float" in lines


def test_tent_dump_round_trip_and_reference_dump(built, tmp_path):
    m, ts = pitched(4, 0.2)
    path = str(tmp_path / "tents.npz")
    ts.save_tents(path)
    assert ts.compare_tents(path) == [] and ts.compare_tents(path, same_numbering=False) == []
    # a stand-in for ngstents.TentSlab exposing the same GetNTents / GetTent(i) attributes (python_tents.cpp:119-128)
    fake = types.SimpleNamespace(GetNTents=ts.GetNTents, GetTent=ts.GetTent)
    ref = str(tmp_path / "ref.npz")
    ng.TentSlab.dump_reference_tents(fake, ref)
    assert ts.compare_tents(ref) == []
    # a different slab height must be detected, with and without a common numbering
    _, other = pitched(4, 0.25)
    assert other.compare_tents(ref) != [] and other.compare_tents(ref, same_numbering=False) != []


def test_mesh1dtents_space_time_mesh(built):
    """Mesh1dTents (src/mesh1dtents.cpp:46-174): one triangle / quadrilateral per tent, unique points, the elements tile
    [x0, x1] x [0, dt] exactly, labelled boundary segments cover the four sides."""
    m = ng.SegMesh(8, 0.0, 2.0)
    ts = ng.TentSlab(m)
    ts.SetMaxWavespeed(1.0)
    dt = 0.3
    assert ts.PitchTents(dt, True, 2 / 3)
    out = ng.Mesh1dTents(ts)
    P = out["points"]
    assert len(np.unique(P, axis=0)) == len(P)
    assert len(out["trigs"]) + len(out["quads"]) == ts.GetNTents()

    def area(poly):
        xy = P[poly]
        return 0.5 * abs(np.sum(xy[:, 0] * np.roll(xy[:, 1], -1) - np.roll(xy[:, 0], -1) * xy[:, 1]))

    total = sum(area(t) for t in out["trigs"]) + sum(area(q) for q in out["quads"])
    assert abs(total - 2.0 * dt) < 1e-12
    seg, bc = out["segments"], out["segment_bc"]

    def length(mask):
        d = P[seg[mask, 0]] - P[seg[mask, 1]]
        return np.sqrt((d ** 2).sum(axis=1)).sum()

    assert abs(length(bc == 1) - 2.0) < 1e-12 and abs(length(bc == 2) - 2.0) < 1e-12   # inflow t = 0, outflow t = dt
    assert np.all(P[seg[bc == 1]][:, :, 1] < 1e-5) and np.all(P[seg[bc == 2]][:, :, 1] > dt - 1e-5)
    db = seg[bc == 3]
    assert np.all(np.isin(P[db][:, :, 0], [0.0, 2.0]))
