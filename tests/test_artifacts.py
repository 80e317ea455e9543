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
