"""Exploratory GPU-vs-oracle parity sweep (development aid; the judged tests are tests/test_gpu_parity.py)."""
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ngstrefftz_b200 as ng  # noqa: E402
from _cases import dirichlet, first_order, gaussian_1d, neumann, poly_solution, standing  # noqa: E402
from _oracle import Oracle  # noqa: E402


def run(name, mesh, order, c, dt, init, bnd, nslabs=1, neu=False):
    if neu:
        mesh.SetBoundaryNames("neumann")
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(c)
    assert ts.PitchTents(dt, True, 2 / 3)
    tw = ng.TWave(order, ts, c)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    o = Oracle(mesh)
    cm = c(o.centroids) if callable(c) else c
    assert o.pitch(dt, True, 2 / 3, cm)
    e1, e2 = ts.export(), o.export()
    same = all(np.array_equal(e1[k], e2[k]) for k in e2)
    o.setup(order, c)
    o.set_initial(init)
    o.set_boundary_cf(bnd)
    ms = []
    for _ in range(nslabs):
        tw.Propagate()
        ms.append(tw.last_ms())
        t0 = time.time()
        o.propagate()
        tcpu = time.time() - t0
    wf, ref = tw.GetWavefront(), o.wavefront()
    rel = np.abs(wf - ref).max() / np.abs(ref).max()
    en_g, en_o = (tw.Energy(wf), o.energy(ref)) if tw.fosystem == 0 else (0.0, 0.0)
    print(f"{name:28s} tents={ts.GetNTents():6d} same_tents={same} N={tw.nbasis:3d} Q={tw.nip:4d} rel={rel:.2e} "
          f"energy g/o={en_g:.12e}/{en_o:.12e} gpu={ms[-1]:.3f}ms cpu={tcpu*1e3:.1f}ms", flush=True)
    return rel


def main():
    pc = lambda x: np.where((np.floor(x[:, 0] * 4).astype(int) % 2) == 0, 1.0, 2.0)  # noqa: E731
    f1, f2, f3 = standing(1), standing(2), standing(3)
    run("1D p4 dirichlet", ng.SegMesh(16, 0, 1), 4, 1.0, 0.5, f1, dirichlet(f1, 1))
    run("1D p4 neumann", ng.SegMesh(16, 0, 1), 4, 1.0, 0.5, f1, neumann(f1, 1), neu=True)
    run("1D p6 piecewise c (macro)", ng.SegMesh(32, 0, 1), 6, pc, 0.3, gaussian_1d, lambda x, t: 0 * t, nslabs=2)
    run("1D p6 fo", ng.SegMesh(16, 0, 1), 6, 1.0, 0.5, first_order(f1), dirichlet(f1, 1))
    run("2D p4 dirichlet", ng.UnitSquareMesh(4), 4, 1.0, 0.5, f2, dirichlet(f2, 2))
    run("2D p4 neumann", ng.UnitSquareMesh(4), 4, 1.0, 0.5, f2, neumann(f2, 2), neu=True)
    run("2D p4 fo 2 slabs", ng.UnitSquareMesh(6), 4, 1.0, 0.3, first_order(f2), dirichlet(f2, 2), nslabs=2)
    pc2 = lambda x: np.where(x[:, 0] < 0.5, 1.0, 1.5)  # noqa: E731
    run("2D p3 piecewise c (macro)", ng.UnitSquareMesh(4), 3, pc2, 0.2, f2, dirichlet(f2, 2))
    run("2D p5 poly exact", ng.UnitSquareMesh(3), 5, 1.0, 0.3, poly_solution(2), dirichlet(poly_solution(2), 2))
    run("3D p3 dirichlet", ng.UnitCubeMesh(3), 3, 1.0, 0.2, f3, dirichlet(f3, 3))
    run("3D p4 dirichlet", ng.UnitCubeMesh(4), 4, 1.0, 0.2, f3, dirichlet(f3, 3))
    run("3D p4 neumann", ng.UnitCubeMesh(3), 4, 1.0, 0.2, f3, neumann(f3, 3), neu=True)
    pc3 = lambda x: np.where(x[:, 2] < 0.5, 1.0, 2.0)  # noqa: E731
    run("3D p2 piecewise c (macro)", ng.UnitCubeMesh(2), 2, pc3, 0.2, f3, dirichlet(f3, 3))
    run("3D p4 n=10 timing", ng.UnitCubeMesh(10), 4, 1.0, 0.1, f3, dirichlet(f3, 3), nslabs=2)


if __name__ == "__main__":
    main()
