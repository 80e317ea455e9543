"""B200-native tent-pitched Trefftz-DG wave propagator with the Python surface of NGSTrefftz' tent path.

Mirrors (reference file:line, relative to the NGSTrefftz tree):
  TentSlab / Tent        external_dependencies/ngstents/src/python_tents.cpp:7-131
  TWave, TrefftzTents,   src/twavetents.cpp:1157-1229
  TWaveTents1/2/3
  GetWave                src/__init__.py:7-53

NGSolve is not available, so `CoefficientFunction`s become plain Python callables / arrays:
  wavespeed : float | array(nel) | callable(x[n,D]) -> array(n)        (evaluated at element centroids)
  initial   : callable(x[n,D], t) -> array(n, ncomp)  | array(nel, ncomp*nip)
  boundary  : callable(x[n,D], t[n]) -> array(n) (Dirichlet: dt u) or array(n, D) (Neumann: grad u)
  initial / boundary may also be an AnalyticData object (StandingWave, PlanePulse, PolynomialWave): those are sampled
  on the device (cuda/sample.cu), with no host evaluation and no H2D copy of samples
and netgen meshes become `Mesh` objects (structured generators SegMesh / UnitSquareMesh / UnitCubeMesh,
or explicit point/element arrays). All arithmetic of Propagate runs in libngst_b200.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import NgstError, check, dptr, f64, i32, iptr, lib
from ._poly import PolynomialCF
from ._analytic import AnalyticData, PlanePulse, PolynomialWave, StandingWave

__all__ = ["Mesh", "SegMesh", "UnitSquareMesh", "UnitCubeMesh", "StructuredMesh", "TentSlab", "Tent", "TWave",
           "TrefftzTents", "TWaveTents", "TaskManager", "SetNumThreads", "NgstError", "PolynomialCF",
           "AnalyticData", "StandingWave", "PlanePulse", "PolynomialWave",
           "GetIntegrationRule", "SetIntegrationRule"]


class TaskManager:
    """No-op stand-in for `with TaskManager():` (required by the reference, paralleldepend.hpp:70)."""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def SetNumThreads(n):
    """Accepted for script compatibility; tents are scheduled on the GPU."""


class Mesh:
    """Simplex mesh (replaces ngsolve.Mesh for this path)."""

    def __init__(self, points=None, elements=None, boundary=None, boundary_names=None, _handle=None):
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
        else:
            pts = f64(points)
            if pts.ndim == 1:
                pts = pts.reshape(-1, 1)
            el = i32(elements)
            D = pts.shape[1]
            bnd = i32(boundary if boundary is not None else np.zeros((0, D)))
            if bnd.ndim == 1:
                bnd = bnd.reshape(-1, 1)
            kind = self._kinds(boundary_names, bnd.shape[0])
            check(lib.ngst_mesh_create(D, pts.shape[0], dptr(pts), el.shape[0], iptr(el), bnd.shape[0], iptr(bnd),
                                       iptr(kind), C.byref(self._h)))
        cnt = np.zeros(8, dtype=np.int32)
        check(lib.ngst_mesh_counts(self._h, iptr(cnt)))
        self.dim, self.nv, self.ne, self.nse, self.nedge, self.nfacet = (int(x) for x in cnt[:6])
        self._tables = None

    @staticmethod
    def _kinds(names, n):
        if names is None:
            return np.zeros(n, dtype=np.int32)
        if isinstance(names, str):
            names = [names] * n
        return i32([1 if (nm == "neumann" or nm == 1) else 0 for nm in names])

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            lib.ngst_mesh_destroy(self._h)
            self._h = C.c_void_p()

    def SetBoundaryNames(self, names):
        """Boundary named "neumann" selects the Neumann flux (src/twavetents.cpp:380); anything else Dirichlet."""
        kind = self._kinds(names, self.nse)
        check(lib.ngst_mesh_set_boundary_kind(self._h, iptr(kind)))
        self._tables = None

    def tables(self):
        if self._tables is None:
            D = self.dim
            nepe = {1: 1, 2: 3, 3: 6}[D]
            t = dict(pts=np.zeros((self.nv, D)), el2v=np.zeros((self.ne, D + 1), np.int32),
                     bnd2v=np.zeros((self.nse, D), np.int32), bnd_kind=np.zeros(self.nse, np.int32),
                     edge2v=np.zeros((self.nedge, 2), np.int32), facet2v=np.zeros((self.nfacet, D), np.int32),
                     el2facet=np.zeros((self.ne, D + 1), np.int32), facet2el=np.zeros((self.nfacet, 2), np.int32),
                     facet2bnd=np.zeros(self.nfacet, np.int32), el2edge=np.zeros((self.ne, nepe), np.int32))
            check(lib.ngst_mesh_tables(self._h, dptr(t["pts"]), iptr(t["el2v"]), iptr(t["bnd2v"]), iptr(t["bnd_kind"]),
                                       iptr(t["edge2v"]), iptr(t["facet2v"]), iptr(t["el2facet"]), iptr(t["facet2el"]),
                                       iptr(t["facet2bnd"]), iptr(t["el2edge"])))
            self._tables = t
        return self._tables

    @property
    def points(self):
        return self.tables()["pts"]

    @property
    def elements(self):
        return self.tables()["el2v"]

    def centroids(self):
        t = self.tables()
        return t["pts"][t["el2v"]].mean(axis=1)


def StructuredMesh(ncells, lo, hi):
    n = i32(np.atleast_1d(ncells))
    lo = f64(np.atleast_1d(lo))
    hi = f64(np.atleast_1d(hi))
    h = C.c_void_p()
    check(lib.ngst_mesh_structured(len(n), iptr(n), dptr(lo), dptr(hi), C.byref(h)))
    return Mesh(_handle=h)


def SegMesh(n, x0, x1):
    """Uniform 1-D mesh, as test/tents.py:38 `SegMesh(4,0,math.pi)`."""
    return StructuredMesh([n], [x0], [x1])


def UnitSquareMesh(n):
    return StructuredMesh([n, n], [0.0, 0.0], [1.0, 1.0])


def UnitCubeMesh(n):
    return StructuredMesh([n, n, n], [0.0, 0.0, 0.0], [1.0, 1.0, 1.0])


class Tent:
    """struct Tent fields exported at python_tents.cpp:119-128."""

    def __init__(self, vertex, ttop, tbot, nbv, nbtime, els, level, internal_facets, maxslope):
        self.vertex, self.ttop, self.tbot = int(vertex), float(ttop), float(tbot)
        self.nbv, self.nbtime, self.els = nbv, nbtime, els
        self.level, self.internal_facets, self._maxslope = int(level), internal_facets, float(maxslope)

    def MaxSlope(self):
        return self._maxslope


class TentSlab:
    """TentSlab(mesh, method="edge", heapsize=1000000)  (python_tents.cpp:13-29)."""

    def __init__(self, mesh, method="edge", heapsize=1000000):
        if method not in ("edge", "vol"):
            print("Invalid method! Setting edge algorithm as default...")
            method = "edge"
        self.mesh = mesh
        self._h = C.c_void_p()
        check(lib.ngst_slab_create(mesh._h, 2 if method == "edge" else 1, C.byref(self._h)))
        self._export = None

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            lib.ngst_slab_destroy(self._h)
            self._h = C.c_void_p()

    def SetMaxWavespeed(self, wavespeed):
        if callable(wavespeed):
            c = f64(wavespeed(self.mesh.centroids())).reshape(-1)
        else:
            c = f64(np.atleast_1d(wavespeed)).reshape(-1)
        check(lib.ngst_slab_set_max_wavespeed(self._h, dptr(c), len(c)))

    def PitchTents(self, dt, local_ct=False, global_ct=1.0, level_sync=False):
        """python_tents.cpp:44-65. level_sync=True (extension): the level-synchronous parallel pitcher with the pick rule
        "minimal (level, vertex number)" -- other tents than the reference's order (ngst_slab_pitch_level_sync)."""
        ok = C.c_int(0)
        self._export = None
        try:
            fn = lib.ngst_slab_pitch_level_sync if level_sync else lib.ngst_slab_pitch
            check(fn(self._h, float(dt), int(bool(local_ct)), float(global_ct), C.byref(ok)))
        except NgstError as e:
            if e.code == -2:
                raise RuntimeError("Wavespeed has not been set!") from e
            raise
        return bool(ok.value)

    def GetNTents(self):
        return lib.ngst_slab_ntents(self._h)

    def GetNLayers(self):
        return lib.ngst_slab_nlayers(self._h)

    def GetSlabHeight(self):
        return lib.ngst_slab_height(self._h)

    def MaxSlope(self):
        return lib.ngst_slab_max_slope(self._h)

    def export(self):
        """Flattened tents as a dict of numpy arrays (CSR offsets *_off)."""
        if self._export is None:
            nt = self.GetNTents()
            sz = np.zeros(4, dtype=np.int64)
            check(lib.ngst_slab_sizes(self._h, sz.ctypes.data_as(_capi.c_llp)))
            e = dict(vertex=np.zeros(nt, np.int32), level=np.zeros(nt, np.int32), tbot=np.zeros(nt), ttop=np.zeros(nt),
                     maxslope=np.zeros(nt), nb_off=np.zeros(nt + 1, np.int32), nbv=np.zeros(sz[0], np.int32),
                     nbtime=np.zeros(sz[0]), el_off=np.zeros(nt + 1, np.int32), els=np.zeros(sz[1], np.int32),
                     facet_off=np.zeros(nt + 1, np.int32), facets=np.zeros(sz[2], np.int32),
                     dag_off=np.zeros(nt + 1, np.int32), dag=np.zeros(sz[3], np.int32))
            check(lib.ngst_slab_export(self._h, iptr(e["vertex"]), iptr(e["level"]), dptr(e["tbot"]), dptr(e["ttop"]),
                                       dptr(e["maxslope"]), iptr(e["nb_off"]), iptr(e["nbv"]), dptr(e["nbtime"]),
                                       iptr(e["el_off"]), iptr(e["els"]), iptr(e["facet_off"]), iptr(e["facets"]),
                                       iptr(e["dag_off"]), iptr(e["dag"])))
            nl = C.c_int(0)
            check(lib.ngst_slab_levels(self._h, C.byref(nl), None, None))
            e["level_off"] = np.zeros(nl.value + 1, np.int32)
            e["level_tents"] = np.zeros(nt, np.int32)
            check(lib.ngst_slab_levels(self._h, C.byref(nl), iptr(e["level_off"]), iptr(e["level_tents"])))
            self._export = e
        return self._export

    def DrawPitchedTentsVTK(self, filename="output"):
        """Headless writer of TentPitchedSlab::DrawPitchedTentsVTK (ngstents/src/tents.cpp:982-1058): 2-D meshes, one
        tetrahedron per (tent, element), point data tentlevel / tentnumber. Writes <filename>.vtk (legacy ASCII)."""
        mesh = self.mesh
        if mesh.dim != 2:
            raise ValueError("DrawPitchedTentsVTK: 2-D meshes only (as the reference)")
        e = self.export()
        pts, el2v = mesh.points, mesh.elements
        P, cells, level, tentnr = [], [], [], []
        for i in range(len(e["vertex"])):
            v = int(e["vertex"][i])
            first = len(P)
            P.append((pts[v, 0], pts[v, 1], e["tbot"][i]))
            P.append((pts[v, 0], pts[v, 1], e["ttop"][i]))
            nb = {int(n): float(t) for n, t in zip(e["nbv"][e["nb_off"][i]:e["nb_off"][i + 1]],
                                                   e["nbtime"][e["nb_off"][i]:e["nb_off"][i + 1]])}
            for el in e["els"][e["el_off"][i]:e["el_off"][i + 1]]:
                tet = [first, first + 1]
                for w in el2v[el]:
                    if int(w) != v:
                        tet.append(len(P))
                        P.append((pts[w, 0], pts[w, 1], nb[int(w)]))
                cells.append(tet)
            n = len(P) - first
            level += [int(e["level"][i])] * n
            tentnr += [i] * n
        with open(filename + ".vtk", "w") as out:
            out.write("# vtk DataFile Version 3.0\nvtk output\nASCII\nDATASET UNSTRUCTURED_GRID\n")
            out.write(f"POINTS {len(P)} float\n")
            out.writelines(f"{x:.9g} {y:.9g} {t:.9g}\n" for x, y, t in P)
            out.write(f"CELLS {len(cells)} {5 * len(cells)}\n")
            out.writelines("4 " + " ".join(str(k) for k in c) + "\n" for c in cells)
            out.write(f"CELL_TYPES {len(cells)}\n" + "10 \n" * len(cells))
            out.write(f"CELL_DATA {len(cells)}\nPOINT_DATA {len(P)}\nFIELD FieldData 2\n")
            out.write(f"tentlevel 1 {len(level)} float\n" + " ".join(str(k) for k in level) + " \n")
            out.write(f"tentnumber 1 {len(tentnr)} float\n" + " ".join(str(k) for k in tentnr) + " \n")
        return len(P), len(cells)

    # ---- exchange of tent lists with a real NGSTrefftz / ngstents install
    def save_tents(self, path):
        """All flattened tent tables (export()) as one .npz; dump_reference_tents() writes the same keys from ngstents."""
        np.savez_compressed(path, **self.export())

    @staticmethod
    def dump_reference_tents(ref_slab, path):
        """Run inside a real ngstents session: ref_slab is ngstents.TentSlab after PitchTents. Flattens
        GetTent(i).{vertex, ttop, tbot, nbv, nbtime, els, level, internal_facets} (python_tents.cpp:119-128) to the .npz
        layout of save_tents (no dag / maxslope / level sets: they are derived)."""
        cols = {k: [] for k in ("vertex", "level", "tbot", "ttop", "nbv", "nbtime", "els", "facets")}
        off = {"nb_off": [0], "el_off": [0], "facet_off": [0]}
        for i in range(ref_slab.GetNTents()):
            t = ref_slab.GetTent(i)
            cols["vertex"].append(t.vertex)
            cols["level"].append(t.level)
            cols["tbot"].append(t.tbot)
            cols["ttop"].append(t.ttop)
            cols["nbv"] += list(t.nbv)
            cols["nbtime"] += list(t.nbtime)
            cols["els"] += list(t.els)
            cols["facets"] += list(t.internal_facets)
            off["nb_off"].append(len(cols["nbv"]))
            off["el_off"].append(len(cols["els"]))
            off["facet_off"].append(len(cols["facets"]))
        out = {k: np.asarray(v, dtype=np.float64 if k in ("tbot", "ttop", "nbtime") else np.int32) for k, v in cols.items()}
        out.update({k: np.asarray(v, dtype=np.int32) for k, v in off.items()})
        np.savez_compressed(path, **out)

    def import_tents(self, tents, dt=None):
        """Take the tents of another pitcher instead of pitching: `tents` is a dict (or .npz path) with the keys of
        save_tents / dump_reference_tents -- vertex, tbot, ttop, nb_off, nbv, nbtime, el_off, els and optionally level,
        facet_off, facets (ngst_slab_import_tents). dt defaults to the largest ttop."""
        if isinstance(tents, (str, bytes)) or hasattr(tents, "__fspath__"):
            with np.load(tents) as z:
                tents = {k: z[k] for k in z.files}
        g = lambda k, conv: conv(tents[k]) if k in tents and tents[k] is not None else None  # noqa: E731
        vertex, level = i32(tents["vertex"]), g("level", i32)
        tbot, ttop, nbtime = f64(tents["tbot"]), f64(tents["ttop"]), f64(tents["nbtime"])
        nb_off, nbv, el_off, els = i32(tents["nb_off"]), i32(tents["nbv"]), i32(tents["el_off"]), i32(tents["els"])
        facet_off, facets = g("facet_off", i32), g("facets", i32)
        n = len(vertex)
        if not (len(tbot) == len(ttop) == n and len(nb_off) == len(el_off) == n + 1 and len(nbv) == len(nbtime)
                and (level is None or len(level) == n) and (facet_off is None or len(facet_off) == n + 1)
                and (n == 0 or (len(nbv) == nb_off[-1] and len(els) == el_off[-1]))
                and (facet_off is None or facets is not None and len(facets) == facet_off[-1])):
            raise ValueError("import_tents: array lengths do not fit together")
        dt = float(ttop.max()) if dt is None else float(dt)
        self._export = None
        check(lib.ngst_slab_import_tents(self._h, n, dt, iptr(vertex), iptr(level), dptr(tbot), dptr(ttop), iptr(nb_off),
                                         iptr(nbv), dptr(nbtime), iptr(el_off), iptr(els), iptr(facet_off), iptr(facets)))
        return self

    @classmethod
    def from_reference(cls, mesh, tents, dt=None, method="edge"):
        """TentSlab over `mesh` holding the tents of a dump written by dump_reference_tents inside a real ngstents session
        (or by save_tents): the counterpart of handing the reference's TentPitchedSlab to TWave."""
        return cls(mesh, method).import_tents(tents, dt)

    def compare_tents(self, path, same_numbering=True):
        """Compare this slab with a tent dump (save_tents / dump_reference_tents). same_numbering=True: bitwise equality of
        every shared table (needs identical vertex / element / facet numbering). False: numbering-free invariants only --
        tent count, level histogram, sorted (tbot, ttop) per level -- for meshes numbered by another program.
        Returns the list of differing items (empty = match)."""
        ref = np.load(path)
        mine = self.export()
        diffs = []
        if same_numbering:
            return [k for k in ref.files if k in mine and not np.array_equal(ref[k], mine[k])]
        if len(ref["vertex"]) != len(mine["vertex"]):
            diffs.append("ntents")
        if not np.array_equal(np.bincount(ref["level"]), np.bincount(mine["level"])):
            diffs.append("level histogram")
        else:
            for name in ("tbot", "ttop"):
                a = np.lexsort((ref[name], ref["level"]))
                b = np.lexsort((mine[name], mine["level"]))
                if not np.allclose(ref[name][a], mine[name][b], rtol=1e-13, atol=1e-15):
                    diffs.append(name)
        return diffs

    def GetTent(self, i):
        e = self.export()
        if not 0 <= i < len(e["vertex"]):
            raise IndexError("tent index out of range")
        nb = slice(e["nb_off"][i], e["nb_off"][i + 1])
        el = slice(e["el_off"][i], e["el_off"][i + 1])
        fc = slice(e["facet_off"][i], e["facet_off"][i + 1])
        return Tent(e["vertex"][i], e["ttop"][i], e["tbot"][i], e["nbv"][nb].tolist(), e["nbtime"][nb].tolist(),
                    e["els"][el].tolist(), e["level"][i], e["facets"][fc].tolist(), e["maxslope"][i])

    def TentData1D(self):
        """python_tents.cpp:87-101"""
        ret = []
        for i in range(self.GetNTents()):
            t = self.GetTent(i)
            ret.append([(t.vertex, t.ttop, t.tbot, t.level)] + list(zip(t.nbv, t.nbtime)))
        return ret


def Mesh1dTents(tps):
    """Headless form of Mesh1dTents / NgsTP1dmesh (src/mesh1dtents.cpp:46-174): the 1+1-D tent slab as a 2-D space-time
    mesh -- one triangle (boundary tent) or quadrilateral per tent, unique points (x, t), boundary segments labelled
    "inflow" (t = 0), "outflow" (t = dt) and "dirichlet" (spatial boundary). The reference builds a netgen Mesh; here the
    same entities come back as arrays: dict(points[n,2], trigs[m,3], quads[k,4], segments[s,2], segment_bc[s] in
    {1: inflow, 2: outflow, 3: dirichlet}). Degenerate tents (ttop < tbot + 1e-5) are skipped as in the reference."""
    mesh = tps.mesh
    if mesh.dim != 1:
        raise ValueError("Mesh1dTents: 1-D meshes only")
    dt_eps = 0.00001
    dt = tps.GetSlabHeight()
    x = mesh.points[:, 0]
    bd = [float(x[v]) for v in np.unique(mesh.tables()["bnd2v"])] if "bnd2v" in mesh.tables() else [float(x.min()), float(x.max())]
    pim, pts, trigs, quads, segs, bcs = {}, [], [], [], [], []

    def add(px, pt):
        key = (px, pt)
        if key not in pim:
            pim[key] = len(pts)
            pts.append(key)
        return pim[key]

    e = tps.export()
    for i in range(len(e["vertex"])):
        v, tbot, ttop = int(e["vertex"][i]), float(e["tbot"][i]), float(e["ttop"][i])
        if ttop < tbot + dt_eps:
            continue
        nbv = e["nbv"][e["nb_off"][i]:e["nb_off"][i + 1]]
        nbt = e["nbtime"][e["nb_off"][i]:e["nb_off"][i + 1]]
        vert = [None] * (len(nbv) + 2)
        ibot, itop = (0, 2) if nbv[0] > v else (2, 0)
        vert[ibot] = add(float(x[v]), tbot)
        vert[itop] = add(float(x[v]), ttop)
        for k in range(len(nbv)):
            vert[2 * k + 1] = add(float(x[nbv[k]]), float(nbt[k]))
        (trigs if len(nbv) == 1 else quads).append(vert[:3] if len(nbv) == 1 else vert[:4])
        for k in range(len(nbv)):
            if tbot < dt_eps and nbt[k] < dt_eps:
                segs.append((vert[ibot], vert[2 * k + 1]) if v < nbv[k] else (vert[2 * k + 1], vert[ibot]))
                bcs.append(1)
            if ttop > dt - dt_eps and nbt[k] > dt - dt_eps:
                segs.append((vert[2 * k + 1], vert[itop]) if v < nbv[k] else (vert[itop], vert[2 * k + 1]))
                bcs.append(2)
            if float(x[v]) in bd:
                segs.append((vert[2], vert[0]))
                bcs.append(3)
    return {"points": np.array(pts, dtype=float).reshape(-1, 2), "trigs": np.array(trigs, dtype=np.int32).reshape(-1, 3),
            "quads": np.array(quads, dtype=np.int32).reshape(-1, 4), "segments": np.array(segs, dtype=np.int32).reshape(-1, 2),
            "segment_bc": np.array(bcs, dtype=np.int32)}


class L2Wave:
    """Result of GetWave: coef[el, component, k] in the basis psi_k = sum_j mono[k, j] xi^expo[j] (orthonormal on the
    reference simplex; xi = reference coordinates, x = sum_i lambda_i P_i with lambda = (xi, 1 - sum xi))."""

    def __init__(self, coef, order, psi_at_points, mono, expo):
        self.coef, self.order, self.psi_at_points, self.mono, self.expo = coef, order, psi_at_points, mono, expo

    def at_quadrature(self):
        """values[el, component, q] of the projected field at the rule points of every element"""
        return np.einsum("eck,qk->ecq", self.coef, self.psi_at_points)

    def __call__(self, el, xi):
        """values[component, point] on element(s) el at reference points xi (npts x D)"""
        xi = np.atleast_2d(np.asarray(xi, dtype=float))
        m = np.prod(xi[:, None, :] ** self.expo[None, :, :], axis=2)  # npts x ndof monomials
        psi = m @ self.mono.T
        return self.coef[el] @ psi.T


class TrefftzTents:
    """Base class (src/twavetents.hpp:78-93)."""

    def Propagate(self):
        raise RuntimeError("TrefftzTents virtual!")

    def SetInitial(self, init):
        raise RuntimeError("TrefftzTents virtual!")

    def SetBoundaryCF(self, bddatum):
        raise RuntimeError("TrefftzTents virtual!")


class TWaveTents(TrefftzTents):
    """TWaveTents<D> (src/twavetents.hpp:95-224), state resident on one B200."""

    def __init__(self, order, tps, wavespeedcf, device=0, partition=None, comm_id=None, BBcf=None):
        """partition: None (one GPU) or (rank, nranks, owner) with owner either an int array (one rank per mesh vertex)
        or a tuple (gx, gy, ..) = brick grid over the mesh's bounding box. comm_id: 128-byte NCCL id shared by all ranks
        (default: created on rank 0 and broadcast with torch.distributed, which must be initialised)."""
        self.tps = tps
        self.mesh = tps.mesh
        self.order = int(order)
        mesh = self.mesh
        c_vertex = None
        if callable(wavespeedcf):
            c_el = f64(wavespeedcf(mesh.centroids())).reshape(-1)
            c_vertex = f64(wavespeedcf(mesh.points)).reshape(-1)
        else:
            c = f64(np.atleast_1d(wavespeedcf)).reshape(-1)
            c_el = np.full(mesh.ne, c[0]) if len(c) == 1 else c
        if len(c_el) != mesh.ne:
            raise ValueError("wavespeed: need a scalar, one value per element, or a callable")
        self.c_el = c_el
        self._h = C.c_void_p()
        self.partition = None
        self.qt = BBcf is not None
        if self.qt:
            # TWave(order, tps, wavespeedcf, BBcf) -> QTWaveTents (src/twavetents.cpp:1201-1216)
            if not isinstance(wavespeedcf, PolynomialCF):
                raise TypeError("quasi-Trefftz: the wavespeed must be a ngstrefftz_b200.PolynomialCF (polynomial in x)")
            if not isinstance(BBcf, PolynomialCF):
                BBcf = PolynomialCF.constant(mesh.dim, float(BBcf))
            aa = f64(wavespeedcf.taylor_inverse_square(mesh.points, max(self.order - 2, 0)))
            bb = f64(BBcf.taylor(mesh.points, self.order - 1))
            cc = f64(wavespeedcf.coef_array())
            if partition is None:
                check(lib.ngst_qtwave_create(tps._h, self.order, dptr(c_el), dptr(c_vertex), int(device), wavespeedcf.deg,
                                             dptr(cc), dptr(aa), dptr(bb), C.byref(self._h)))
            else:
                rank, nranks, owner = partition
                owner = brick_vertex_owner(mesh, owner) if isinstance(owner, tuple) else i32(owner)
                if len(owner) != mesh.nv:
                    raise ValueError("partition: need one owner per mesh vertex")
                check(lib.ngst_qtwave_create_partitioned(tps._h, self.order, dptr(c_el), dptr(c_vertex), int(device),
                                                         wavespeedcf.deg, dptr(cc), dptr(aa), dptr(bb), iptr(owner),
                                                         int(nranks), int(rank), C.byref(self._h)))
                self.partition = (int(rank), int(nranks), owner)
        elif partition is None:
            check(lib.ngst_twave_create(tps._h, self.order, dptr(c_el), dptr(c_vertex), int(device), C.byref(self._h)))
        else:
            rank, nranks, owner = partition
            owner = brick_vertex_owner(mesh, owner) if isinstance(owner, tuple) else i32(owner)
            if len(owner) != mesh.nv:
                raise ValueError("partition: need one owner per mesh vertex")
            check(lib.ngst_twave_create_partitioned(tps._h, self.order, dptr(c_el), dptr(c_vertex), int(device),
                                                    iptr(owner), int(nranks), int(rank), C.byref(self._h)))
            self.partition = (int(rank), int(nranks), owner)
        self._bddatum = None
        self._bnd_on_device = False
        self._bnd_dirty = False
        self._bnd_pts = None
        self._quad = None
        self._refresh()
        if self.partition is not None:
            self._own = np.zeros(self.ntents, dtype=np.int32)
            self.loc2glob = np.zeros(self.nel, dtype=np.int32)
            self.owned = np.zeros(self.nel, dtype=np.int32)
            check(lib.ngst_twave_partition(self._h, iptr(self._own), iptr(self.loc2glob), iptr(self.owned)))
            if self.partition[1] > 1 and comm_id is not False:
                self.init_comm(comm_id)

    def init_comm(self, comm_id=None):
        """NCCL bootstrap (ngst_comm_unique_id / ngst_twave_comm_init)."""
        rank = self.partition[0]
        if comm_id is None:
            import torch
            import torch.distributed as dist
            buf = np.zeros(128, dtype=np.uint8)
            if rank == 0:
                check(lib.ngst_comm_unique_id(buf.ctypes.data_as(C.c_void_p)))
            t = torch.from_numpy(buf)
            if dist.get_backend() == "nccl":
                t = t.cuda()
            dist.broadcast(t, src=0)
            comm_id = t.cpu().numpy().tobytes()
        idbuf = C.create_string_buffer(bytes(comm_id), 128)
        check(lib.ngst_twave_comm_init(self._h, C.cast(idbuf, C.c_void_p)))

    def owned_tents(self):
        """Global ids of the tents this rank solves (all tents without a partition)."""
        return self._own if self.partition is not None else np.arange(self.ntents)

    def _refresh(self):
        info = np.zeros(8, dtype=np.int32)
        check(lib.ngst_twave_info(self._h, iptr(info)))
        self.D, _, self.nbasis, self.nip, self.ncomp, self.fosystem, self.nel, self.ntents = (int(x) for x in info)

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            lib.ngst_twave_destroy(self._h)
            self._h = C.c_void_p()

    # ---- sampling helpers
    def quadrature(self):
        """(x[nel,nip,D], weight[nel,nip]) mapped points and |detJ|*w weights."""
        if self._quad is None:
            x = np.zeros((self.nel, self.nip, self.D))
            w = np.zeros((self.nel, self.nip))
            check(lib.ngst_twave_quadrature(self._h, dptr(x), dptr(w)))
            self._quad = (x, w)
        return self._quad

    def MakeWavefront(self, cf, time=0.0):
        """src/twavetents.cpp:734-764: sample cf at all element quadrature points at a fixed time."""
        if isinstance(cf, AnalyticData):  # sampled on the device (cuda/sample.cu)
            self._check_data(cf)
            ncomp = self.D + (1 if cf.first_order else 2)
            wf = np.zeros((self.nel, ncomp * self.nip))
            check(lib.ngst_twave_make_wavefront_analytic(self._h, cf.family, dptr(cf.params), len(cf.params), float(time),
                                                         int(cf.first_order), dptr(wf)))
            return wf
        if not callable(cf):
            wf = f64(cf)
            if wf.ndim != 2 or wf.shape[0] != self.nel or wf.shape[1] % self.nip:
                raise ValueError("wavefront array must be nel x (ncomp*nip)")
            return wf
        x, _ = self.quadrature()
        vals = f64(cf(x.reshape(-1, self.D), float(time)))
        if vals.ndim == 1:
            vals = vals.reshape(-1, 1)
        ncomp = vals.shape[1]
        return np.ascontiguousarray(vals.reshape(self.nel, self.nip, ncomp).transpose(0, 2, 1)).reshape(
            self.nel, ncomp * self.nip)

    # ---- TrefftzTents interface
    def _check_data(self, data):
        if data.D != self.D:
            raise ValueError(f"analytic data is {data.D}-dimensional, the solver {self.D}-dimensional")

    def SetInitial(self, init):
        if isinstance(init, AnalyticData):
            self._check_data(init)
            check(lib.ngst_twave_set_initial_analytic(self._h, init.family, dptr(init.params), len(init.params),
                                                      int(init.first_order)))
            self._refresh()
            return
        wf = self.MakeWavefront(init, 0.0)
        ncomp = wf.shape[1] // self.nip
        check(lib.ngst_twave_set_initial(self._h, dptr(wf), ncomp))
        self._refresh()

    def SetInitialStreamed(self, wf):
        """SetInitial from a host array whose rows the device fetches in first-use order while the first tent levels of the
        next Propagate run (ngst_twave_set_initial_streamed). `wf` (C-contiguous float64, nel x ncomp*nip) must stay
        unchanged until that Propagate has returned; the solver keeps a reference."""
        if not (isinstance(wf, np.ndarray) and wf.dtype == np.float64 and wf.flags.c_contiguous and wf.ndim == 2
                and wf.shape[0] == self.nel and wf.shape[1] % self.nip == 0):
            raise ValueError("SetInitialStreamed: need a C-contiguous float64 array of nel x (ncomp*nip) entries")
        self._streamed = wf
        check(lib.ngst_twave_set_initial_streamed(self._h, dptr(wf), wf.shape[1] // self.nip))
        self._refresh()

    def PrefetchInput(self, wf, bnd_values=None):
        """Start copying the NEXT step's initial rows (and pre-sampled boundary values) to the device while the current step
        runs; UsePrefetched() then makes them the state. Arrays must be page-locked (e.g. torch pinned tensors' .numpy())
        and stay unchanged until UsePrefetched() has been called."""
        self._prefetched = (wf, bnd_values)
        n = 0 if bnd_values is None else len(bnd_values)
        check(lib.ngst_twave_prefetch_input(self._h, dptr(wf), wf.shape[1] // self.nip, dptr(bnd_values), n))

    def UsePrefetched(self):
        check(lib.ngst_twave_use_prefetched(self._h))
        self._bnd_dirty = True
        self._refresh()

    def GetWavefrontAsync(self, out):
        """Copy of the state into the page-locked array `out`, asynchronously; WaitOutput() blocks until it has landed."""
        self._async_out = out
        check(lib.ngst_twave_get_wavefront_async(self._h, dptr(out)))

    def WaitOutput(self):
        check(lib.ngst_twave_wait_output(self._h))

    def SetBoundaryCF(self, bddatum):
        """An AnalyticData datum is sampled on the device by every Propagate (Dirichlet faces take d_t u, Neumann faces
        grad u); any other callable is evaluated on the host per slab as before."""
        self._bddatum = bddatum
        self._bnd_on_device = isinstance(bddatum, AnalyticData)
        self._bnd_dirty = False
        if self._bnd_on_device:
            self._check_data(bddatum)
            check(lib.ngst_twave_set_boundary_analytic(self._h, bddatum.family, dptr(bddatum.params), len(bddatum.params)))

    def _boundary_values(self, timeshift=None):
        npts = lib.ngst_twave_boundary_npoints(self._h)
        if self._bddatum is None or npts == 0:
            return None
        if self._bnd_pts is None:
            xt = np.zeros((npts, self.D + 1))
            kind = np.zeros(npts, dtype=np.int32)
            check(lib.ngst_twave_boundary_points(self._h, dptr(xt), iptr(kind)))
            # the space coordinates do not change from slab to slab: one read-only copy, so that a datum may keep
            # what depends on x alone between calls (tests/_cases.py does)
            x = np.ascontiguousarray(xt[:, :self.D])
            x.flags.writeable = False
            self._bnd_pts = (xt, kind, x)
        xt, kind, x = self._bnd_pts
        t = xt[:, self.D] + (lib.ngst_twave_timeshift(self._h) if timeshift is None else float(timeshift))
        cf = self._bddatum
        Q, D = self.nip, self.D
        fk = kind.reshape(-1, Q)[:, 0]  # kind per boundary face
        out = []
        if isinstance(cf, dict):
            vd = f64(cf["dirichlet"](x, t)).reshape(npts, -1)[:, 0] if (fk == 0).any() else None
            vn = f64(cf["neumann"](x, t)).reshape(npts, D) if (fk == 1).any() else None
        else:
            v = f64(cf(x, t)).reshape(npts, -1)
            vd = v[:, 0]
            vn = v if v.shape[1] == D else None
        if (fk == 0).all():
            return np.ascontiguousarray(vd)
        if (fk == 1).all():
            if vn is None:
                raise ValueError("Neumann boundary needs a D-vector datum")
            return np.ascontiguousarray(vn.reshape(-1, Q, D).transpose(0, 2, 1)).reshape(-1)
        for f, k in enumerate(fk):
            sl = slice(f * Q, (f + 1) * Q)
            out.append(vd[sl] if k == 0 else vn[sl].T.reshape(-1))
        return np.ascontiguousarray(np.concatenate(out))

    def Propagate(self):
        if getattr(self, "_bnd_on_device", False):
            if self._bnd_dirty:  # propagate_with_values replaced the device-side datum in between
                self.SetBoundaryCF(self._bddatum)
            check(lib.ngst_twave_propagate(self._h))
            return
        self.propagate_with_values(self._boundary_values())

    def sample_boundary(self, timeshift):
        """Boundary datum sampled for the slab that starts at `timeshift` (src/twavetents.cpp:373-374)."""
        return self._boundary_values(timeshift)

    def propagate_with_values(self, vals):
        """Propagate with pre-sampled boundary values (the array ngst_twave_set_boundary_values takes)."""
        n = 0 if vals is None else len(vals)
        self._bnd_dirty = True
        check(lib.ngst_twave_set_boundary_values(self._h, dptr(vals), n))
        check(lib.ngst_twave_propagate(self._h))

    def propagate_levels(self, nlevels, vals="auto"):
        """Diagnostics: run only the tents of the first `nlevels` tent levels (timeshift unchanged), with the boundary
        datum of the current slab; see ngst_twave_propagate_levels."""
        if not getattr(self, "_bnd_on_device", False):
            if isinstance(vals, str):
                vals = self._boundary_values()
            n = 0 if vals is None else len(vals)
            check(lib.ngst_twave_set_boundary_values(self._h, dptr(vals), n))
        check(lib.ngst_twave_propagate_levels(self._h, int(nlevels)))

    def reset_time(self, timeshift=0.0):
        check(lib.ngst_twave_set_timeshift(self._h, float(timeshift)))

    def kernel_name(self):
        return lib.ngst_twave_kernel_name(self._h).decode()

    def set_point_wavespeed(self, cf):
        """Energy / Error with the wavespeed polynomial evaluated at every quadrature point (src/twavetents.cpp:785-793)."""
        if cf is None:
            check(lib.ngst_twave_set_wavespeed_poly(self._h, 0, None))
        else:
            cc = f64(cf.coef_array())
            check(lib.ngst_twave_set_wavespeed_poly(self._h, cf.deg, dptr(cc)))

    def set_kernel_mode(self, mode):
        """0 = automatic routing (DMMA kernel for single-wavespeed tents), 1 = generic DFMA kernel for every tent."""
        check(lib.ngst_twave_set_kernel_mode(self._h, int(mode)))

    def set_compact_state(self, on=True):
        """Tents read / write the internal compact-rule copy of the state (default) or the public rows (A/B)."""
        check(lib.ngst_twave_set_compact_state(self._h, int(bool(on))))

    def compact_state(self):
        return bool(lib.ngst_twave_compact_state(self._h))

    def set_peer_exchange(self, on=True):
        """Multi-GPU: exchange inside the tent kernel over peer memory (default where available) or per level with NCCL."""
        check(lib.ngst_twave_set_peer_exchange(self._h, int(on)))  # 2 = halo tents of a level first (A/B)

    def peer_exchange(self):
        """(active, note)"""
        buf = C.create_string_buffer(256)
        on = lib.ngst_twave_peer_exchange(self._h, buf, 256)
        return bool(on), buf.value.decode()

    def set_factor_cache(self, on=True):
        """Keep the inverse of every tent matrix in HBM after the first Propagate (default) or assemble and solve in every
        Propagate (ngst_twave_set_factor_cache)."""
        # 2..6 select among bitwise-equivalent apply kernels (A/B runs), see include/ngst_b200.h
        check(lib.ngst_twave_set_factor_cache(self._h, int(on)))

    def factor_cache(self):
        """(state, bytes): state 0 = not in use, 1 = next Propagate builds it, 2 = next Propagate applies it."""
        b = C.c_longlong(0)
        st = lib.ngst_twave_factor_cache(self._h, C.byref(b))
        return int(st), int(b.value)

    def ntents_fast(self):
        return lib.ngst_twave_ntents_fast(self._h)

    # ---- DeclareETClass methods (src/twavetents.cpp:1160-1178)
    def import_reference_wavefront(self, wf, simd_width=None):
        """A GetWavefront() matrix of the real reference has ncomp * snip columns with snip = nip rounded up to the SIMD
        width (zero-weight pad points, src/twavetents.cpp:192): drop the pad columns of every component block so that it can
        be compared with / loaded into this solver (SetInitial). Point ORDER inside a block is NGSolve's."""
        wf = np.asarray(wf, dtype=float)
        for ncomp in (self.D + 2, self.D + 1):
            if wf.shape[1] % ncomp == 0:
                snip = wf.shape[1] // ncomp
                if snip >= self.nip and (simd_width is None or snip == -(-self.nip // simd_width) * simd_width):
                    return np.ascontiguousarray(wf.reshape(len(wf), ncomp, snip)[:, :, :self.nip].reshape(len(wf), -1))
        raise ValueError("wavefront width does not match (D+1 or D+2) component blocks of >= nip points")

    def SetOutputMirror(self, out):
        """Stream the state rows into the host array `out` (C-contiguous float64, nel x ncomp*nip) during Propagate, each
        row as soon as it is final (cuda/mirror.cu): after Propagate `out` equals GetWavefront() without a separate
        device-to-host copy. None switches the mirror off. Call after SetInitial."""
        if out is None:
            check(lib.ngst_twave_set_output_mirror(self._h, None))
            self._mirror = None
            return
        if not (isinstance(out, np.ndarray) and out.dtype == np.float64 and out.flags.c_contiguous and out.flags.writeable
                and out.size == self.nel * self.ncomp * self.nip):
            raise ValueError("SetOutputMirror: need a writeable C-contiguous float64 array of nel * ncomp * nip entries")
        check(lib.ngst_twave_set_output_mirror(self._h, dptr(out)))
        self._mirror = out  # keeps the page-locked buffer alive as long as the solver writes to it

    def set_mirror_ctas(self, ctas):
        """Copy CTAs the output mirror keeps in flight (default 24)."""
        check(lib.ngst_twave_set_mirror_ctas(self._h, int(ctas)))

    def GetWavefront(self):
        wf = np.zeros((self.nel, self.ncomp * self.nip))
        check(lib.ngst_twave_get_wavefront(self._h, dptr(wf)))
        return wf

    def GetWave(self, U=None):
        """GetWave (src/__init__.py:7-43): element-wise L2 projection of the current state onto polynomials -- u onto
        degree p for a second-order state (the reference's U.dim == 1 branch), (grad u, d_t u) onto degree p-1 for a
        first-order state. Returns an L2Wave (modal coefficients in a basis orthonormal on the reference simplex); when U
        is given (any object with a writable `.vec` numpy array of the right size) the coefficients are also stored there."""
        sizes = np.zeros(3, dtype=np.int32)
        check(lib.ngst_twave_get_wave(self._h, iptr(sizes), None, None, None, None))
        comps, ndof, deg = (int(x) for x in sizes)
        coef = np.zeros((self.nel, comps, ndof))
        psi = np.zeros((self.nip, ndof))
        mono = np.zeros((ndof, ndof))
        expo = np.zeros((ndof, self.D), dtype=np.int32)
        check(lib.ngst_twave_get_wave(self._h, iptr(sizes), dptr(coef), dptr(psi), dptr(mono), iptr(expo)))
        if U is not None:
            np.asarray(U.vec).reshape(-1)[:] = coef.reshape(-1)
        return L2Wave(coef, deg, psi, mono, expo)

    def _error(self, wavefront, wavefront_corr, time, l2only):
        out = C.c_double(0)
        if isinstance(wavefront_corr, AnalyticData):
            self._check_data(wavefront_corr)
            time = self.timeshift() if time is None else float(time)
            if wavefront is None:  # resident state against device-sampled data: nothing crosses PCIe
                check(lib.ngst_twave_error_analytic(self._h, wavefront_corr.family, dptr(wavefront_corr.params),
                                                    len(wavefront_corr.params), time, int(l2only), C.byref(out)))
                return out.value
            d = wavefront_corr.FirstOrder() if self.fosystem else wavefront_corr
            wavefront_corr = self.MakeWavefront(d, time)
        if wavefront is None:
            wavefront = self.GetWavefront()
        fn = lib.ngst_twave_l2error if l2only else lib.ngst_twave_error
        check(fn(self._h, dptr(f64(wavefront)), dptr(f64(wavefront_corr)), C.byref(out)))
        return out.value

    def Error(self, wavefront, wavefront_corr, time=None):
        """src/twavetents.cpp:766-800. wavefront None = the state resident on the device; wavefront_corr may be an
        AnalyticData object, evaluated on the device at `time` (default: the solver's current time)."""
        return self._error(wavefront, wavefront_corr, time, False)

    def L2Error(self, wavefront, wavefront_corr, time=None):
        """src/twavetents.cpp:802-822, same conventions as Error."""
        return self._error(wavefront, wavefront_corr, time, True)

    def Energy(self, wavefront=None):
        out = C.c_double(0)
        wf = None if wavefront is None else f64(wavefront)
        check(lib.ngst_twave_energy(self._h, dptr(wf), C.byref(out)))
        return out.value

    def EnergyNoAlloc(self):
        """Energy of the resident state through ngst_twave_energy_resident_noalloc: static device scratch, no allocation
        (no counterpart in the reference; at most ngst_twave_energy_noalloc_max_elements() elements, single-GPU solvers)."""
        out = C.c_double(0)
        check(lib.ngst_twave_energy_resident_noalloc(self._h, C.byref(out)))
        return out.value

    def MaxAdiam(self):
        out = C.c_double(0)
        check(lib.ngst_twave_max_adiam(self._h, C.byref(out)))
        return out.value

    def LocalDofs(self):
        return self.nbasis

    def GetOrder(self):
        return self.order

    def GetSpaceDim(self):
        return self.D

    def GetInitmesh(self):
        return self.mesh

    def timeshift(self):
        return lib.ngst_twave_timeshift(self._h)

    def last_ms(self):
        return lib.ngst_twave_last_ms(self._h)

    def last_launches(self):
        return lib.ngst_twave_last_launches(self._h)


def GetIntegrationRule(dim, order):
    """(points[nip, dim], weights[nip]) of the rule in force for the reference simplex of dimension `dim` and the given
    order (the solver of polynomial order p samples its state at the rule of order 2p)."""
    nip = C.c_int(0)
    check(lib.ngst_rule(int(dim), int(order), C.byref(nip), None, None))
    xi, w = np.zeros((nip.value, int(dim))), np.zeros(nip.value)
    check(lib.ngst_rule(int(dim), int(order), C.byref(nip), dptr(xi), dptr(w)))
    return xi, w


def SetIntegrationRule(dim, order, points=None, weights=None):
    """Install a rule table dumped from a real NGSolve install, e.g.
        ir = IntegrationRule(TET, 8); points = [list(p.point) for p in ir]; weights = [p.weight for p in ir]
    for every solver created afterwards (ngst_rule_override): wavefront matrices then correspond
    column for column to the reference's. points=None restores the built-in rule. The table is validated (point count of
    the built-in rule, positive weights, exact to `order`); a rejected table raises NgstError."""
    if int(dim) not in (1, 2, 3):
        raise ValueError("SetIntegrationRule: dim must be 1, 2 or 3")
    if points is None:
        check(lib.ngst_rule_override(int(dim), int(order), 0, None, None))
        return
    xi = np.ascontiguousarray(f64(points).reshape(-1, int(dim)))
    w = f64(weights).reshape(-1)
    if len(w) != len(xi):
        raise ValueError("SetIntegrationRule: as many weights as points")
    check(lib.ngst_rule_override(int(dim), int(order), len(w), dptr(xi), dptr(w)))


def brick_vertex_owner(mesh, grid):
    """Vertex owners for a brick partition: the bounding box is cut into grid[0] x grid[1] x .. equal bricks, a vertex
    belongs to the brick that contains it (upper brick on a cut plane); rank = bx + gx * (by + gy * bz)."""
    pts = mesh.points
    lo, hi = pts.min(axis=0), pts.max(axis=0)
    owner = np.zeros(len(pts), dtype=np.int64)
    stride = 1
    for d, g in enumerate(grid):
        if d >= pts.shape[1]:
            if g != 1:
                raise ValueError("partition grid has more directions than the mesh")
            continue
        b = np.floor((pts[:, d] - lo[d]) / (hi[d] - lo[d]) * g + 1e-9).astype(np.int64)
        owner += np.clip(b, 0, g - 1) * stride
        stride *= g
    return i32(owner)


def balanced_vertex_owner(tps, grid):
    """Vertex owners for a cost-balanced brick partition: recursive coordinate bisection into grid[0] x grid[1] x ..
    parts, each cut placed where the accumulated tent cost (tent-elements of the tents pitched at a vertex) splits evenly.
    Vertices are ordered lexicographically (cut direction first), so a cut may run through a vertex layer -- with 2k+1
    layers the plain brick cut leaves one side a whole layer heavier. rank = bx + gx * (by + gy * bz)."""
    mesh = tps.mesh
    pts = mesh.points
    D = pts.shape[1]
    e = tps.export()
    cost = np.zeros(len(pts))
    np.add.at(cost, e["vertex"], np.diff(e["el_off"]).astype(float))
    cost += 1e-9
    owner = np.zeros(len(pts), dtype=np.int64)

    def split(idx, d, stride):
        if d >= len(grid):
            return
        g = grid[d]
        if d >= D:
            if g != 1:
                raise ValueError("partition grid has more directions than the mesh")
            return split(idx, d + 1, stride)
        keys = [pts[idx, (d + k) % D] for k in range(D)]
        order = np.lexsort(keys[::-1])  # primary key: coordinate d
        idx = idx[order]
        cum = np.cumsum(cost[idx])
        cuts = np.searchsorted(cum, cum[-1] * np.arange(1, g) / g, side="left")
        for b, part in enumerate(np.split(idx, cuts)):
            owner[part] += b * stride
            split(part, d + 1, stride * g)

    split(np.arange(len(pts)), 0, 1)
    return i32(owner)


def partition_plan(tps, owner, nranks, rank):
    """Host-only view of a rank's partition plan (ngst_partition_plan): dict of numpy arrays."""
    owner = i32(owner)
    counts = np.zeros(5, dtype=np.int32)
    check(lib.ngst_partition_plan(tps._h, iptr(owner), nranks, rank, iptr(counts), None, None, None, None, None, None, None))
    nown, nloc, nsend, nrecv, nlev = (int(x) for x in counts)
    out = {"own_tents": np.zeros(nown, np.int32), "loc2glob": np.zeros(nloc, np.int32), "owned": np.zeros(nloc, np.int32),
           "send_off": np.zeros(nlev * nranks + 1, np.int32), "send_el": np.zeros(nsend, np.int32),
           "recv_off": np.zeros(nlev * nranks + 1, np.int32), "recv_el": np.zeros(nrecv, np.int32)}
    check(lib.ngst_partition_plan(tps._h, iptr(owner), nranks, rank, iptr(counts), iptr(out["own_tents"]),
                                  iptr(out["loc2glob"]), iptr(out["owned"]), iptr(out["send_off"]), iptr(out["send_el"]),
                                  iptr(out["recv_off"]), iptr(out["recv_el"])))
    out["nlev"] = nlev
    return out


def partition_p2p(tps, owner, nranks, rank):
    """Host-only view of a rank's cross-rank hazard graph and push lists (ngst_partition_p2p): dict of numpy arrays."""
    owner = i32(owner)
    counts = np.zeros(5, dtype=np.int64)
    cp = counts.ctypes.data_as(_capi.c_llp)
    check(lib.ngst_partition_p2p(tps._h, iptr(owner), nranks, rank, cp, None, None, None, None, None, None, None, None, None))
    nown, nl, nr, npush, nte = (int(x) for x in counts)
    out = {"indeg": np.zeros(nown, np.int32), "lsucc_off": np.zeros(nown + 1, np.int32), "lsucc": np.zeros(nl, np.int32),
           "rsucc_off": np.zeros(nown + 1, np.int32), "rsucc_peer": np.zeros(nr, np.int32), "rsucc_tent": np.zeros(nr, np.int32),
           "push_off": np.zeros(nte + 1, np.int32), "push_peer": np.zeros(npush, np.int32), "push_el": np.zeros(npush, np.int32)}
    check(lib.ngst_partition_p2p(tps._h, iptr(owner), nranks, rank, cp, iptr(out["indeg"]), iptr(out["lsucc_off"]),
                                 iptr(out["lsucc"]), iptr(out["rsucc_off"]), iptr(out["rsucc_peer"]), iptr(out["rsucc_tent"]),
                                 iptr(out["push_off"]), iptr(out["push_peer"]), iptr(out["push_el"])))
    return out


def TWave(order, tps, wavespeedcf, BBcf=None, device=0, partition=None, comm_id=None):
    """Factory of src/twavetents.cpp:1194-1228. partition / comm_id: multi-GPU (see TWaveTents)."""
    return TWaveTents(order, tps, wavespeedcf, device=device, partition=partition, comm_id=comm_id, BBcf=BBcf)
