"""ctypes wrapper of oracle/liboracle.so -- the CPU restatement of the reference path (test infrastructure).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = os.path.join(_ROOT, "oracle", "liboracle.so")

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


def _dp(a):
    return None if a is None else a.ctypes.data_as(c_dp)


def _ip(a):
    return None if a is None else a.ctypes.data_as(c_ip)


def build_oracle():
    subprocess.run(["make", "-C", os.path.join(_ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)


def _load():
    if not os.path.exists(_LIB):
        build_oracle()
    lib = C.CDLL(_LIB)
    lib.oracle_create.restype = C.c_void_p
    lib.oracle_create.argtypes = [C.c_int, C.c_int, c_dp, C.c_int, c_ip, C.c_int, c_ip, c_ip, C.c_int, c_ip, c_ip, C.c_int,
                                  c_ip, c_ip, c_ip, c_ip]
    lib.oracle_destroy.argtypes = [C.c_void_p]
    lib.oracle_last_error.restype = C.c_char_p
    lib.oracle_pitch.argtypes = [C.c_void_p, C.c_int, c_dp, C.c_double, C.c_int, C.c_double]
    lib.oracle_pitch_rule.argtypes = [C.c_void_p, C.c_int, c_dp, C.c_double, C.c_int, C.c_double, C.c_int]
    lib.oracle_ntents.argtypes = [C.c_void_p]
    lib.oracle_nlayers.argtypes = [C.c_void_p]
    lib.oracle_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
    lib.oracle_export.argtypes = [C.c_void_p, c_ip, c_ip, c_dp, c_dp, c_dp, c_ip, c_ip, c_dp, c_ip, c_ip, c_ip, c_ip, c_ip,
                                  c_ip]
    lib.oracle_import_tents.argtypes = [C.c_void_p, C.c_int, C.c_double, c_ip, c_ip, c_dp, c_dp, c_ip, c_ip, c_dp, c_ip,
                                        c_ip, c_ip, c_ip]
    lib.oracle_twave_setup.argtypes = [C.c_void_p, C.c_int, c_dp, c_dp, C.c_int, c_dp, c_dp]
    lib.oracle_nbasis.argtypes = [C.c_void_p]
    lib.oracle_set_initial.argtypes = [C.c_void_p, c_dp, C.c_int]
    lib.oracle_boundary_nvalues.restype = C.c_longlong
    lib.oracle_boundary_nvalues.argtypes = [C.c_void_p]
    lib.oracle_boundary_npoints.restype = C.c_longlong
    lib.oracle_boundary_npoints.argtypes = [C.c_void_p]
    lib.oracle_boundary_points.argtypes = [C.c_void_p, c_dp, c_ip]
    lib.oracle_set_boundary_values.argtypes = [C.c_void_p, c_dp, C.c_longlong]
    lib.oracle_propagate.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.oracle_max_cond.restype = C.c_double
    lib.oracle_max_cond.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_propagate_sample.argtypes = [C.c_void_p, C.c_int, c_ip, C.c_int]
    lib.oracle_get_wavefront.argtypes = [C.c_void_p, c_dp]
    lib.oracle_set_phase_timing.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_set_compact_rhs.argtypes = [C.c_void_p, C.c_int, c_dp, c_dp, c_dp]
    lib.oracle_get_phase_times.argtypes = [C.c_void_p, c_dp]
    for f in ("oracle_timeshift", "oracle_max_adiam"):
        getattr(lib, f).restype = C.c_double
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.oracle_energy.restype = C.c_double
    lib.oracle_energy.argtypes = [C.c_void_p, c_dp, C.c_int]
    for f in ("oracle_error", "oracle_l2error"):
        getattr(lib, f).restype = C.c_double
        getattr(lib, f).argtypes = [C.c_void_p, c_dp, c_dp, C.c_int]
    lib.oracle_basis.argtypes = [C.c_int, C.c_int, C.c_int, c_ip, c_ip, c_dp]
    lib.oracle_qt_setup.argtypes = [C.c_void_p, C.c_int, c_dp, c_dp, C.c_int, c_dp, c_dp, C.c_int, c_dp, c_dp, c_dp, C.c_int,
                                    c_dp, c_dp]
    lib.oracle_set_point_wavespeed.argtypes = [C.c_void_p, c_dp]
    lib.oracle_qt_basis.argtypes = [C.c_int, C.c_int, c_dp, c_dp, c_ip, c_ip, c_dp]
    return lib


lib = _load()


def reference_rule(dim, order):
    """Simplex rule of SURVEY.md section 8.3 built from scipy's Gauss-Jacobi nodes (independent of the product)."""
    from scipy.special import roots_jacobi
    n = order // 2 + 1

    def gj(alpha):
        x, w = roots_jacobi(n, alpha, 0)
        return 0.5 * (1 + x), w / 2 ** (alpha + 1)

    if dim == 1:
        x, w = gj(0)
        return x.reshape(-1, 1).copy(), w.copy()
    if dim == 2:
        x0, w0 = gj(1)
        x1, w1 = gj(0)
        pts, ws = [], []
        for i in range(n):
            for j in range(n):
                pts.append((x0[i], x1[j] * (1 - x0[i])))
                ws.append(w0[i] * w1[j])
        return np.array(pts), np.array(ws)
    x0, w0 = gj(2)
    x1, w1 = gj(1)
    x2, w2 = gj(0)
    pts, ws = [], []
    for i in range(n):
        for j in range(n):
            for k in range(n):
                pts.append((x0[i], x1[j] * (1 - x0[i]), x2[k] * (1 - x0[i]) * (1 - x1[j])))
                ws.append(w0[i] * w1[j] * w2[k])
    return np.array(pts), np.array(ws)


class Oracle:
    """CPU oracle over the mesh tables of a ngstrefftz_b200.Mesh (tables stand for NGSolve's MeshAccess)."""

    def __init__(self, mesh):
        t = mesh.tables()
        self.D, self.nel, self.nv = mesh.dim, mesh.ne, mesh.nv
        self._keep = t
        self.centroids = t["pts"][t["el2v"]].mean(axis=1)
        self.h = C.c_void_p(lib.oracle_create(
            mesh.dim, mesh.nv, _dp(t["pts"]), mesh.ne, _ip(t["el2v"]), mesh.nse, _ip(t["bnd2v"]), _ip(t["bnd_kind"]),
            mesh.nedge, _ip(t["edge2v"]), _ip(t["el2edge"]), mesh.nfacet, _ip(t["facet2v"]), _ip(t["el2facet"]),
            _ip(t["facet2el"]), _ip(t["facet2bnd"])))
        self.mesh = mesh
        self.bd_cf = None

    def __del__(self):
        if getattr(self, "h", None):
            lib.oracle_destroy(self.h)
            self.h = None

    def pitch(self, dt, local_ct=False, global_ct=1.0, cmax=1.0, method="edge", rule=0):
        """rule 0: the reference's pick rule; rule 1: minimal (level, vertex number) -- the level-synchronous pitcher's."""
        c = np.ascontiguousarray(np.broadcast_to(np.asarray(cmax, dtype=np.float64), (self.nel,)))
        rc = lib.oracle_pitch_rule(self.h, 2 if method == "edge" else 1, _dp(c), float(dt), int(local_ct), float(global_ct), int(rule))
        if rc < 0:
            raise RuntimeError(lib.oracle_last_error().decode())
        return bool(rc)

    def ntents(self):
        return lib.oracle_ntents(self.h)

    def nlayers(self):
        return lib.oracle_nlayers(self.h)

    def export(self):
        nt = self.ntents()
        sz = (C.c_longlong * 4)()
        lib.oracle_sizes(self.h, sz)
        e = dict(vertex=np.zeros(nt, np.int32), level=np.zeros(nt, np.int32), tbot=np.zeros(nt), ttop=np.zeros(nt),
                 maxslope=np.zeros(nt), nb_off=np.zeros(nt + 1, np.int32), nbv=np.zeros(sz[0], np.int32),
                 nbtime=np.zeros(sz[0]), el_off=np.zeros(nt + 1, np.int32), els=np.zeros(sz[1], np.int32),
                 facet_off=np.zeros(nt + 1, np.int32), facets=np.zeros(sz[2], np.int32),
                 dag_off=np.zeros(nt + 1, np.int32), dag=np.zeros(sz[3], np.int32))
        lib.oracle_export(self.h, _ip(e["vertex"]), _ip(e["level"]), _dp(e["tbot"]), _dp(e["ttop"]), _dp(e["maxslope"]),
                          _ip(e["nb_off"]), _ip(e["nbv"]), _dp(e["nbtime"]), _ip(e["el_off"]), _ip(e["els"]),
                          _ip(e["facet_off"]), _ip(e["facets"]), _ip(e["dag_off"]), _ip(e["dag"]))
        return e

    def import_tents(self, e, dt):
        rc = lib.oracle_import_tents(self.h, len(e["vertex"]), float(dt), _ip(e["vertex"]), _ip(e["level"]), _dp(e["tbot"]),
                                     _dp(e["ttop"]), _ip(e["nb_off"]), _ip(e["nbv"]), _dp(e["nbtime"]), _ip(e["el_off"]),
                                     _ip(e["els"]), _ip(e["facet_off"]), _ip(e["facets"]))
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    def setup(self, order, wavespeed=1.0, rule=None):
        if callable(wavespeed):
            c_el = np.ascontiguousarray(wavespeed(self.centroids), dtype=np.float64).reshape(-1)
            c_v = np.ascontiguousarray(wavespeed(self._keep["pts"]), dtype=np.float64).reshape(-1)
        else:
            c_el = np.ascontiguousarray(np.broadcast_to(np.asarray(wavespeed, dtype=np.float64), (self.nel,)))
            c_v = None
        xi, w = rule if rule is not None else reference_rule(self.D, 2 * order)
        xi = np.ascontiguousarray(xi, dtype=np.float64)
        w = np.ascontiguousarray(w, dtype=np.float64)
        self.order, self.nip, self.xi, self.w, self.c_el = order, len(w), xi, w, c_el
        rc = lib.oracle_twave_setup(self.h, order, _dp(c_el), _dp(c_v), len(w), _dp(xi), _dp(w))
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    def setup_qt(self, order, cpoly, bpoly=None):
        """QTWaveTents (src/twavetents.hpp:225-252): wavespeed polynomial cpoly, optional coefficient B (default 1)."""
        from ngstrefftz_b200._poly import PolynomialCF
        D = self.D
        pts = self._keep["pts"]
        bpoly = bpoly or PolynomialCF.constant(D, 1.0)
        c_el = np.ascontiguousarray(cpoly(self.centroids))
        c_v = np.ascontiguousarray(cpoly(pts))
        xi, w = reference_rule(D, 2 * order)
        vxi, vw = reference_rule(D + 1, 2 * order)
        xi, w, vxi, vw = (np.ascontiguousarray(a, dtype=np.float64) for a in (xi, w, vxi, vw))
        aa = cpoly.taylor_inverse_square(pts, max(order - 2, 0))
        bb = bpoly.taylor(pts, order - 1)
        cc = np.ascontiguousarray(cpoly.coef_array())
        self.order, self.nip, self.xi, self.w, self.c_el = order, len(w), xi, w, c_el
        rc = lib.oracle_qt_setup(self.h, order, _dp(c_el), _dp(c_v), len(w), _dp(xi), _dp(w), cpoly.deg, _dp(cc), _dp(aa),
                                 _dp(bb), len(vw), _dp(vxi), _dp(vw))
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())
        self.set_point_wavespeed(cpoly)

    def set_point_wavespeed(self, cf):
        """Energy / Error with the wavespeed evaluated at every quadrature point (src/twavetents.cpp:785-793, 835-843)."""
        if cf is None:
            lib.oracle_set_point_wavespeed(self.h, None)
            return
        c = np.ascontiguousarray(cf(self.quadrature_points().reshape(-1, self.D)), dtype=np.float64)
        lib.oracle_set_point_wavespeed(self.h, _dp(c))

    def quadrature_points(self):
        t = self._keep
        lam = np.concatenate([self.xi, 1 - self.xi.sum(axis=1, keepdims=True)], axis=1)  # nip x (D+1)
        P = t["pts"][t["el2v"]]  # nel x (D+1) x D
        return np.einsum("qk,ekd->eqd", lam, P)

    def make_wavefront(self, cf, time=0.0):
        x = self.quadrature_points().reshape(-1, self.D)
        v = np.asarray(cf(x, float(time)), dtype=np.float64)
        if v.ndim == 1:
            v = v.reshape(-1, 1)
        nc = v.shape[1]
        return np.ascontiguousarray(v.reshape(self.nel, self.nip, nc).transpose(0, 2, 1)).reshape(self.nel, nc * self.nip)

    def set_initial(self, wf_or_cf):
        wf = self.make_wavefront(wf_or_cf) if callable(wf_or_cf) else np.ascontiguousarray(wf_or_cf, dtype=np.float64)
        self.ncomp = wf.shape[1] // self.nip
        rc = lib.oracle_set_initial(self.h, _dp(wf), self.ncomp)
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    def set_boundary_cf(self, cf):
        self.bd_cf = cf

    def boundary_points(self):
        n = lib.oracle_boundary_npoints(self.h)
        xt = np.zeros((n, self.D + 1))
        kind = np.zeros(n, np.int32)
        if n:
            lib.oracle_boundary_points(self.h, _dp(xt), _ip(kind))
        return xt, kind

    def _boundary_values(self):
        if self.bd_cf is None:
            return None
        xt, kind = self.boundary_points()
        if len(kind) == 0:
            return None
        D, Q = self.D, self.nip
        t = xt[:, D] + lib.oracle_timeshift(self.h)
        x = np.ascontiguousarray(xt[:, :D])
        fk = kind.reshape(-1, Q)[:, 0]
        cf = self.bd_cf
        if isinstance(cf, dict):
            vd = np.asarray(cf["dirichlet"](x, t), dtype=np.float64).reshape(len(t), -1)[:, 0] if (fk == 0).any() else None
            vn = np.asarray(cf["neumann"](x, t), dtype=np.float64).reshape(len(t), D) if (fk == 1).any() else None
        else:
            v = np.asarray(cf(x, t), dtype=np.float64).reshape(len(t), -1)
            vd = v[:, 0]
            vn = v if v.shape[1] == D else None
        out = []
        for f, k in enumerate(fk):
            sl = slice(f * Q, (f + 1) * Q)
            out.append(vd[sl] if k == 0 else vn[sl].T.reshape(-1))
        return np.ascontiguousarray(np.concatenate(out))

    def set_boundary_values(self, vals):
        """Pre-sampled boundary values (what _boundary_values() returns for the current timeshift)."""
        lib.oracle_set_boundary_values(self.h, _dp(vals), 0 if vals is None else len(vals))

    def propagate_sample(self, tents, nthreads=0):
        """TIMING ONLY (bench.py CPU baseline): the listed tents with the boundary values already set, dependencies
        between tents ignored -- the state afterwards is meaningless."""
        tents = np.ascontiguousarray(tents, dtype=np.int32)
        rc = lib.oracle_propagate_sample(self.h, int(nthreads or os.cpu_count() or 1), _ip(tents), len(tents))
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    def propagate(self, nthreads=0, ntents_limit=0, vals="sample"):
        if isinstance(vals, str):
            vals = self._boundary_values()
        lib.oracle_set_boundary_values(self.h, _dp(vals), 0 if vals is None else len(vals))
        nthreads = nthreads or os.cpu_count() or 1
        rc = lib.oracle_propagate(self.h, int(nthreads), int(ntents_limit))
        if rc == -2:
            raise ArithmeticError("oracle: singular local system")
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    def set_compact_rhs(self, xi=None, w=None, M=None):
        """EXPERIMENT: integrate the bottom-face terms on the compact rule (xi, w) after transferring
        the state row with M (ngst_rule_transfer); no arguments = the restated reference algorithm."""
        if xi is None:
            rc = lib.oracle_set_compact_rhs(self.h, 0, None, None, None)
        else:
            xi, w, M = (np.ascontiguousarray(a, dtype=np.float64) for a in (xi, w, M))
            assert M.shape == (len(w), self.nip)
            rc = lib.oracle_set_compact_rhs(self.h, len(w), _dp(xi), _dp(w), _dp(M))
        if rc:
            raise RuntimeError(lib.oracle_last_error().decode())

    PHASES = ("tent top calcshape", "tent top AAt", "tent top bilinearform", "tentel", "solve", "tenteval", "tent")

    def phase_timing(self, on=True):
        """Switch the per-phase timers (the reference's Timer names, src/twavetents.cpp:84-88, 186-188) on / off and clear them."""
        lib.oracle_set_phase_timing(self.h, int(bool(on)))

    def phase_times(self):
        """Thread-seconds per phase since phase_timing(True)."""
        out = np.zeros(7)
        lib.oracle_get_phase_times(self.h, _dp(out))
        return dict(zip(self.PHASES, out.tolist()))

    def wavefront(self):
        out = np.zeros((self.nel, self.ncomp * self.nip))
        lib.oracle_get_wavefront(self.h, _dp(out))
        return out

    def energy(self, wf):
        wf = np.ascontiguousarray(wf, dtype=np.float64)
        return lib.oracle_energy(self.h, _dp(wf), self.ncomp)

    def error(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        return lib.oracle_error(self.h, _dp(a), _dp(b), self.ncomp)

    def l2error(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        return lib.oracle_l2error(self.h, _dp(a), _dp(b), self.ncomp)

    def nbasis(self):
        return lib.oracle_nbasis(self.h)

    def max_cond(self, reset=True):
        """Largest 1-norm condition number of a tent system since the last reset (plain Trefftz path)."""
        return lib.oracle_max_cond(self.h, int(reset))

    def max_adiam(self):
        return lib.oracle_max_adiam(self.h)


def oracle_basis(D, order, fowave):
    N = lib.oracle_nbasis  # noqa: F841 (symbol presence)
    nnz = lib.oracle_basis(D, order, fowave, None, None, None)
    import math
    nb = math.comb(D + order, order) + math.comb(D + order - 1, order - 1) - (1 if fowave else 0)
    rowptr = np.zeros(nb + 1, np.int32)
    col = np.zeros(nnz, np.int32)
    val = np.zeros(nnz)
    lib.oracle_basis(D, order, fowave, _ip(rowptr), _ip(col), _dp(val))
    return rowptr, col, val
