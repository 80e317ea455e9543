#!/usr/bin/env python
"""Fully symmetric positive interior quadrature rules on the triangle / tetrahedron with (near-)minimal point counts, found
numerically: unknown orbit parameters and weights are solved from the moment equations of the S_(D+1)-invariant
polynomials up to the wanted degree (Newton / least squares from random starts), then VERIFIED against the exact integrals
of ALL monomials up to that degree. Used for the top-face Gram of the fast tent kernel, whose integrand is a polynomial of
degree 2p-2: any exact rule gives the same matrix up to round-off.

  python scripts/gen_compact_rules.py > ngstrefftz_b200/csrc/host/compact_rules.inc
"""
import itertools
import math
import sys

import numpy as np
from scipy.optimize import least_squares

# orbit types by dimension: name -> (number of free parameters, generator of barycentric points)
def orbit_points(dim, kind, par):
    n = dim + 1
    if kind == "c":      # centroid
        base = [1.0 / n] * n
    elif kind == "a":    # (a, .., a, 1 - dim a)
        base = [par[0]] * dim + [1 - dim * par[0]]
    elif kind == "ab":   # triangle: (a, b, 1-a-b); tetrahedron: (a, a, b, 1-2a-b)
        base = [par[0], par[1], 1 - par[0] - par[1]] if dim == 2 else [par[0], par[0], par[1], 1 - 2 * par[0] - par[1]]
    elif kind == "aabb":  # tetrahedron: (a, a, 1/2-a, 1/2-a)
        base = [par[0], par[0], 0.5 - par[0], 0.5 - par[0]]
    else:
        raise ValueError(kind)
    return sorted(set(itertools.permutations(base)))


NPAR = {"c": 0, "a": 1, "ab": 2, "aabb": 1}


def exact_monomial(dim, alpha):
    """integral of prod lambda_i^alpha_i over the reference simplex (volume 1/dim!)"""
    return math.prod(math.factorial(a) for a in alpha) * 1.0 / math.factorial(sum(alpha) + dim)


def invariants(dim, degree):
    """exponent tuples (i2, i3[, i4]) of e2^i2 e3^i3 [e4^i4] with weighted degree <= degree (e1 = 1)"""
    out = []
    rng = range(degree + 1)
    for t in itertools.product(rng, repeat=dim):
        if sum((k + 2) * e for k, e in enumerate(t)) <= degree:
            out.append(t)
    return out


def elem_sym(lam):
    n = lam.shape[1]
    return [sum(np.prod(lam[:, list(c)], axis=1) for c in itertools.combinations(range(n), k)) for k in range(2, n + 1)]


def build(dim, orbits, x):
    pts, wts, k = [], [], 0
    for kind in orbits:
        par = x[k:k + NPAR[kind]]
        w = x[k + NPAR[kind]]
        k += NPAR[kind] + 1
        p = orbit_points(dim, kind, par)
        pts += p
        wts += [w] * len(p)
    return np.array(pts), np.array(wts)


def exact_invariant(dim, expo, cache={}):
    """exact integral of prod e_k^i_k by expanding into monomials (symbolic expansion with dict polynomials)"""
    key = (dim, expo)
    if key in cache:
        return cache[key]
    n = dim + 1
    def mul(p, q):
        r = {}
        for a, ca in p.items():
            for b, cb in q.items():
                c = tuple(x + y for x, y in zip(a, b))
                r[c] = r.get(c, 0) + ca * cb
        return r
    poly = {tuple([0] * n): 1}
    for k, e in enumerate(expo):
        ek = {}
        for c in itertools.combinations(range(n), k + 2):
            ek[tuple(1 if i in c else 0 for i in range(n))] = 1
        for _ in range(e):
            poly = mul(poly, ek)
    val = sum(c * exact_monomial(dim, a) for a, c in poly.items())
    cache[key] = val
    return val


def solve(dim, degree, orbits, seed=0, tries=400):
    inv = invariants(dim, degree)
    target = np.array([exact_invariant(dim, e) for e in inv])
    nunk = sum(NPAR[k] + 1 for k in orbits)
    rng = np.random.default_rng(seed)

    def resid(x):
        pts, wts = build(dim, orbits, x)
        es = elem_sym(pts)
        return np.array([np.sum(wts * np.prod([es[k] ** e for k, e in enumerate(ex)], axis=0)) for ex in inv]) / target - 1

    best = None
    for _ in range(tries):
        x0 = []
        for kind in orbits:
            if kind == "a":
                x0 += [rng.uniform(0.02, 1.0 / dim - 0.02)]
            elif kind == "ab":
                a = rng.uniform(0.02, 0.45)
                x0 += [a, rng.uniform(0.02, max(0.03, (1 - (2 if dim == 3 else 1) * a) - 0.05))]
            elif kind == "aabb":
                x0 += [rng.uniform(0.02, 0.48)]
            x0 += [1.0 / math.factorial(dim) / 30]
        r = least_squares(resid, np.array(x0), xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=400)
        if np.abs(r.fun).max() > 1e-13:
            continue
        pts, wts = build(dim, orbits, r.x)
        if wts.min() <= 0 or pts.min() <= 1e-4:
            continue
        if len(pts) != sum(len(orbit_points(dim, k, [0.1, 0.23][:NPAR[k]])) for k in orbits):
            continue
        score = wts.min() / wts.max()
        if best is None or score > best[0]:
            best = (score, pts, wts)
    if best is None:
        raise RuntimeError(f"no rule found for dim {dim} degree {degree} orbits {orbits}")
    return best[1], best[2]


def duffy_rule(dim, n):
    """dense positive reference rule (Duffy-collapsed Gauss-Jacobi, n points per direction) used only to orthonormalise"""
    from scipy.special import roots_jacobi
    pts, wts = [np.zeros(0)], np.ones(1)
    pts = np.zeros((1, 0))
    scale = np.ones(1)
    for k in range(dim):
        x, w = roots_jacobi(n, dim - 1 - k, 0)
        x, w = (x + 1) / 2, w / 2 ** (dim - k)
        newp = np.concatenate([np.repeat(pts, n, axis=0), (np.tile(x, len(pts)) * np.repeat(scale, n))[:, None]], axis=1)
        wts = np.repeat(wts, n) * np.tile(w, len(pts))
        scale = np.repeat(scale, n) * (1 - np.tile(x, len(pts)))
        pts = newp
    return np.hstack([pts, 1 - pts.sum(axis=1, keepdims=True)]), wts


def solve_conditioned(dim, degree, orbits, seed=0, tries=400):
    """As solve(), for larger systems: the invariant polynomials are orthonormalised in the inner product of a dense exact
    rule (well-conditioned moment equations), the unknowns are box-constrained to interior points / positive weights, and
    the result is polished by minimum-norm Gauss-Newton steps."""
    inv = invariants(dim, degree)

    def invmat(pts):
        es = elem_sym(pts)
        return np.stack([np.prod([es[k] ** e for k, e in enumerate(ex)], axis=0) for ex in inv], axis=1)

    lam, w = duffy_rule(dim, degree)
    R = np.linalg.qr(np.sqrt(w)[:, None] * invmat(lam))[1]
    Rinv = np.linalg.inv(R)
    tgt = (w @ invmat(lam)) @ Rinv

    def resid(x):
        pts, wts = build(dim, orbits, x)
        return (wts @ invmat(pts)) @ Rinv - tgt

    lo, hi = [], []
    for kind in orbits:
        lo += {"c": [], "a": [0.01], "aabb": [0.01], "ab": [0.01, 0.01]}[kind] + [1e-6]
        hi += {"c": [], "a": [1.0 / dim - 0.003], "aabb": [0.49], "ab": [0.49, 0.97]}[kind] + [0.1]
    lo, hi = np.array(lo), np.array(hi)
    rng = np.random.default_rng(seed)
    npts = sum(len(orbit_points(dim, k, [0.1, 0.23][:NPAR[k]])) for k in orbits)
    for _ in range(tries):
        x0 = rng.uniform(lo, hi)
        r = least_squares(resid, x0, bounds=(lo, hi), xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=300)
        if np.abs(r.fun).max() > 1e-11:
            continue
        x = r.x
        for _it in range(6):
            r0 = resid(x)
            J = np.zeros((len(r0), len(x)))
            for k in range(len(x)):
                h = 1e-6 * max(abs(x[k]), 1e-3)
                xp, xm = x.copy(), x.copy()
                xp[k] += h
                xm[k] -= h
                J[:, k] = (resid(xp) - resid(xm)) / (2 * h)
            x = x - np.linalg.lstsq(J, r0, rcond=None)[0]
        pts, wts = build(dim, orbits, x)
        if wts.min() > 0 and pts.min() > 5e-3 and len(np.unique(np.round(pts, 8), axis=0)) == npts == len(pts):
            return pts, wts
    raise RuntimeError(f"no rule found for dim {dim} degree {degree} orbits {orbits}")


def verify(dim, degree, pts, wts):
    worst = 0.0
    for alpha in itertools.product(range(degree + 1), repeat=dim + 1):
        if sum(alpha) > degree:
            continue
        ex = exact_monomial(dim, alpha)
        got = np.sum(wts * np.prod(pts ** np.array(alpha), axis=1))
        worst = max(worst, abs(got - ex) / ex)
    return worst


RULES = [  # (dim, degree, orbits)
    (2, 4, ["a", "a"]),                                # 6 points
    (2, 6, ["a", "a", "ab"]),                          # 12
    (2, 8, ["c", "a", "a", "a", "ab"]),                # 16
    (2, 10, ["c", "a", "a", "ab", "ab", "ab"]),        # 25
    (3, 2, ["a"]),                                     # 4
    (3, 4, ["a", "a", "aabb"]),                        # 14 (exact to degree 5)
    (3, 6, ["a", "a", "a", "ab"]),                     # 24
    (3, 8, ["a", "a", "a", "a", "aabb", "ab", "ab"]),  # 46 (one free parameter; solve_conditioned)
]

if __name__ == "__main__":
    print("// generated by scripts/gen_compact_rules.py -- do not edit. {dim, degree, npoints, barycentric points..., weights...}")
    for dim, degree, orbits in RULES:
        pts, wts = solve_conditioned(dim, degree, orbits) if (dim, degree) == (3, 8) else solve(dim, degree, orbits)
        err = verify(dim, degree, pts, wts)
        assert err < 5e-14, (dim, degree, err)
        print(f"// dim {dim} degree {degree}: {len(wts)} points, orbits {'+'.join(orbits)}, max rel. moment error {err:.1e}, "
              f"min/max weight {wts.min() / wts.max():.3f}", file=sys.stderr)
        print(f"{{{dim}, {degree}, {len(wts)}, {{")
        for p in pts:
            print("  " + ", ".join(f"{v:.17g}" for v in p) + ",")
        print("}, {")
        print("  " + ", ".join(f"{v:.17g}" for v in wts))
        print("}},")
