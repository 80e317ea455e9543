"""Polynomial coefficient functions: the stand-in for NGSolve CoefficientFunctions that the quasi-Trefftz path needs.

QTWaveTents (src/twavetents.cpp:967-1144) evaluates the wavespeed CF at arbitrary space-time points inside a tent and
QTWaveBasis (src/trefftzfespace.hpp:298-332, src/trefftzfespace.cpp:1076-1089) needs the Taylor coefficients of 1/c^2 and
of B at every tent vertex, which the reference gets from symbolic differentiation (ComputeDerivs,
src/trefftzfespace.hpp:59-126). Without NGSolve the wavespeed of the quasi-Trefftz solver is a polynomial in x: it can be
evaluated exactly on the device and its Taylor tables follow from truncated power-series arithmetic.
"""
import itertools
import math

import numpy as np


def monomial_exponents(D, deg):
    """Exponent tuples with |alpha| <= deg, first coordinate outermost (the order of PolBasis::IndexMap2,
    src/trefftzfespace.hpp:128-142, and of the monomial evaluation, src/scalarmappedfe.cpp:526-531)."""
    return [e for e in itertools.product(range(deg + 1), repeat=D) if sum(e) <= deg]


class PolynomialCF:
    """c(x) = sum_alpha coef[alpha] x^alpha in D space dimensions; callable on (n, D) point arrays (time is ignored)."""

    def __init__(self, D, coef):
        self.D = int(D)
        self.coef = {tuple(int(k) for k in a): float(v) for a, v in coef.items() if v != 0.0}
        if any(len(a) != self.D for a in self.coef):
            raise ValueError("PolynomialCF: exponent tuples must have D entries")
        self.deg = max((sum(a) for a in self.coef), default=0)

    @staticmethod
    def affine(a, b):
        """c(x) = a . x + b"""
        a = np.atleast_1d(np.asarray(a, dtype=float))
        D = len(a)
        coef = {tuple(0 for _ in range(D)): float(b)}
        for d in range(D):
            coef[tuple(1 if k == d else 0 for k in range(D))] = float(a[d])
        return PolynomialCF(D, coef)

    @staticmethod
    def constant(D, value):
        return PolynomialCF(D, {tuple(0 for _ in range(D)): float(value)})

    def __call__(self, x, t=None):
        x = np.asarray(x, dtype=float).reshape(-1, self.D)
        out = np.zeros(len(x))
        for a, v in self.coef.items():
            out += v * np.prod(x ** np.asarray(a), axis=1)
        return out

    def coef_array(self):
        """Coefficients in monomial_exponents(D, deg) order (the C ABI's wavespeed-polynomial layout)."""
        return np.array([self.coef.get(e, 0.0) for e in monomial_exponents(self.D, self.deg)])

    # ---- truncated power series at expansion points: dict exponent -> array over points
    def series(self, points, deg):
        """Taylor coefficients d^beta c(x0) / beta!, |beta| <= deg, at every row x0 of `points`."""
        x0 = np.asarray(points, dtype=float).reshape(-1, self.D)
        out = {b: np.zeros(len(x0)) for b in monomial_exponents(self.D, deg)}
        for a, v in self.coef.items():
            for b in out:
                if any(bk > ak for ak, bk in zip(a, b)):
                    continue
                binom = math.prod(math.comb(ak, bk) for ak, bk in zip(a, b))
                out[b] += v * binom * np.prod(x0 ** (np.asarray(a) - np.asarray(b)), axis=1)
        return out

    def taylor(self, points, deg):
        """(npoints, nmono) table of d^beta c(x0)/beta! in monomial_exponents(D, deg) order."""
        s = self.series(points, deg)
        return np.ascontiguousarray(np.stack([s[b] for b in monomial_exponents(self.D, deg)], axis=1))

    def taylor_inverse_square(self, points, deg):
        """(npoints, nmono) table of d^beta (1/c^2)(x0)/beta!: reciprocal of the truncated series of c^2."""
        expo = monomial_exponents(self.D, deg)
        c = self.series(points, deg)
        sq = _series_mul(c, c, expo)
        inv = _series_reciprocal(sq, expo)
        return np.ascontiguousarray(np.stack([inv[b] for b in expo], axis=1))


def _series_mul(a, b, expo):
    out = {e: 0.0 for e in expo}
    for e in expo:
        acc = 0.0
        for f in expo:
            if all(fk <= ek for fk, ek in zip(f, e)):
                g = tuple(ek - fk for fk, ek in zip(f, e))
                acc = acc + a[f] * b[g]
        out[e] = acc
    return out


def _series_reciprocal(s, expo):
    zero = tuple(0 for _ in expo[0])
    r = {zero: 1.0 / s[zero]}
    for e in sorted(expo, key=sum):
        if e == zero:
            continue
        acc = 0.0
        for f in expo:
            if f != zero and all(fk <= ek for fk, ek in zip(f, e)):
                g = tuple(ek - fk for fk, ek in zip(f, e))
                acc = acc + s[f] * r[g]
        r[e] = -acc / s[zero]
    return r
