"""Closed-form wave data the solver samples on the device (cuda/sample.cu).

In the reference, initial and boundary data are NGSolve CoefficientFunctions evaluated on the host at every quadrature
point (src/twavetents.cpp:734-764, 373-411). For the data families below `TWave.SetInitial`, `SetBoundaryCF` and
`MakeWavefront` take the object itself and the samples are produced directly in HBM (`ngst_twave_*_analytic`).
Calling the object, `data(x, t)`, evaluates the same formulas with numpy and returns the columns (u, grad u, d_t u):
that is what user scripts pass to the oracle and to `Error`; the solver never goes through it.
"""
import math

import numpy as np


class AnalyticData:
    family = -1

    def __init__(self, D, params, first_order=False):
        self.D = int(D)
        self.params = np.ascontiguousarray(params, dtype=np.float64)
        self.first_order = bool(first_order)

    def FirstOrder(self):
        """The same data in first-order form (grad u, d_t u): SetInitial then switches to the first-order system
        (src/twavetents.hpp:198-203)."""
        other = type(self).__new__(type(self))
        other.__dict__.update(self.__dict__)
        other.first_order = True
        return other

    def full(self, x, t):
        raise NotImplementedError

    def __call__(self, x, t):
        x = np.asarray(x, dtype=float).reshape(-1, self.D)
        t = np.broadcast_to(np.asarray(t, dtype=float), (x.shape[0],))
        v = self.full(x, t)
        return v[:, 1:] if self.first_order else v


class StandingWave(AnalyticData):
    """u = prod_d sin(pi x_d) sin(pi sqrt(D) c t) / (sqrt(D) pi) (test/tents.py:83-105)."""
    family = 0

    def __init__(self, D, c=1.0):
        super().__init__(D, [float(c)])
        self.c = float(c)

    def full(self, x, t):
        D, c = self.D, self.c
        sq, pi = math.sqrt(D), math.pi
        s = [np.sin(pi * x[:, d]) for d in range(D)]
        co = [np.cos(pi * x[:, d]) for d in range(D)]
        prod = np.prod(s, axis=0)
        st, ct = np.sin(pi * t * c * sq), np.cos(pi * t * c * sq)
        out = [prod * st / (sq * pi)]
        for d in range(D):
            g = np.ones_like(prod)
            for k in range(D):
                g = g * (co[k] if k == d else s[k])
            out.append(g * st / sq)
        out.append(prod * ct * c)
        return np.stack(out, axis=1)


class PlanePulse(AnalyticData):
    """u = exp(-a (k.x - c t - s0)^2) with the direction k normalised: a travelling Gaussian pulse."""
    family = 1

    def __init__(self, D, c=1.0, a=50.0, s0=0.0, direction=None):
        k = np.zeros(D) if direction is None else np.asarray(direction, dtype=float).reshape(D)
        if direction is None:
            k[0] = 1.0
        k = k / np.linalg.norm(k)
        super().__init__(D, [float(c), float(a), float(s0), *k])
        self.c, self.a, self.s0, self.k = float(c), float(a), float(s0), k

    def full(self, x, t):
        s = x @ self.k - self.c * t - self.s0
        u = np.exp(-self.a * s * s)
        cols = [u] + [-2.0 * self.a * s * u * self.k[d] for d in range(self.D)] + [2.0 * self.a * s * self.c * u]
        return np.stack(cols, axis=1)


class PolynomialWave(AnalyticData):
    """u = sum_m coef_m x^e_m t^et_m; terms = [(coef, (e_0, .., e_{D-1}, e_t)), ...]. The caller is responsible for u being
    a solution of u_tt = c^2 Laplace(u) where that matters (Trefftz exactness tests)."""
    family = 2

    def __init__(self, D, terms):
        flat = [float(len(terms))]
        self.terms = []
        for coef, expo in terms:
            expo = tuple(int(e) for e in expo)
            if len(expo) != D + 1 or min(expo) < 0:
                raise ValueError("PolynomialWave: every term needs D+1 non-negative exponents (x..., t)")
            self.terms.append((float(coef), expo))
            flat += [float(coef), *map(float, expo)]
        super().__init__(D, flat)

    def full(self, x, t):
        D = self.D
        z = np.concatenate([x, t.reshape(-1, 1)], axis=1)
        out = np.zeros((x.shape[0], D + 2))
        for coef, e in self.terms:
            pw = [z[:, d] ** e[d] for d in range(D + 1)]
            dpw = [e[d] * z[:, d] ** (e[d] - 1) if e[d] > 0 else np.zeros(x.shape[0]) for d in range(D + 1)]
            out[:, 0] += coef * np.prod(pw, axis=0)
            for d in range(D + 1):
                g = coef * dpw[d]
                for k in range(D + 1):
                    if k != d:
                        g = g * pw[k]
                out[:, 1 + d] += g
        return out


def poly_solution_terms(D, c=1.0):
    """tests/_cases.poly_solution as PolynomialWave terms: u = x0^3 + 3 x0 (ct)^2 + x0 x_last ct + 0.5."""
    def e(**kw):
        v = [0] * (D + 1)
        for k, p in kw.items():
            v[D if k == "t" else int(k[1:])] += p
        return tuple(v)
    terms = [(1.0, e(x0=3)), (3.0 * c * c, e(x0=1, t=2)), (0.5, e())]
    if D > 1:
        ex = [0] * (D + 1)
        ex[0] += 1
        ex[D - 1] += 1
        ex[D] += 1
        terms.append((c, tuple(ex)))
    return terms
