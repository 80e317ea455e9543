"""Shared analytic data for the tests (standing waves of test/tents.py:83-105, polynomial Trefftz solutions)."""
import math

import numpy as np


def standing(D, c=1.0):
    """(u, grad u, dt u) of the standing wave used by the reference's SolveWaveTents (test/tents.py:83-105)."""
    sq, pi = math.sqrt(D), math.pi

    def f(x, t):
        t = np.broadcast_to(np.asarray(t, dtype=float), (x.shape[0],))
        s = [np.sin(pi * x[:, d]) for d in range(D)]
        co = [np.cos(pi * x[:, d]) for d in range(D)]
        prod = np.prod(s, axis=0)
        out = [prod * np.sin(pi * t * c * sq) / (sq * pi)]
        for d in range(D):
            g = np.ones_like(prod)
            for k in range(D):
                g = g * (co[k] if k == d else s[k])
            out.append(g * np.sin(pi * t * c * sq) / sq)
        out.append(prod * np.cos(pi * t * c * sq) * c)
        return np.stack(out, axis=1)

    kept = {}

    def spatial_product(x):
        # prod_d sin(pi x_d); kept between calls for a READ-ONLY array (the boundary points a solver hands out for every
        # slab: same object, cannot change), recomputed otherwise
        if not x.flags.writeable and kept.get("x") is x:
            return kept["prod"]
        prod = np.prod([np.sin(pi * x[:, d]) for d in range(D)], axis=0)
        if not x.flags.writeable:
            kept["x"], kept["prod"] = x, prod
        return prod

    def dt_u(x, t):
        """Column D + 1 of f(x, t) alone (the Dirichlet datum): same expression, same bits, a fifth of the work."""
        t = np.broadcast_to(np.asarray(t, dtype=float), (x.shape[0],))
        return spatial_product(x) * np.cos(pi * t * c * sq) * c

    f.dt_u = dt_u
    return f


def poly_solution(D, c=1.0):
    """A degree-3 polynomial solution of u_tt = c^2 Laplace(u): u = x0^3 + 3 x0 (c t)^2 + x0*x_last*ct + 0.5."""
    def f(x, t):
        t = np.broadcast_to(np.asarray(t, dtype=float), (x.shape[0],))
        x0, xl, ct = x[:, 0], x[:, D - 1], c * t
        u = x0 ** 3 + 3 * x0 * ct ** 2 + 0.5
        grad = [np.zeros_like(x0) for _ in range(D)]
        grad[0] = 3 * x0 ** 2 + 3 * ct ** 2
        ut = 6 * x0 * ct * c
        if D > 1:
            u = u + x0 * xl * ct
            grad[0] = grad[0] + xl * ct
            grad[D - 1] = grad[D - 1] + x0 * ct
            ut = ut + x0 * xl * c
        return np.stack([u] + grad + [ut], axis=1)

    return f


def gaussian_1d(x, t):
    """Initial pulse of docs/notebooks/twavetents.ipynb cell 6 shape: exp(-500 (x-1/2)^2), at rest."""
    u = np.exp(-500.0 * (x[:, 0] - 0.5) ** 2)
    return np.stack([u, -1000.0 * (x[:, 0] - 0.5) * u, np.zeros_like(u)], axis=1)


def second_order(f):
    return f


def first_order(f):
    return lambda x, t: f(x, t)[:, 1:]


def dirichlet(f, D):
    dt_u = getattr(f, "dt_u", None)   # data that can evaluate d_t u alone (bench.py samples 17 M boundary points per slab)
    if dt_u is not None:
        return lambda x, t: dt_u(x, t)
    return lambda x, t: f(x, t)[:, D + 1]


def neumann(f, D):
    return lambda x, t: f(x, t)[:, 1:D + 1]
