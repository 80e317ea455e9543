"""Analytic data families (ngstrefftz_b200/_analytic.py): the numpy restatements agree with the test data of
tests/_cases.py, derivatives are consistent, and the wave equation holds where claimed. CPU only."""
import numpy as np
import pytest

import ngstrefftz_b200 as ng
from ngstrefftz_b200._analytic import poly_solution_terms
from _cases import poly_solution, standing


def pts(D, n=40, seed=3):
    rng = np.random.default_rng(seed)
    return rng.random((n, D)), rng.random(n)


@pytest.mark.parametrize("D", [1, 2, 3])
@pytest.mark.parametrize("c", [1.0, 1.3])
def test_standing_matches_cases(D, c):
    x, t = pts(D)
    assert np.allclose(ng.StandingWave(D, c)(x, t), standing(D, c)(x, t), rtol=0, atol=1e-15)
    assert np.array_equal(ng.StandingWave(D, c).FirstOrder()(x, t), ng.StandingWave(D, c)(x, t)[:, 1:])


@pytest.mark.parametrize("D", [1, 2, 3])
def test_dirichlet_datum_of_the_standing_wave_is_the_full_evaluation(D):
    """dirichlet(standing) evaluates d_t u alone and keeps its spatial factor for a read-only point array (what a solver
    hands out for every slab): bitwise the column of the full evaluation, also when the times change between calls."""
    from _cases import dirichlet
    f = standing(D, 1.3)
    bd = dirichlet(f, D)
    x, t = pts(D, n=500)
    assert np.array_equal(bd(x, t), f(x, t)[:, D + 1])
    xr = x.copy()
    xr.flags.writeable = False
    for shift in (0.0, 0.25, 0.5):
        assert np.array_equal(bd(xr, t + shift), f(x, t + shift)[:, D + 1])
    x2 = np.ascontiguousarray(x[::-1])   # another array: nothing kept from the first one is used
    x2.flags.writeable = False
    assert np.array_equal(bd(x2, t), f(x2, t)[:, D + 1])
    assert np.array_equal(bd(xr, 0.3), f(x, 0.3)[:, D + 1])


@pytest.mark.parametrize("D", [1, 2, 3])
def test_polynomial_matches_cases(D):
    x, t = pts(D)
    c = 1.2
    got = ng.PolynomialWave(D, poly_solution_terms(D, c))(x, t)
    assert np.allclose(got, poly_solution(D, c)(x, t), rtol=1e-14, atol=1e-14)


@pytest.mark.parametrize("data", [ng.StandingWave(2, 1.5), ng.PlanePulse(2, 1.5, 7.0, 0.2, [1.0, 2.0]),
                                  ng.PlanePulse(3, 0.7, 3.0, -0.1, [1.0, -1.0, 0.5]),
                                  ng.PolynomialWave(3, poly_solution_terms(3, 0.9))])
def test_derivatives_and_wave_equation(data):
    D = data.D
    c = data.params[0] if data.family != 2 else 0.9
    x, t = pts(D, 10)
    h = 1e-4
    f = data(x, t)
    lap = np.zeros(len(x))
    for d in range(D):
        e = np.zeros(D)
        e[d] = h
        fp, fm = data(x + e, t), data(x - e, t)
        assert np.allclose((fp[:, 0] - fm[:, 0]) / (2 * h), f[:, 1 + d], rtol=1e-6, atol=1e-7)
        lap += (fp[:, 1 + d] - fm[:, 1 + d]) / (2 * h)
    fp, fm = data(x, t + h), data(x, t - h)
    assert np.allclose((fp[:, 0] - fm[:, 0]) / (2 * h), f[:, D + 1], rtol=1e-6, atol=1e-7)
    utt = (fp[:, D + 1] - fm[:, D + 1]) / (2 * h)
    assert np.allclose(utt, c * c * lap, rtol=1e-5, atol=1e-5)


def test_polynomial_argument_checks():
    with pytest.raises(ValueError):
        ng.PolynomialWave(2, [(1.0, (1, 2))])
    with pytest.raises(ValueError):
        ng.PolynomialWave(2, [(1.0, (1, -1, 0))])
