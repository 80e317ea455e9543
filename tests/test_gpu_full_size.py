"""BASELINE.json configs[0..3] at their FULL sizes on the GPU (the oracle does not finish these in seconds): each run must
satisfy the size-independent properties of the method -- energy non-increasing from slab to slab with homogeneous or
exact data, error against the analytic solution at the level the discretisation predicts, every tent routed to the
expected kernel. scripts/bench_configs.py is the driver (it also reports tents/s for DESIGN.md)."""
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "scripts"))

pytestmark = pytest.mark.gpu


def run(idx, slabs=2):
    from bench_configs import run_config
    return run_config(idx, slabs=slabs, scale=1.0)


def non_increasing(en, tol=1e-12):
    return all(b <= a * (1 + tol) for a, b in zip(en, en[1:]))


def test_config0_unit_square_order4(built):
    """2+1D unit square, h = 0.25, order 4, one slab (test/tents.py path)."""
    r = run(0, slabs=1)
    assert r["elements"] == 32 and r["nbasis"] == 25 and r["kernel"] == "tent_kernel_fast"
    assert non_increasing(r["energy"])
    assert r["error_vs_exact"] < 2e-2 * r["energy_norm_exact"]


def test_config1_1d_order6_piecewise_wavespeed_1e6_tents(built):
    """1+1D, order 6, piecewise-constant wavespeed (material interfaces -> CalcTentMacroEl), >= 1e6 tents per slab."""
    r = run(1)
    assert r["tents"] >= 1_000_000 and r["elements"] == 100_000 and r["nbasis"] == 13
    assert non_increasing(r["energy"]) and r["energy"][-1] > 0.999 * r["energy"][0]


def test_config2_unit_cube_100k_tets_order3(built):
    """3+1D unit cube, 26^3 x 6 = 105 456 tets, order 3."""
    r = run(2)
    assert r["elements"] == 105_456 and r["nbasis"] == 30 and r["kernel"] == "tent_kernel_fast"
    assert non_increasing(r["energy"]) and r["energy"][-1] > 0.99 * r["energy"][0]
    assert r["error_vs_exact"] < 1e-3 * r["energy_norm_exact"]


def test_config3_quasi_trefftz_1m_triangles_order5(built):
    """2+1D quasi-Trefftz, c = x + y + 1, order 5, 708^2 x 2 = 1 002 528 triangles."""
    r = run(3)
    assert r["elements"] == 1_002_528 and r["nbasis"] == 36 and r["kernel"] == "tent_kernel_qt_fast"
    assert r["energy"][2] < r["energy"][1] < r["energy"][0]  # the exact solution decays like exp(-k t)
    assert r["error_vs_exact"] < 1e-9 * r["energy_norm_exact"]
