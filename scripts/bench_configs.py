#!/usr/bin/env python
"""Side measurements at BASELINE.json's full sizes for the configs that are NOT the bench line (configs[0..3]):
tents/s, DOF-updates/s and the size-independent properties each run must satisfy (energy decay, error against the
analytic solution, determinism). bench.py stays the contract line (configs[4]'s per-GPU share); this script is evidence
for DESIGN.md section 6 and is what tests/test_gpu_full_size.py runs at reduced repeat counts.

  python scripts/bench_configs.py [--config 0 1 2 3] [--slabs K] [--scale S]   -> one JSON line per config

--scale S shrinks the number of elements by about S (for quick checks); S = 1 is the BASELINE size.
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run_config(idx, slabs=2, scale=1.0, generic_apply=None):
    import ngstrefftz_b200 as ng
    from _cases import dirichlet, gaussian_1d, standing

    qt = False
    if idx == 0:  # 2+1D unit square h = 0.25, order 4, c = 1, one slab (test/tents.py path)
        D, order, c, dt = 2, 4, 1.0, 0.5
        mesh = ng.UnitSquareMesh(4)
        f = standing(D, c)
        init, bnd, cw = f, dirichlet(f, D), c
        exact = f
    elif idx == 1:  # 1+1D, order 6, piecewise-constant c, >= 1e6 tents per slab
        D, order = 1, 6
        n = max(int(100000 / scale), 2000)
        mesh = ng.SegMesh(n, 0.0, 1.0)
        cw = np.where((np.arange(n) // 1000) % 2 == 0, 1.0, 2.0)
        dt = 8e-5 * (100000 / n)
        init, bnd, exact = gaussian_1d, (lambda x, t: 0 * t), None
    elif idx == 2:  # 3+1D unit cube ~100k tets, order 3
        D, order, c = 3, 3, 1.0
        n = max(int(round(26 / scale ** (1 / 3))), 4)
        mesh = ng.UnitCubeMesh(n)
        dt = 0.05 * 26 / n
        f = standing(D, c)
        init, bnd, cw, exact = f, dirichlet(f, D), c, f
    elif idx == 3:  # 2+1D quasi-Trefftz, c = x + y + 1, order 5, 1M triangles
        from test_oracle_qt import qt_exact
        D, order, qt = 2, 5, True
        n = max(int(round(708 / math.sqrt(scale))), 8)
        mesh = ng.UnitSquareMesh(n)
        cw = ng.PolynomialCF.affine([1.0] * D, 1.0)
        dt = 0.0005 * 708 / n
        f = qt_exact(D)
        init, bnd, exact = f, (lambda x, t: f(x, t)[:, D + 1]), f
    else:
        raise SystemExit("config must be 0..3 (configs[4] is bench.py)")

    t0 = time.perf_counter()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(cw)
    assert ts.PitchTents(dt, True, 2 / 3)
    t_pitch = time.perf_counter() - t0
    t0 = time.perf_counter()
    tw = ng.TWave(order, ts, cw, BBcf=1) if qt else ng.TWave(order, ts, cw)
    t_stage = time.perf_counter() - t0
    if generic_apply is not None:   # stored inverses of the generic kernel applied by a warp / a CTA per tent (A/B)
        tw.set_factor_cache({"warp": 3, "cta": 4, "warp-generic": 5}[generic_apply])
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    e0 = tw.Energy()
    ms, en = [], [e0]
    for _ in range(slabs):
        tw.Propagate()
        ms.append(tw.last_ms())
        en.append(tw.Energy())
    best = min(ms)  # with the factor cache: a slab that applies the stored inverses (ms[0] forms them)
    cache_state, cache_bytes = tw.factor_cache()
    ex = ts.export()
    ndof = int(tw.nbasis) * int(ts.GetNTents())
    out = {
        "config": idx, "D": D, "order": order, "quasi_trefftz": qt, "elements": int(mesh.ne),
        "tents": int(ts.GetNTents()), "levels": int(ts.GetNLayers()), "tent_elements": int(len(ex["els"])),
        "nbasis": int(tw.nbasis), "nip": int(tw.nip), "kernel": tw.kernel_name(), "launches": int(tw.last_launches()),
        "pitch_s": t_pitch, "stage_s": t_stage, "ms_per_slab": ms, "tents_per_s": ts.GetNTents() / (best * 1e-3),
        "dof_updates_per_s_single_domain": ndof / (best * 1e-3), "energy": en,
        "factor_cache": {"state": cache_state, "bytes": cache_bytes, "ms_first_slab": ms[0]},
        "wavefront_sum_xor": [int(np.frombuffer(tw.GetWavefront().tobytes(), dtype=np.uint64).sum(dtype=np.uint64)),
                              int(np.bitwise_xor.reduce(np.frombuffer(tw.GetWavefront().tobytes(), dtype=np.uint64)))],
    }
    if exact is not None:
        wf = tw.GetWavefront()
        out["error_vs_exact"] = tw.Error(wf, tw.MakeWavefront(exact, slabs * dt))
        out["energy_norm_exact"] = math.sqrt(2 * tw.Energy(tw.MakeWavefront(exact, slabs * dt)))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, nargs="*", default=[0, 1, 2, 3])
    ap.add_argument("--slabs", type=int, default=2)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--generic-apply", choices=["warp", "cta", "warp-generic"], default=None)
    a = ap.parse_args()
    for idx in a.config:
        print(json.dumps(run_config(idx, a.slabs, a.scale, a.generic_apply)), flush=True)


if __name__ == "__main__":
    main()
