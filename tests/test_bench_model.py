"""bench.py's work model and its CPU arm, without a GPU: the algorithmic flop / byte figures are SURVEY.md 8.4's
(the numbers roofline.achieved is computed from), the weak-scaling grids are BASELINE configs[4]'s, and
`bench.py --impl reference` prints one well-formed JSON line on a tiny workload."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_flop_and_byte_model_match_survey_8_4(built):
    import ngstrefftz_b200 as ng
    from ngstrefftz_b200._capi import check, iptr, lib
    one = np.array([1.0])
    # (D, p) -> MFLOP per tent-element, bytes per tent-element quoted in SURVEY.md 8.4 (Qr = unpadded nip, second order)
    want = {(2, 4): (0.18e6, 1600), (1, 6): (13e3, 336), (3, 3): (0.80e6, 5120), (3, 4): (4.65e6, 10000)}
    for (D, p), (f_want, b_want) in want.items():
        sizes = np.zeros(3, dtype=np.int32)
        check(lib.ngst_basis(D, p, 0, iptr(sizes), None, None, None, None))
        N, npoly, nnz = (int(x) for x in sizes)
        nip = np.zeros(1, dtype=np.int32)
        check(lib.ngst_rule(D, 2 * p, iptr(nip), None, None))
        Q = int(nip[0])
        _, f_el = bench.algorithmic_flops(D, p, N, npoly, nnz, Q, one, one)
        _, b_el = bench.algorithmic_bytes(D, Q, D + 2, one, one)
        assert abs(f_el - f_want) <= 0.06 * f_want, (D, p, f_el)
        assert b_el == b_want, (D, p, b_el)
    # config 4 exactly: N = 55, npoly = 70, nnz = 103, Q = 125
    assert bench.algorithmic_flops(3, 4, 55, 70, 103, 125, one, one)[1] == 4646250
    assert ng is not None


def test_weak_scaling_grids():
    import types
    a = types.SimpleNamespace(cells=28)
    assert [int(np.prod(bench.workload(a, n)[0])) * 6 for n in (1, 2, 4, 8)] == [131712, 263424, 526848, 1053696]


def test_reference_arm_prints_one_json_line(built):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--cells", "4", "--cpu-sample-tents", "60"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "tents/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["e2e"] == {"value": d["value"], "unit": "tents/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"]
    ph = cb["phases"]
    assert set(ph) >= {"tent top calcshape", "tent top AAt", "tent top bilinearform", "tentel", "tenteval", "tent"}
    assert ph["tent"]["share_of_tent"] == 1.0
