"""The generated straight-line basis evaluators (basis_eval_gen.cuh, used by tent_kernel_fast) against a numpy
evaluation of the CSR Trefftz basis (src/trefftzfespace.cpp:611-682, src/scalarmappedfe.cpp:437-674). No GPU needed:
the generated header is compiled for the host with g++."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from test_basis_rule import product_basis

HERE = os.path.dirname(os.path.abspath(__file__))
CONFIGS = [(3, 4), (3, 3), (3, 2), (2, 4), (2, 5), (2, 3), (2, 6)]


@pytest.fixture(scope="module")
def genlib(built):
    so = os.path.join(HERE, "_build", "libgen_basis_host.so")
    src = os.path.join(HERE, "gen_basis_host.cpp")
    hdr = os.path.join(HERE, "..", "ngstrefftz_b200", "csrc", "cuda", "basis_eval_gen.cuh")
    os.makedirs(os.path.dirname(so), exist_ok=True)
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", src, "-o", so], check=True)
    lib = C.CDLL(so)
    lib.gen_basis_eval.restype = C.c_int
    lib.gen_basis_eval.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    return lib


def test_generated_header_is_current(built):
    """scripts/gen_basis_eval.py reproduces the committed header byte for byte."""
    import importlib.util
    path = os.path.join(HERE, "..", "scripts", "gen_basis_eval.py")
    spec = importlib.util.spec_from_file_location("gen_basis_eval", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    hdr = os.path.join(HERE, "..", "ngstrefftz_b200", "csrc", "cuda", "basis_eval_gen.cuh")
    before = open(hdr).read()
    mod.main()
    assert open(hdr).read() == before


@pytest.mark.parametrize("D,p", CONFIGS)
@pytest.mark.parametrize("fo", [0, 1])
def test_generated_evaluators(genlib, D, p, fo):
    N, npoly, rowptr, col, val, expo = product_basis(D, p, fo)
    d1 = D + 1
    rng = np.random.default_rng(10 * D + p)
    for _ in range(5):
        x = np.zeros(4)
        x[:d1] = rng.uniform(-1, 1, d1)
        out = np.full((N, 5), np.nan)
        n = genlib.gen_basis_eval(D, p, fo, x.ctypes.data_as(C.POINTER(C.c_double)), out.ctypes.data_as(C.POINTER(C.c_double)))
        assert n == N
        mono = np.prod(x[None, :d1] ** expo, axis=1)
        ref = np.zeros((N, 5))
        for i in range(N):
            for j in range(rowptr[i], rowptr[i + 1]):
                e = expo[col[j]]
                ref[i, 0] += val[j] * mono[col[j]]
                for u in range(d1):
                    if e[u] > 0:
                        eu = e.copy()
                        eu[u] -= 1
                        ref[i, 1 + u] += val[j] * e[u] * np.prod(x[:d1] ** eu)
        assert np.abs(out - ref).max() <= 1e-14 * max(1.0, np.abs(ref).max())
