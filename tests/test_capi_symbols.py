"""The C-ABI library loads and exports every symbol include/ngst_b200.h declares (no compute calls: no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ngst_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ngst_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(built):
    lib = C.CDLL(os.path.join(ROOT, "ngstrefftz_b200", "libngst_b200.so"))
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/ngst_b200.h but not exported"


def test_python_binding_covers_header(built):
    from ngstrefftz_b200 import _capi
    assert set(declared_symbols()) == set(_capi.EXPORTED_SYMBOLS)


def test_no_oracle_in_product():
    """The product must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "ngstrefftz_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cpp", ".hpp", ".cu", ".cuh", ".h")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "liboracle" not in txt and "oracle_" not in txt and "_oracle" not in txt, os.path.join(dp, fn)


def test_error_codes_and_messages(built):
    import ngstrefftz_b200 as ng
    from ngstrefftz_b200._capi import lib, NgstError
    with pytest.raises(NgstError) as ei:
        ng.Mesh(points=np.zeros((1, 2)), elements=np.zeros((0, 3), dtype=np.int32))
    assert ei.value.code == -1
    mesh = ng.UnitSquareMesh(2)
    ts = ng.TentSlab(mesh)
    with pytest.raises(RuntimeError):  # "Wavespeed has not been set!" (python_tents.cpp:46-47)
        ts.PitchTents(0.1)
    assert lib.ngst_version().decode().startswith("ngst_b200")


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback(built):
    """Without a CUDA device the solver must refuse to exist (NGST_ERR_CUDA), not fall back to the CPU."""
    import ngstrefftz_b200 as ng
    from ngstrefftz_b200._capi import NgstError
    mesh = ng.UnitSquareMesh(2)
    ts = ng.TentSlab(mesh)
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.1, True, 2 / 3)
    with pytest.raises(NgstError) as ei:
        ng.TWave(2, ts, 1.0)
    assert ei.value.code == -3
    assert "no CPU fallback" in str(ei.value) or "CUDA" in str(ei.value)
