"""ctypes binding of libngst_b200.so (C ABI declared in include/ngst_b200.h).

The library is the product: if it is missing this module raises ImportError -- there is no
Python or CPU fallback for the hot path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libngst_b200.so")


class NgstError(RuntimeError):
    """Raised when a C-ABI call returns a negative status (ngcore::Exception in the reference)."""

    def __init__(self, code, msg):
        super().__init__(f"[ngst {code}] {msg}")
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C ngstrefftz_b200/csrc` (there is no fallback implementation)")
    return C.CDLL(LIB_PATH)


lib = _load()

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_llp = C.POINTER(C.c_longlong)
c_vp = C.c_void_p

_SIGS = {
    "ngst_last_error": (C.c_char_p, []),
    "ngst_version": (C.c_char_p, []),
    "ngst_mesh_create": (C.c_int, [C.c_int, C.c_int, c_dp, C.c_int, c_ip, C.c_int, c_ip, c_ip, C.POINTER(c_vp)]),
    "ngst_mesh_structured": (C.c_int, [C.c_int, c_ip, c_dp, c_dp, C.POINTER(c_vp)]),
    "ngst_mesh_set_boundary_kind": (C.c_int, [c_vp, c_ip]),
    "ngst_mesh_counts": (C.c_int, [c_vp, c_ip]),
    "ngst_mesh_tables": (C.c_int, [c_vp, c_dp, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip]),
    "ngst_mesh_destroy": (None, [c_vp]),
    "ngst_slab_create": (C.c_int, [c_vp, C.c_int, C.POINTER(c_vp)]),
    "ngst_slab_set_max_wavespeed": (C.c_int, [c_vp, c_dp, C.c_int]),
    "ngst_slab_pitch": (C.c_int, [c_vp, C.c_double, C.c_int, C.c_double, c_ip]),
    "ngst_slab_import_tents": (C.c_int, [c_vp, C.c_int, C.c_double, c_ip, c_ip, c_dp, c_dp, c_ip, c_ip, c_dp, c_ip, c_ip,
                                         c_ip, c_ip]),
    "ngst_slab_pitch_level_sync": (C.c_int, [c_vp, C.c_double, C.c_int, C.c_double, c_ip]),
    "ngst_slab_ntents": (C.c_int, [c_vp]),
    "ngst_slab_nlayers": (C.c_int, [c_vp]),
    "ngst_slab_height": (C.c_double, [c_vp]),
    "ngst_slab_max_slope": (C.c_double, [c_vp]),
    "ngst_slab_sizes": (C.c_int, [c_vp, c_llp]),
    "ngst_slab_export": (C.c_int, [c_vp, c_ip, c_ip, c_dp, c_dp, c_dp, c_ip, c_ip, c_dp, c_ip, c_ip, c_ip, c_ip,
                                   c_ip, c_ip]),
    "ngst_slab_levels": (C.c_int, [c_vp, c_ip, c_ip, c_ip]),
    "ngst_slab_destroy": (None, [c_vp]),
    "ngst_twave_create": (C.c_int, [c_vp, C.c_int, c_dp, c_dp, C.c_int, C.POINTER(c_vp)]),
    "ngst_partition_plan": (C.c_int, [c_vp, c_ip, C.c_int, C.c_int, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip]),
    "ngst_partition_p2p": (C.c_int, [c_vp, c_ip, C.c_int, C.c_int, c_llp, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip]),
    "ngst_twave_create_partitioned": (C.c_int, [c_vp, C.c_int, c_dp, c_dp, C.c_int, c_ip, C.c_int, C.c_int,
                                                C.POINTER(c_vp)]),
    "ngst_qtwave_create": (C.c_int, [c_vp, C.c_int, c_dp, c_dp, C.c_int, C.c_int, c_dp, c_dp, c_dp, C.POINTER(c_vp)]),
    "ngst_qtwave_create_partitioned": (C.c_int, [c_vp, C.c_int, c_dp, c_dp, C.c_int, C.c_int, c_dp, c_dp, c_dp, c_ip, C.c_int,
                                                 C.c_int, C.POINTER(c_vp)]),
    "ngst_twave_set_wavespeed_poly": (C.c_int, [c_vp, C.c_int, c_dp]),
    "ngst_comm_unique_id": (C.c_int, [c_vp]),
    "ngst_twave_comm_init": (C.c_int, [c_vp, c_vp]),
    "ngst_twave_set_peer_exchange": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_peer_exchange": (C.c_int, [c_vp, C.c_char_p, C.c_int]),
    "ngst_twave_partition": (C.c_int, [c_vp, c_ip, c_ip, c_ip]),
    "ngst_twave_info": (C.c_int, [c_vp, c_ip]),
    "ngst_twave_quadrature": (C.c_int, [c_vp, c_dp, c_dp]),
    "ngst_twave_set_initial": (C.c_int, [c_vp, c_dp, C.c_int]),
    "ngst_twave_set_initial_streamed": (C.c_int, [c_vp, c_dp, C.c_int]),
    "ngst_twave_prefetch_input": (C.c_int, [c_vp, c_dp, C.c_int, c_dp, C.c_longlong]),
    "ngst_twave_use_prefetched": (C.c_int, [c_vp]),
    "ngst_twave_get_wavefront_async": (C.c_int, [c_vp, c_dp]),
    "ngst_twave_wait_output": (C.c_int, [c_vp]),
    "ngst_twave_boundary_npoints": (C.c_longlong, [c_vp]),
    "ngst_twave_boundary_nvalues": (C.c_longlong, [c_vp]),
    "ngst_twave_boundary_points": (C.c_int, [c_vp, c_dp, c_ip]),
    "ngst_twave_set_boundary_values": (C.c_int, [c_vp, c_dp, C.c_longlong]),
    "ngst_twave_set_initial_analytic": (C.c_int, [c_vp, C.c_int, c_dp, C.c_int, C.c_int]),
    "ngst_twave_make_wavefront_analytic": (C.c_int, [c_vp, C.c_int, c_dp, C.c_int, C.c_double, C.c_int, c_dp]),
    "ngst_twave_set_boundary_analytic": (C.c_int, [c_vp, C.c_int, c_dp, C.c_int]),
    "ngst_twave_set_output_mirror": (C.c_int, [c_vp, c_dp]),
    "ngst_twave_set_mirror_ctas": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_error_analytic": (C.c_int, [c_vp, C.c_int, c_dp, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_double)]),
    "ngst_twave_propagate": (C.c_int, [c_vp]),
    "ngst_twave_propagate_levels": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_timeshift": (C.c_double, [c_vp]),
    "ngst_twave_set_timeshift": (C.c_int, [c_vp, C.c_double]),
    "ngst_twave_kernel_name": (C.c_char_p, [c_vp]),
    "ngst_twave_set_kernel_mode": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_ntents_fast": (C.c_int, [c_vp]),
    "ngst_twave_set_compact_state": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_compact_state": (C.c_int, [c_vp]),
    "ngst_twave_set_factor_cache": (C.c_int, [c_vp, C.c_int]),
    "ngst_twave_factor_cache": (C.c_int, [c_vp, c_llp]),
    "ngst_twave_last_ms": (C.c_double, [c_vp]),
    "ngst_twave_last_launches": (C.c_int, [c_vp]),
    "ngst_twave_get_wavefront": (C.c_int, [c_vp, c_dp]),
    "ngst_twave_get_wave": (C.c_int, [c_vp, c_ip, c_dp, c_dp, c_dp, c_ip]),
    "ngst_twave_energy": (C.c_int, [c_vp, c_dp, c_dp]),
    "ngst_twave_energy_resident_noalloc": (C.c_int, [c_vp, c_dp]),
    "ngst_twave_energy_noalloc_max_elements": (C.c_int, []),
    "ngst_twave_error": (C.c_int, [c_vp, c_dp, c_dp, c_dp]),
    "ngst_twave_l2error": (C.c_int, [c_vp, c_dp, c_dp, c_dp]),
    "ngst_twave_max_adiam": (C.c_int, [c_vp, c_dp]),
    "ngst_twave_basis": (C.c_int, [c_vp, c_ip, c_ip, c_ip, c_dp]),
    "ngst_twave_destroy": (None, [c_vp]),
    "ngst_device_fp64_peak": (C.c_int, [C.c_int, C.c_int, c_dp]),
    "ngst_basis": (C.c_int, [C.c_int, C.c_int, C.c_int, c_ip, c_ip, c_ip, c_dp, c_ip]),
    "ngst_rule": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp]),
    "ngst_rule_transfer": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, c_ip, c_dp]),
    "ngst_rule_override": (C.c_int, [C.c_int, C.c_int, C.c_int, c_dp, c_dp]),
    "ngst_compact_rule": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp]),
}

EXPORTED_SYMBOLS = tuple(_SIGS)

for _name, (_res, _args) in _SIGS.items():
    _f = getattr(lib, _name)  # AttributeError here = the library does not export a declared symbol
    _f.restype = _res
    _f.argtypes = _args


def check(code):
    if code != 0:
        raise NgstError(code, lib.ngst_last_error().decode())


def dptr(a):
    return None if a is None else a.ctypes.data_as(c_dp)


def iptr(a):
    return None if a is None else a.ctypes.data_as(c_ip)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)
