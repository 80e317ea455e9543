"""Per-phase timers of the oracle (the reference's Timer names, src/twavetents.cpp:84-88, 186-188; reported by bench.py
beside the CPU baseline): they nest as in the reference and do not change the result."""
import numpy as np

import ngstrefftz_b200 as ng
from _cases import dirichlet, standing
from _oracle import Oracle


def run(timing):
    mesh = ng.UnitSquareMesh(4)
    o = Oracle(mesh)
    assert o.pitch(0.3, True, 2 / 3, 1.0)
    o.setup(4, 1.0)
    o.set_initial(standing(2))
    o.set_boundary_cf(dirichlet(standing(2), 2))
    if timing:
        o.phase_timing(True)
    o.propagate(nthreads=2)
    return o


def test_phase_timers_nest_and_leave_the_result_alone(built):
    a, b = run(False), run(True)
    assert np.array_equal(a.wavefront(), b.wavefront())
    assert all(v == 0.0 for v in a.phase_times().values())
    ph = b.phase_times()
    assert set(ph) == set(Oracle.PHASES)
    assert ph["tent"] > 0
    top = ph["tent top calcshape"] + ph["tent top AAt"] + ph["tent top bilinearform"]
    assert 0 < top <= ph["tentel"] <= ph["tent"]
    assert ph["tentel"] + ph["solve"] + ph["tenteval"] <= ph["tent"] * (1 + 1e-9)
    b.phase_timing(True)  # switching on again clears
    assert all(v == 0.0 for v in b.phase_times().values())
