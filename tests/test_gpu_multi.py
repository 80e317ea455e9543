"""Multi-GPU parity: the partitioned Propagate (one process per GPU, per-level NCCL halo exchange) must reproduce the
single-GPU state bit for bit -- same tents, same arithmetic per tent -- and the all-reduced Energy / Error to round-off.
Needs >= 2 GPUs on the box (skipped otherwise; run with `gpurun --gpus 2`)."""
import os
import socket
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import ngstrefftz_b200 as ng
from _cases import dirichlet, standing

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def ngpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("case,world,how", [("cube_p4", 2, "brick"), ("square_p4", 2, "brick"), ("cube_p3", 2, "balanced"),
                                            ("cube_p4", 2, "balanced"), ("cube_p4", 2, "brick-nooverlap"), ("cube_p4", 2, "balanced-nccl"),
                                            ("qt_square_p4", 2, "brick"), ("qt_square_p2", 2, "brick"),
                                            ("cube_p4", 4, "balanced"), ("cube_p3", 8, "brick"), ("cube_p4", 8, "balanced")])
def test_partitioned_matches_single_gpu(built, case, world, how):
    if ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    sys.path.insert(0, HERE)
    from _multi_gpu_worker import CASES, make_solver, problem
    mk, grids, order, dt, nslabs = CASES[case]
    with tempfile.TemporaryDirectory() as out:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
               "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "_multi_gpu_worker.py"), case, out] + \
              (["brick", "2"] if how == "brick-nooverlap" else [how])
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        ranks = [np.load(os.path.join(out, f"rank{k}.npz")) for k in range(world)]
    mesh = mk()
    D = mesh.dim
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(problem(case, mesh)[0])
    assert ts.PitchTents(dt, True, 2 / 3)
    tw, init, bnd, f = make_solver(case, mesh, ts, order)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    e0 = tw.Energy()
    for _ in range(nslabs):
        tw.Propagate()
    ref = tw.GetWavefront()
    en = tw.Energy()
    err = tw.Error(ref, tw.MakeWavefront(f, nslabs * dt))
    ref_dev = None
    if not case.startswith("qt_"):
        data = ng.StandingWave(D, 1.0)
        tw.reset_time()
        tw.SetInitial(data)
        tw.SetBoundaryCF(data)
        for _ in range(nslabs):
            tw.Propagate()
        ref_dev = tw.GetWavefront()
    print(case, world, how, "peer exchange:", [(bool(d["p2p"]), str(d["p2p_note"])) for d in ranks][:1])
    covered = np.zeros(mesh.ne, dtype=int)
    own = []
    for k, d in enumerate(ranks):
        # every local row (owned or ghost) is current after the last level's exchange
        assert np.array_equal(d["wf"], ref[d["loc2glob"]]), f"rank {k}: state differs from the single-GPU run"
        assert bool(d["mirror_ok"]), f"rank {k}: output mirror differs from GetWavefront"
        if ref_dev is not None:
            assert np.array_equal(d["wf_dev"], ref_dev[d["loc2glob"]]), f"rank {k}: device-sampled run differs"
        assert bool(d["compact"]) == tw.compact_state() and int(d["cache"]) == tw.factor_cache()[0]
        np.add.at(covered, d["loc2glob"][d["owned"] == 1], 1)
        own.append(d["own_tents"])
        assert abs(float(d["energy0"]) - e0) <= 1e-13 * e0
        assert abs(float(d["energy"]) - en) <= 1e-13 * en
        assert abs(float(d["error"]) - err) <= 1e-10 * err
    assert np.all(covered == 1)
    assert sorted(np.concatenate(own)) == list(range(ts.GetNTents()))
