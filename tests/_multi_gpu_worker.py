"""Worker of tests/test_gpu_multi.py: one process per GPU (NCCL), partitioned Propagate, results written per rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import ngstrefftz_b200 as ng  # noqa: E402
from _cases import dirichlet, standing  # noqa: E402

CASES = {
    # name: (mesh factory, grid by world size, order, dt, nslabs)
    "cube_p4": (lambda: ng.UnitCubeMesh(6), {2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}, 4, 0.1, 2),
    "cube_p3": (lambda: ng.StructuredMesh([8, 4, 4], [0, 0, 0], [2, 1, 1]), {2: (2, 1, 1), 4: (4, 1, 1), 8: (4, 2, 1)}, 3, 0.1, 2),
    "square_p4": (lambda: ng.UnitSquareMesh(12), {2: (2, 1), 4: (2, 2), 8: (4, 2)}, 4, 0.1, 3),
    # quasi-Trefftz (QTWaveTents), c = x + y + 1: fast monomial-space kernel (order 4) and generic kernel (order 2)
    "qt_square_p4": (lambda: ng.UnitSquareMesh(10), {2: (2, 1), 4: (2, 2), 8: (4, 2)}, 4, 0.05, 2),
    "qt_square_p2": (lambda: ng.UnitSquareMesh(8), {2: (2, 1), 4: (2, 2), 8: (4, 2)}, 2, 0.05, 2),
}


def problem(case, mesh):
    """(wavespeed, BBcf or None, initial data, boundary datum, exact solution) of a case."""
    D = mesh.dim
    if case.startswith("qt_"):
        from test_oracle_qt import qt_exact
        f = qt_exact(D)
        return ng.PolynomialCF.affine([1.0] * D, 1.0), 1, f, (lambda x, t: f(x, t)[:, D + 1]), f
    f = standing(D, 1.0)
    return 1.0, None, f, dirichlet(f, D), f


def make_solver(case, mesh, ts, order, **kw):
    c, bb, init, bnd, exact = problem(case, mesh)
    tw = ng.TWave(order, ts, c, BBcf=bb, **kw) if bb is not None else ng.TWave(order, ts, c, **kw)
    return tw, init, bnd, exact


def main():
    case, outdir = sys.argv[1], sys.argv[2]
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    mk, grids, order, dt, nslabs = CASES[case]
    mesh = mk()
    D = mesh.dim
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(problem(case, mesh)[0])
    assert ts.PitchTents(dt, True, 2 / 3)
    # argv[3]: "brick" (default) or "balanced" vertex ownership; argv[4]: kernel mode (2 = exchange not overlapped)
    how = sys.argv[3] if len(sys.argv) > 3 else "brick"
    nccl_only = how.endswith("-nccl")   # per-level NCCL exchange instead of the peer-to-peer exchange inside the kernel
    how = how.replace("-nccl", "")
    owner = ng.balanced_vertex_owner(ts, grids[world]) if how == "balanced" else grids[world]
    tw, init, bnd, f = make_solver(case, mesh, ts, order, device=local, partition=(rank, world, owner))
    if len(sys.argv) > 4:
        tw.set_kernel_mode(int(sys.argv[4]))
    if nccl_only:
        tw.set_peer_exchange(False)
    tw.SetInitial(init)
    tw.SetBoundaryCF(bnd)
    e0 = tw.Energy()
    # output mirror of a partitioned solver: local rows (owned and ghost) streamed to the host during Propagate
    mirror = np.full((tw.nel, tw.ncomp * tw.nip), np.nan)
    tw.SetOutputMirror(mirror)
    for _ in range(nslabs):
        tw.Propagate()
    wf = tw.GetWavefront()
    mirror_ok = bool(np.array_equal(mirror, wf))
    tw.SetOutputMirror(None)
    en = tw.Energy()
    err = tw.Error(wf, tw.MakeWavefront(f, nslabs * dt))
    # the same slabs from data sampled on the device (what bench.py's parity_check runs)
    wf_dev = np.zeros(0)
    if not case.startswith("qt_"):
        data = ng.StandingWave(D, 1.0)
        tw.reset_time()
        tw.SetInitial(data)
        tw.SetBoundaryCF(data)
        for _ in range(nslabs):
            tw.Propagate()
        wf_dev = tw.GetWavefront()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), wf=wf, loc2glob=tw.loc2glob, owned=tw.owned, energy0=e0, energy=en,
             error=err, own_tents=tw.owned_tents(), launches=tw.last_launches(), ms=tw.last_ms(), mirror_ok=mirror_ok,
             wf_dev=wf_dev, compact=tw.compact_state(), cache=tw.factor_cache()[0], p2p=tw.peer_exchange()[0],
             p2p_note=tw.peer_exchange()[1])
    dist.barrier()
    dist.destroy_process_group()
    print("MULTI_OK", rank)


if __name__ == "__main__":
    main()
