"""Worker of tests/test_partition.py::test_halo_exchange_two_ranks_gloo (one process per rank, gloo backend)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ngstrefftz_b200 as ng  # noqa: E402


def main():
    case = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    if case == "cube":
        mesh, grid, dt = ng.UnitCubeMesh(4), (2, 1, 1), 0.15
    else:
        mesh, grid, dt = ng.UnitSquareMesh(8), (1, 2), 0.2
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    owner = ng.brick_vertex_owner(mesh, grid)
    plan = ng.partition_plan(ts, owner, world, rank)
    e = ts.export()
    nlev = plan["nlev"]
    width = 3
    # serial truth: row value = id of the last tent that wrote the element (-1 = initial data)
    truth = -np.ones((mesh.ne, width))
    local = -np.ones((len(plan["loc2glob"]), width))
    glob2loc = -np.ones(mesh.ne, dtype=int)
    glob2loc[plan["loc2glob"]] = np.arange(len(plan["loc2glob"]))
    mine = set(int(t) for t in plan["own_tents"])
    order = np.argsort(e["level"], kind="stable")
    lev_of = e["level"]
    for l in range(nlev):
        for t in order[lev_of[order] == l]:
            els = e["els"][e["el_off"][t]:e["el_off"][t + 1]]
            if int(t) in mine:
                # a tent reads its patch rows before overwriting them: they must be current
                assert np.array_equal(local[glob2loc[els]], truth[els]), f"rank {rank}: stale row read by tent {t}"
                local[glob2loc[els]] = t
            truth[els] = t
        reqs, recvs = [], []
        for q in range(world):
            s = plan["send_el"][plan["send_off"][l * world + q]:plan["send_off"][l * world + q + 1]]
            r = plan["recv_el"][plan["recv_off"][l * world + q]:plan["recv_off"][l * world + q + 1]]
            if len(s):
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(local[s])), q, tag=l))
            if len(r):
                buf = torch.empty((len(r), width), dtype=torch.float64)
                reqs.append(dist.irecv(buf, q, tag=l))
                recvs.append((r, buf))
        for rq in reqs:
            rq.wait()
        for r, buf in recvs:
            local[r] = buf.numpy()
        assert np.array_equal(local, truth[plan["loc2glob"]]), f"rank {rank}: local state differs after level {l}"
    dist.barrier()
    print("HALO_OK", rank, nlev, int(plan["send_off"][-1]), int(plan["recv_off"][-1]))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
