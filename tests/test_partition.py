"""Multi-GPU host logic without a GPU: the vertex-ownership partition plan and the per-level
halo exchange lists. The N>1 path is exercised with world_size-2 gloo processes that really exchange the rows."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import ngstrefftz_b200 as ng

HERE = os.path.dirname(os.path.abspath(__file__))


def pitched(mesh, dt=0.15):
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(dt, True, 2 / 3)
    return ts


@pytest.mark.parametrize("mk,grid", [(lambda: ng.UnitCubeMesh(4), (2, 1, 1)), (lambda: ng.UnitCubeMesh(4), (2, 2, 1)),
                                     (lambda: ng.UnitSquareMesh(6), (2, 2)), (lambda: ng.SegMesh(16, 0, 1), (4,)),
                                     (lambda: ng.StructuredMesh([4, 2, 2], [0, 0, 0], [2, 1, 1]), (2, 2, 2))])
@pytest.mark.parametrize("how", ["brick", "balanced"])
def test_plan_is_consistent(built, mk, grid, how):
    mesh = mk()
    ts = pitched(mesh)
    nranks = int(np.prod(grid))
    owner = ng.brick_vertex_owner(mesh, grid) if how == "brick" else ng.balanced_vertex_owner(ts, grid)
    assert set(owner) == set(range(nranks))
    e = ts.export()
    el2v = mesh.elements
    plans = [ng.partition_plan(ts, owner, nranks, r) for r in range(nranks)]
    # every tent has exactly one owner; every element is owned (counted) by exactly one rank
    own = np.concatenate([p["own_tents"] for p in plans])
    assert sorted(own) == list(range(ts.GetNTents()))
    counted = np.zeros(mesh.ne, dtype=int)
    for p in plans:
        np.add.at(counted, p["loc2glob"][p["owned"] == 1], 1)
        assert np.all(np.diff(p["loc2glob"]) > 0)
    assert np.all(counted == 1)
    # local element set = elements with a vertex owned by the rank = union of the patches of its tents
    for r, p in enumerate(plans):
        want = np.nonzero((owner[el2v] == r).any(axis=1))[0]
        assert np.array_equal(p["loc2glob"], want)
        patches = set()
        for t in p["own_tents"]:
            patches.update(e["els"][e["el_off"][t]:e["el_off"][t + 1]])
        assert patches == set(want)
    # send list r -> q at level l is the recv list of q from r, as global element ids in the same order
    nlev = plans[0]["nlev"]
    for l in range(nlev):
        for r in range(nranks):
            for q in range(nranks):
                a, b = plans[r], plans[q]
                s = a["loc2glob"][a["send_el"][a["send_off"][l * nranks + q]:a["send_off"][l * nranks + q + 1]]]
                v = b["loc2glob"][b["recv_el"][b["recv_off"][l * nranks + r]:b["recv_off"][l * nranks + r + 1]]]
                assert np.array_equal(s, v)
                if r == q:
                    assert len(s) == 0


def test_balanced_owner_balances_cost(built):
    """An even number of cells per direction has an odd number of vertex layers: the brick cut leaves the upper bricks a
    whole layer heavier, the cost-balanced bisection cuts through the middle layer."""
    mesh = ng.UnitCubeMesh(8)
    ts = pitched(mesh, 0.1)
    e = ts.export()
    cost = np.zeros(mesh.nv)
    np.add.at(cost, e["vertex"], np.diff(e["el_off"]).astype(float))

    def imbalance(owner):
        w = np.array([cost[owner == r].sum() for r in range(8)])
        return w.max() / w.mean()

    assert imbalance(ng.balanced_vertex_owner(ts, (2, 2, 2))) < 1.03 < imbalance(ng.brick_vertex_owner(mesh, (2, 2, 2)))


def test_plan_argument_errors(built):
    from ngstrefftz_b200._capi import NgstError
    mesh = ng.UnitSquareMesh(2)
    ts = pitched(mesh)
    owner = np.zeros(mesh.nv, dtype=np.int32)
    with pytest.raises(NgstError):
        ng.partition_plan(ts, owner + 5, 2, 0)
    with pytest.raises(NgstError):
        ng.partition_plan(ts, owner, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("case", ["cube", "square"])
def test_halo_exchange_two_ranks_gloo(built, case):
    """Two gloo ranks replay a slab level by level: each rank 'solves' its own tents (writes the tent id into the rows of
    the tent's elements), then sends/receives the planned rows with torch.distributed. After every level every local row of
    every rank must equal the serial truth -- i.e. no tent ever reads a stale ghost row."""
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "_halo_worker.py"), case], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "HALO_OK" in o, o


# ---------------------------------------------------------------------------------------------- peer-to-peer exchange plan
@pytest.mark.parametrize("D,world,grid", [(2, 2, (2, 1)), (2, 4, (2, 2)), (3, 2, (2, 1, 1)), (3, 8, (2, 2, 2)), (1, 2, (2,))])
def test_peer_exchange_plan_replays_the_slab(built, D, world, grid):
    """The cross-rank hazard graph and push lists the tent kernel uses for the peer-to-peer exchange (ngst_partition_p2p),
    replayed on the host with one row per (rank, local element): every rank runs its tents in level order, a tent may only
    run when its counter is zero and every row it reads carries the version the serial order prescribes, pushes reach
    exactly the other holders of an element, and the counters drain."""
    mesh = {1: lambda: ng.SegMesh(24, 0, 1), 2: lambda: ng.UnitSquareMesh(6), 3: lambda: ng.UnitCubeMesh(4)}[D]()
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    assert ts.PitchTents(0.2, True, 2 / 3)
    e = ts.export()
    owner = ng.brick_vertex_owner(mesh, grid)
    plans = [ng.partition_plan(ts, owner, world, r) for r in range(world)]
    p2p = [ng.partition_p2p(ts, owner, world, r) for r in range(world)]
    nt = ts.GetNTents()
    # serial reference: version of an element's row = id of the last tent that wrote it
    version = -np.ones(mesh.ne, dtype=np.int64)
    want_read = {}
    for t in range(nt):
        els = e["els"][e["el_off"][t]:e["el_off"][t + 1]]
        want_read[t] = version[els].copy()
        version[els] = t
    rows = [-np.ones(len(pl["loc2glob"]), dtype=np.int64) for pl in plans]          # per rank: version held per local row
    g2l = [{int(g): i for i, g in enumerate(pl["loc2glob"])} for pl in plans]
    remaining = [pp["indeg"].copy() for pp in p2p]
    te0 = [np.concatenate([[0], np.cumsum([e["el_off"][t + 1] - e["el_off"][t] for t in pl["own_tents"]])]) for pl in plans]
    # interleave the ranks level by level (any interleaving that respects the counters must do)
    order = sorted(((int(e["level"][t]), r, i) for r, pl in enumerate(plans) for i, t in enumerate(pl["own_tents"])))
    done = 0
    for lev, r, i in order:
        t = int(plans[r]["own_tents"][i])
        assert remaining[r][i] == 0, f"tent {t} on rank {r} would start before its predecessors are done"
        els = e["els"][e["el_off"][t]:e["el_off"][t + 1]]
        loc = [g2l[r][int(g)] for g in els]
        assert np.array_equal(rows[r][loc], want_read[t]), f"tent {t} on rank {r} reads a stale row"
        rows[r][loc] = t
        pp = p2p[r]
        for k, g in enumerate(els):
            te = te0[r][i] + k
            peers = pp["push_peer"][pp["push_off"][te]:pp["push_off"][te + 1]]
            pels = pp["push_el"][pp["push_off"][te]:pp["push_off"][te + 1]]
            holders = sorted(q for q in range(world) if q != r and int(g) in g2l[q])
            assert sorted(peers.tolist()) == holders
            for q, pel in zip(peers, pels):
                assert plans[q]["loc2glob"][pel] == g
                rows[q][pel] = t
        for s in pp["lsucc"][pp["lsucc_off"][i]:pp["lsucc_off"][i + 1]]:
            remaining[r][s] -= 1
        for q, s in zip(pp["rsucc_peer"][pp["rsucc_off"][i]:pp["rsucc_off"][i + 1]],
                        pp["rsucc_tent"][pp["rsucc_off"][i]:pp["rsucc_off"][i + 1]]):
            assert q != r
            remaining[q][s] -= 1
        done += 1
    assert done == nt and all((x == 0).all() for x in remaining)
    for r in range(world):  # every rank ends with the final version of all its rows
        assert np.array_equal(rows[r], version[plans[r]["loc2glob"]])
