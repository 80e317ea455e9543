#!/usr/bin/env python
"""Benchmark of the tent-pitched Trefftz-DG wave propagator (BASELINE.json: tents/s and DOF-updates/s,
3+1D order-4, 1/2/4/8 B200 vs host CPU).

A "step" is one Propagate() of one tent slab (TWaveTents<3>::Propagate, src/twavetents.cpp:72-175) over the
per-GPU share of BASELINE configs[4]: unit-cube bricks of 28^3 cells x 6 tets = 131 712 tets per GPU,
order 4, c = 1, standing-wave data, Dirichlet boundary (input 4). Weak scaling:
1 GPU 28x28x28 cells, 2 GPUs 56x28x28, 4 GPUs 56x56x28, 8 GPUs 56x56x56.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (CUDA, libngst_b200.so through the C ABI)
  python bench.py --impl reference [--steps K] [--warmup W]      CPU arm: the oracle restatement of the reference's
                                                                 Propagate on all host cores (the reference itself
                                                                 needs NGSolve and cannot be built)
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

_RESULT = []  # the JSON line, printed by main() on the real stdout
METRIC = "tents/s (3+1D order-4 Trefftz-DG tent slab, one Propagate per step)"
WEAK_GRIDS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None,
                    help="timed steps (default 10; 3 for --impl reference, whose step is seconds of CPU work). A step takes "
                         "12-40 ms here; 10 steps amortise the unoverlapped first input / last output of the pipelined e2e arm "
                         "and give the clock sampler several samples")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=28, help="cells per side of the per-GPU brick")
    ap.add_argument("--order", type=int, default=4)
    ap.add_argument("--dt", type=float, default=0.05, help="slab height")
    ap.add_argument("--cpu-sample-tents", type=int, default=0, help="tents in the bounded CPU sample (0 = auto)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 28^3 cells per GPU (default); strong: the 56^3-cell, 1 053 696-tet slab on every N")
    ap.add_argument("--no-parity-check", action="store_true",
                    help="N>1: skip the bitwise comparison of the partitioned run with a single-GPU run on rank 0")
    ap.add_argument("--mirror-ctas", type=int, default=0, help="copy CTAs of the output mirror (0 = library default)")
    ap.add_argument("--no-factor-cache", action="store_true",
                    help="assemble and solve every tent system in every Propagate (no stored inverses): the headline becomes the "
                         "cold figure that is otherwise reported beside it as `assembly_every_step`")
    ap.add_argument("--apply-four-ctas", action="store_true",
                    help="factor cache: apply kernel variant with 128 registers / 55 KB shared memory, four CTAs per SM (A/B)")
    ap.add_argument("--no-compact-state", action="store_true", help="tents read / write the public order-2p rows (A/B)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-mirror", action="store_true", help="e2e arm: explicit get_wavefront copy instead of the output mirror")
    ap.add_argument("--streamed-input", action="store_true",
                    help="sequential e2e arm: ngst_twave_set_initial_streamed (spatial sweep) instead of the blocking set_initial")
    ap.add_argument("--no-pipelined-e2e", action="store_true",
                    help="headline e2e = the sequential call sequence instead of pipelined steps (prefetched input, asynchronous output)")
    ap.add_argument("--no-sequential-mirror", action="store_true", help="sequential e2e arm: explicit get_wavefront instead of the mirror")
    ap.add_argument("--no-device-data", action="store_true", help="skip the informational device-sampled-data arm")
    ap.add_argument("--partition", default="balanced", choices=["balanced", "brick"],
                    help="N>1: cost-balanced recursive bisection (default) or plain brick cut of the bounding box")
    ap.add_argument("--no-peer-exchange", action="store_true",
                    help="N>1: halo tents + grouped ncclSend/ncclRecv per level instead of the exchange inside the tent kernel")
    ap.add_argument("--halo-first", action="store_true", help="N>1, peer-to-peer exchange: ticket the halo tents of a level first (A/B)")
    ap.add_argument("--no-overlap", action="store_true", help="N>1: exchange after each level instead of overlapped")
    ap.add_argument("--level-launches", action="store_true",
                    help="N=1: always one launch per tent level (profiling; default: automatic, see DESIGN.md section 4.5)")
    ap.add_argument("--single-launch", action="store_true",
                    help="N=1: dependency-driven single launch per slab instead of one launch per tent level (A/B)")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = 3 if args.impl == "reference" else 10
    return args


def workload(args, ngpus):
    g = WEAK_GRIDS.get(ngpus)
    if g is None:
        raise SystemExit("--gpus must be 1, 2, 4 or 8")
    if getattr(args, "scaling", "weak") == "strong":
        return [2 * args.cells] * 3, g  # BASELINE configs[4] itself (56^3 cells x 6 = 1 053 696 tets) cut into N parts
    cells = [args.cells * k for k in g]
    return cells, g


def standing3(x, t, c=1.0):
    from _cases import standing
    return standing(3, c)(x, t)


# ------------------------------------------------------------------------------------------------ flop model
def algorithmic_flops(D, p, N, npoly, nnz, Qr, nels_per_tent, ndom_per_tent):
    """per tent-element F_el, per tent sum F_el + (2/3) n^3 + 2 n^2 (LU solve, n = ndom*N)."""
    d1 = D + 1
    f_el = (2 * N * N * (d1 + 1) * Qr + 2 * N * Qr * (2 * (d1 + 1) + (3 * D + 1)) +
            2 * Qr * (d1 * (d1 * npoly + 2 * nnz) + D * npoly + 2 * nnz))
    n = ndom_per_tent * N
    per_tent = nels_per_tent * f_el + (2.0 / 3.0) * n ** 3 + 2.0 * n ** 2
    return float(per_tent.sum()), f_el


def algorithmic_bytes(D, Qr, ncomp, nels_per_tent, nnb_per_tent):
    d1 = D + 1
    b_el = 2 * ncomp * Qr * 8
    per_tent = nels_per_tent * b_el + (1 + nnb_per_tent) * d1 * 8 + 4 * (nels_per_tent * d1 + nnb_per_tent)
    return float(per_tent.sum()), b_el


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                pw.append(float(parts[2]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "power_w_max": float(max(pw)), "samples": len(sm)}
        return out


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_sample(args, cells, sample_tents, steps, warmup, nthreads):
    """Times the oracle's multithreaded per-tent work (restating src/twavetents.cpp:90-171, OpenMP over independent tents
    as the reference's RunParallelDependency does) on a bounded STRATIFIED sample of the same slab: every k-th tent of the
    level-ordered tent list, so that all levels -- small boundary tents of level 0 and full interior patches alike --
    are represented in proportion. The cost of a tent does not depend on the values in its rows, so the sample is run
    without its predecessors (oracle_propagate_sample). The boundary datum is sampled before the timed region (as in
    the GPU arm). Reported in tents/s of the slab's AVERAGE tent: tent-elements/s of the sample / mean tent-elements
    per tent of the slab."""
    import ngstrefftz_b200 as ng
    from _cases import dirichlet, standing
    from _oracle import Oracle

    mesh = ng.StructuredMesh(cells, [0.0] * 3, [c / args.cells for c in cells])
    o = Oracle(mesh)
    assert o.pitch(args.dt, True, 2 / 3, 1.0)
    f = standing(3, 1.0)
    o.setup(args.order, 1.0)
    o.set_initial(f)
    o.set_boundary_cf(dirichlet(f, 3))
    e = o.export()
    nt = o.ntents()
    sample_tents = min(sample_tents, nt)
    order = np.argsort(e["level"], kind="stable")  # level order, ascending id inside a level
    pick = np.unique(np.linspace(0, nt - 1, sample_tents).round().astype(np.int64))
    sample = order[pick].astype(np.int32)
    nels_all = np.diff(e["el_off"])
    nels = nels_all[sample]
    o.set_boundary_values(o._boundary_values())  # outside the timed region
    times = []
    for i in range(warmup + steps):
        if i == warmup:
            o.phase_timing(True)  # per-phase thread-seconds of the timed passes (reference Timer names)
        t0 = time.perf_counter()
        o.propagate_sample(sample, nthreads=nthreads)
        t1 = time.perf_counter()
        if i >= warmup:
            times.append(t1 - t0)
    sec = float(np.sum(times))
    te_per_s = float(nels.sum()) * len(times) / sec
    ph = o.phase_times()
    total = ph["tent"] or 1.0
    phases = {k: {"thread_s": round(v, 4), "share_of_tent": round(v / total, 4)} for k, v in ph.items()}
    return {"tents_per_s": te_per_s / float(nels_all.mean()), "tent_elements_per_s": te_per_s,
            "sample_tents_per_s": len(sample) * len(times) / sec, "sec_per_step": sec / max(len(times), 1),
            "sample_tents": int(len(sample)), "slab_tents": int(nt), "nbasis": o.nbasis(),
            "tent_elements": int(nels.sum()), "elements_per_tent_sample": float(nels.mean()),
            "elements_per_tent_slab": float(nels_all.mean()), "levels": int(e["level"].max()) + 1, "phases": phases}


def cpu_baseline_entry(r, cores, passes):
    return {"value": r["tents_per_s"], "unit": "tents/s", "cores": cores, "kind": "port", "same_config": True,
            "sample": f"{r['sample_tents']} tents spread evenly over the level-ordered tent list (all {r['levels']} levels) "
                      f"of the same {r['slab_tents']}-tent slab, {passes}; {r['elements_per_tent_sample']:.2f} elements "
                      f"per tent in the sample vs {r['elements_per_tent_slab']:.2f} in the slab, value = tent-elements/s / "
                      f"slab mean elements per tent; boundary datum sampled before the timed region; oracle restatement of "
                      f"the reference's per-tent work with OpenMP over independent tents (the reference itself needs "
                      f"NGSolve: unbuildable here)",
            "tent_elements_per_s": r["tent_elements_per_s"], "phases": r["phases"]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cells, grid = workload(args, args.gpus)
    cores = os.cpu_count() or 1
    sample = args.cpu_sample_tents or 4000
    r = cpu_sample(args, cells, sample, args.steps, max(args.warmup, 1), cores)
    N = r["nbasis"]
    line = {
        "impl": "reference", "metric": METRIC, "value": r["tents_per_s"], "unit": "tents/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": 1e3 * r["sec_per_step"],
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "dof_updates_per_s": r["tents_per_s"] * N,
        "config": {"workload": f"3+1D unit-cube brick {cells[0]}x{cells[1]}x{cells[2]} cells x 6 tets, order "
                               f"{args.order}, c=1, standing wave, Dirichlet, dt={args.dt}, edge pitching "
                               f"local_ct global_ct=2/3", "tets": int(np.prod(cells) * 6), "order": args.order,
                   "nbasis": N, "slab_tents": r["slab_tents"]},
        "cpu_baseline": cpu_baseline_entry(r, cores, f"{args.steps} timed pass(es) after {max(args.warmup, 1)} warm-up"),
        "e2e": {"value": r["tents_per_s"], "unit": "tents/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _RESULT.append(json.dumps(line))


def bind_to_gpu_numa(local):
    """Run this rank (and allocate its pinned buffers, by first touch) on the NUMA node its GPU hangs off: with 8 ranks on
    one host every step moves 8 x 1.45 GB through host memory, and a remote node halves the PCIe copy rate."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        addr = bus.lower()[-12:]  # 00000000:1B:00.0 -> 0000:1b:00.0
        with open(f"/sys/bus/pci/devices/{addr}/numa_node") as fh:
            node = int(fh.read().strip())
        if node < 0:
            return {"node": node, "note": "the platform reports no NUMA affinity for this GPU"}
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = set()
            for part in fh.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return {"node": node, "cpus": len(cpus)}
        return {"node": node, "note": "no allowed CPU on that node"}
    except Exception as ex:
        return {"error": repr(ex)[:200]}


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    import ngstrefftz_b200 as ng
    from ngstrefftz_b200 import _capi
    from _cases import dirichlet, standing

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
        raise SystemExit(f"WORLD_SIZE={world} does not match --gpus {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    cells, grid = workload(args, world)
    D, order = 3, args.order
    f = standing(D, 1.0)
    mesh = ng.StructuredMesh(cells, [0.0] * 3, [c / args.cells for c in cells])
    ts = ng.TentSlab(mesh, "edge")
    ts.SetMaxWavespeed(1.0)
    t0 = time.perf_counter()
    assert ts.PitchTents(args.dt, True, 2 / 3)
    pitch_s = time.perf_counter() - t0
    # informational: the level-synchronous pitcher (partition-independent pick rule, other tents) on the same mesh
    pitch_ls = None
    if rank == 0:
        ts2 = ng.TentSlab(mesh, "edge")
        ts2.SetMaxWavespeed(1.0)
        t0 = time.perf_counter()
        ok2 = ts2.PitchTents(args.dt, True, 2 / 3, level_sync=True)
        pitch_ls = {"s": time.perf_counter() - t0, "tents": ts2.GetNTents(), "levels": ts2.GetNLayers(), "ok": bool(ok2),
                    "omp_threads": os.environ.get("OMP_NUM_THREADS", "all")}
        del ts2
    if world > 1:
        owner = ng.balanced_vertex_owner(ts, grid) if args.partition == "balanced" else ng.brick_vertex_owner(mesh, grid)
        tw = ng.TWave(order, ts, 1.0, device=local, partition=(rank, world, owner))
        if args.no_overlap:
            tw.set_kernel_mode(2)
        if args.no_peer_exchange:
            tw.set_peer_exchange(False)
        elif args.halo_first:
            tw.set_peer_exchange(2)
    else:
        tw = ng.TWave(order, ts, 1.0, device=local)
        if args.mirror_ctas:
            tw.set_mirror_ctas(args.mirror_ctas)
        if args.single_launch:
            tw.set_kernel_mode(3)
        if args.level_launches:
            tw.set_kernel_mode(2)
    if args.no_compact_state:
        tw.set_compact_state(False)
    if args.no_factor_cache:
        tw.set_factor_cache(False)
    elif args.apply_four_ctas:
        tw.set_factor_cache(2)
    e = ts.export()
    ntents_global = ts.GetNTents()
    mine = tw.owned_tents() if world > 1 else np.arange(ntents_global)
    nels = np.diff(e["el_off"])[mine]
    nnb = np.diff(e["nb_off"])[mine]
    wf0 = tw.MakeWavefront(f, 0.0)          # host, nel_local x (5*125)
    tw.SetInitial(wf0)
    N, Q, ncomp = tw.nbasis, tw.nip, tw.ncomp
    ndom = np.ones_like(nels)
    sizes = np.zeros(3, dtype=np.int32)
    from ngstrefftz_b200._capi import check, iptr, lib
    check(lib.ngst_twave_basis(tw._h, iptr(sizes), None, None, None))
    npoly, nnz = int(sizes[1]), int(sizes[2])
    flops_step, f_el = algorithmic_flops(D, order, N, npoly, nnz, Q, nels, ndom)
    bytes_step, b_el = algorithmic_bytes(D, Q, ncomp, nels, nnb)

    # boundary datum pre-sampled per slab on the host (it depends on timeshift, src/twavetents.cpp:373-374)
    bd = dirichlet(f, D)
    tw.SetBoundaryCF(bd)
    nsteps_total = args.warmup + args.steps
    bvals = []
    h = ts.GetSlabHeight()
    for i in range(2 * nsteps_total + 3):
        bvals.append(tw.sample_boundary(i * h))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- resident arm: state stays in HBM, boundary samples uploaded before the timed Propagate
    def resident_step(vals):
        """Boundary samples up (host work, outside the device-timed region), ranks aligned, then Propagate. Without the
        alignment the CUDA-event time of a rank that is ready early includes the wait for its neighbours' host-side upload
        (the first exchange / barrier of the slab cannot complete before they arrive) -- host jitter, not device work."""
        n = 0 if vals is None else len(vals)
        check(lib.ngst_twave_set_boundary_values(tw._h, _capi.dptr(vals), n))
        if world > 1:
            dist.barrier()
        check(lib.ngst_twave_propagate(tw._h))

    step = 0
    for _ in range(args.warmup):
        resident_step(bvals[step]); step += 1
    barrier()
    clk = ClockSampler(local)
    dev_ms, launches = [], 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        resident_step(bvals[step]); step += 1
        dev_ms.append(tw.last_ms())
        launches += tw.last_launches()
    barrier()
    wall_s = time.perf_counter() - t0
    clocks = clk.stop()
    dev_s = float(np.sum(dev_ms)) * 1e-3
    energy = tw.Energy()
    cache_state, cache_bytes = tw.factor_cache()
    kernel_variant = ("apply" if cache_state == 2 else "assemble") if tw.kernel_name() == "tent_kernel_fast" else "other"

    # ---------------- the same steps with the factor cache off: every tent system assembled and solved in every Propagate
    # (what the reference does, src/twavetents.cpp:119-164) -- reported beside the headline as `assembly_every_step`
    cold = None
    if cache_state == 2:
        tw.set_factor_cache(False)
        for _ in range(1):
            resident_step(bvals[step]); step += 1
        barrier()
        cold_ms, cold_launches = [], 0
        for _ in range(args.steps):
            resident_step(bvals[step]); step += 1
            cold_ms.append(tw.last_ms())
            cold_launches += tw.last_launches()
        barrier()
        cold = {"sec": float(np.sum(cold_ms)) * 1e-3, "launches": cold_launches}
        tw.set_factor_cache(2 if args.apply_four_ctas else True)

    # ---------------- e2e arm: the plugin call sequence with HOST buffers inside the timed region
    e2e = None
    if not args.no_e2e:
        nbytes = wf0.nbytes
        pin_in = torch.empty(wf0.size, dtype=torch.float64, pin_memory=True)
        pin_out = torch.empty(wf0.size, dtype=torch.float64, pin_memory=True)
        pin_in.numpy()[:] = wf0.reshape(-1)
        bpin = [torch.from_numpy(b).pin_memory() if b is not None else None for b in bvals[:1]]
        b0 = bpin[0]
        in_ptr = _capi.C.cast(pin_in.data_ptr(), _capi.c_dp)
        out_ptr = _capi.C.cast(pin_out.data_ptr(), _capi.c_dp)
        b_ptr = _capi.C.cast(b0.data_ptr(), _capi.c_dp) if b0 is not None else None
        nb = 0 if b0 is None else b0.numel()

        # Sequential form: set_initial, set_boundary_values, propagate, then the result through the output mirror
        # (ngst_twave_set_output_mirror) or an explicit get_wavefront -- every call blocks, the copies of a step cannot overlap
        # its own tents much (all rows are needed by the first levels, most rows are final only in the last ones).
        use_mirror = not args.no_mirror and not args.no_sequential_mirror
        use_stream = args.streamed_input

        def e2e_step(mirror=None, stream=None):
            mirror = use_mirror if mirror is None else mirror
            stream = use_stream if stream is None else stream
            tw.reset_time()
            if stream:
                # spatial sweep (opt-in, measured slower than the blocking copy on this workload: DESIGN.md)
                check(lib.ngst_twave_set_boundary_values(tw._h, b_ptr, nb))
                check(lib.ngst_twave_set_initial_streamed(tw._h, in_ptr, ncomp))
            else:
                check(lib.ngst_twave_set_initial(tw._h, in_ptr, ncomp))           # H2D: state rows
                check(lib.ngst_twave_set_boundary_values(tw._h, b_ptr, nb))        # H2D: boundary samples
            check(lib.ngst_twave_propagate(tw._h))                                 # (+ D2H of final rows with the mirror)
            if not mirror:
                check(lib.ngst_twave_get_wavefront(tw._h, out_ptr))                # D2H: state rows

        # Pipelined form (headline): consecutive steps are independent problems, so the NEXT step's input is prefetched
        # (H2D on the input stream, copy engine) while this step's tents run, and this step's result leaves through an
        # asynchronous copy (D2H on the output stream) while the next step's input arrives -- PCIe is full duplex. Every step
        # still moves its own 794 MB in and 659 MB out inside the timed region; the first input and the last output are not
        # overlapped with anything and are part of the measured time.
        def e2e_pipelined(nsteps):
            check(lib.ngst_twave_prefetch_input(tw._h, in_ptr, ncomp, b_ptr, nb))
            for k in range(nsteps):
                tw.reset_time()
                check(lib.ngst_twave_use_prefetched(tw._h))
                if k + 1 < nsteps:
                    check(lib.ngst_twave_prefetch_input(tw._h, in_ptr, ncomp, b_ptr, nb))
                check(lib.ngst_twave_propagate(tw._h))
                check(lib.ngst_twave_get_wavefront_async(tw._h, out_ptr))
            check(lib.ngst_twave_wait_output(tw._h))

        e2e_step(False, False)               # reference result: blocking copy in, explicit copy out
        want = pin_out.numpy().copy()
        if use_stream:
            pin_out.zero_()
            e2e_step(False, True)
            if not np.array_equal(pin_out.numpy(), want):
                print("bench.py: streamed input differs from ngst_twave_set_initial; timing the blocking copy", file=sys.stderr)
                use_stream = False
        if use_mirror:
            pin_out.zero_()
            check(lib.ngst_twave_set_output_mirror(tw._h, out_ptr))
            e2e_step(True)
            if not np.array_equal(pin_out.numpy(), want):
                # never observed; do not time a path whose result is wrong: go back to the explicit copy and say so
                print("bench.py: output mirror differs from ngst_twave_get_wavefront; timing the explicit copy",
                      file=sys.stderr)
                check(lib.ngst_twave_set_output_mirror(tw._h, None))
                use_mirror = False
        for _ in range(args.warmup):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        barrier()
        seq_s = time.perf_counter() - t0
        if use_mirror:
            check(lib.ngst_twave_set_output_mirror(tw._h, None))
        e2e = {"sec": seq_s, "h2d": int(nbytes + nb * 8), "d2h": int(nbytes), "mirror": use_mirror, "stream": use_stream,
               "pipelined": False, "seq_sec": seq_s}
        if not args.no_pipelined_e2e:
            pin_out.zero_()
            e2e_pipelined(2)
            if np.array_equal(pin_out.numpy(), want):
                e2e_pipelined(max(args.warmup, 1))
                barrier()
                t0 = time.perf_counter()
                e2e_pipelined(args.steps)
                barrier()
                e2e["sec"] = time.perf_counter() - t0
                e2e["pipelined"] = True
            else:
                print("bench.py: pipelined steps differ from the sequential result; reporting the sequential figure", file=sys.stderr)
        del want

    # ---------------- informational: the same call sequence with initial and boundary data sampled ON THE DEVICE
    # (ngst_twave_*_analytic): no H2D of samples, same D2H. Not the e2e figure (that one copies
    # host inputs every step); reported beside it, with the difference of the resulting state to the host-data run.
    device_data = None
    if e2e is not None and world == 1 and not args.no_device_data:
        try:
            data = ng.StandingWave(D, 1.0)
            host_out = pin_out.numpy().copy()

            def dd_step():
                tw.reset_time()
                tw.SetInitial(data)
                tw.SetBoundaryCF(data)
                check(lib.ngst_twave_propagate(tw._h))
                if not use_mirror:
                    check(lib.ngst_twave_get_wavefront(tw._h, out_ptr))

            if use_mirror:
                tw.SetInitial(data)
                check(lib.ngst_twave_set_output_mirror(tw._h, out_ptr))
            for _ in range(args.warmup):
                dd_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                dd_step()
            barrier()
            dd_s = time.perf_counter() - t0
            diff = float(np.abs(pin_out.numpy() - host_out).max() / np.abs(host_out).max())
            device_data = {"value": len(mine) * args.steps / dd_s, "unit": "tents/s", "ms_per_step": dd_s / args.steps * 1e3,
                           "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(nbytes),
                           "max_rel_diff_vs_host_data": diff,
                           "path": "ngst_twave_set_initial_analytic + set_boundary_analytic + propagate + "
                                   + ("output mirror" if use_mirror else "get_wavefront")}
        except Exception as ex:  # informational arm: never takes the bench line down
            device_data = {"error": repr(ex)}
        finally:
            if use_mirror:
                lib.ngst_twave_set_output_mirror(tw._h, None)

    # ---------------- N>1: the partitioned run must equal a single-GPU run of the SAME slab bit for bit (same tents, same
    # arithmetic per tent). Both runs take initial and boundary data sampled on the device (bitwise reproducible per
    # point); rank 0 runs the whole slab on its GPU and broadcasts one checksum pair per global row.
    parity = None
    if world > 1 and not args.no_parity_check:
        try:
            nslabs = 2
            data = ng.StandingWave(D, 1.0)
            local_err = None
            try:  # collectives below must be reached by every rank
                tw.reset_time()
                tw.SetInitial(data)
                tw.SetBoundaryCF(data)
                for _ in range(nslabs):
                    tw.Propagate()
                mine_rows = tw.GetWavefront().view(np.uint64)
            except Exception as ex:
                local_err = repr(ex)
            okf = torch.tensor([0.0 if local_err else 1.0], dtype=torch.float64, device="cuda")
            dist.all_reduce(okf, op=dist.ReduceOp.MIN)
            if float(okf[0]) != 1.0:
                raise RuntimeError(f"partitioned run with device-sampled data failed on some rank (this rank: {local_err})")
            en_part = tw.Energy()
            nel_glob = int(mesh.ne)
            ref = torch.zeros((2, nel_glob), dtype=torch.int64, device="cuda")
            en_ref = torch.zeros(2, dtype=torch.float64, device="cuda")  # [energy, 1 = the single-GPU run succeeded]
            t_ref, ref_err = 0.0, None
            if rank == 0:
                try:  # every rank must reach the broadcasts below, whatever happens here
                    t0 = time.perf_counter()
                    one = ng.TWave(order, ts, 1.0, device=local)
                    one.SetInitial(data)
                    one.SetBoundaryCF(data)
                    for _ in range(nslabs):
                        one.Propagate()
                    rows = one.GetWavefront().view(np.uint64)
                    en_ref[0] = one.Energy()
                    ref[0] = torch.from_numpy(rows.sum(axis=1, dtype=np.uint64).view(np.int64)).cuda()
                    ref[1] = torch.from_numpy(np.bitwise_xor.reduce(rows, axis=1).view(np.int64)).cuda()
                    en_ref[1] = 1.0
                    del one, rows
                    t_ref = time.perf_counter() - t0
                except Exception as ex:
                    ref_err = repr(ex)
            dist.broadcast(ref, 0)
            dist.broadcast(en_ref, 0)
            if float(en_ref[1]) != 1.0:
                raise RuntimeError(f"single-GPU reference run on rank 0 failed: {ref_err}")
            refh = ref.cpu().numpy().view(np.uint64)
            l2g = tw.loc2glob
            bad = int(np.count_nonzero((mine_rows.sum(axis=1, dtype=np.uint64) != refh[0][l2g]) |
                                       (np.bitwise_xor.reduce(mine_rows, axis=1) != refh[1][l2g])))
            cnt = torch.tensor([bad, len(l2g)], dtype=torch.int64, device="cuda")
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
            e1 = float(en_ref[0])
            parity = {"ok": bool(int(cnt[0]) == 0 and abs(en_part - e1) <= 1e-12 * abs(e1)),
                      "kind": "bitwise: every local row (owned and ghost) of every rank vs a single-GPU run of the same "
                              "slab on rank 0 (64-bit sum and xor of each row's bit patterns)",
                      "slabs": nslabs, "rows_compared": int(cnt[1]), "rows_differing": int(cnt[0]),
                      "global_rows": nel_glob, "energy_partitioned": en_part, "energy_single_gpu": e1,
                      "single_gpu_run_s": t_ref, "data": "standing wave sampled on the device (initial + Dirichlet datum)"}
            del mine_rows
        except Exception as ex:  # a failed check is reported, never hidden
            parity = {"ok": False, "error": repr(ex)}

    # ---------------- reduce over ranks (max time, sum units)
    ntents_local = len(mine)
    stats = torch.tensor([dev_s, wall_s, e2e["sec"] if e2e else 0.0, float(ntents_local), flops_step, bytes_step,
                          float(launches), float(nels.sum()), cold["sec"] if cold else 0.0,
                          e2e["seq_sec"] if e2e else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    else:
        mx, sm = stats, stats
    mx, sm = mx.cpu().numpy(), sm.cpu().numpy()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    dev_s, wall_s, e2e_s = mx[0], mx[1], mx[2]
    seq_s_max = mx[9]
    tents_step = sm[3]
    value = tents_step * args.steps / dev_s

    # ---------------- roofline of the dominant kernel (the per-level tent kernel; every launch in the step is it)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peaks = json.load(fh)
    except OSError:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    dfma = _capi.C.c_double(0)
    dmma = _capi.C.c_double(0)
    check(lib.ngst_device_fp64_peak(local, 0, _capi.C.byref(dfma)))
    check(lib.ngst_device_fp64_peak(local, 1, _capi.C.byref(dmma)))
    mixed = _capi.C.c_double(0)
    check(lib.ngst_device_fp64_peak(local, 2, _capi.C.byref(mixed)))
    fp64_peak = max(dfma.value, dmma.value)
    flops_rank0 = flops_step * args.steps
    ach_tf = flops_rank0 / (float(np.sum(dev_ms)) * 1e-3) * 1e-12
    ach_gbs = bytes_step * args.steps / (float(np.sum(dev_ms)) * 1e-3) * 1e-9
    dev_sec0 = float(np.sum(dev_ms)) * 1e-3          # rank 0's device time of the timed steps
    te_rank0 = float(nels.sum()) * args.steps         # tent-elements rank 0 processed in them
    launches_rank0 = max(launches, 1)
    # Executed FP64 work per tent-element of the kernel variant, from ncu counters of ONE capture of this workload
    # (profiles/executed_flops.json, written by scripts/summarize_ncu.py: DMMA warp instructions x 512 + 2 x DFMA + DMUL +
    # DADD thread instructions, divided by the tent-elements of the captured launches), and the DRAM bytes of the same
    # capture. Falls back to the tile-count model of the DMMA work alone when the file does not describe this variant.
    nt_tiles = (N + 7) // 8
    n_rule = _capi.C.c_int(0)
    check(lib.ngst_compact_rule(D, 2 * order - 2, _capi.C.byref(n_rule), None, None))
    Qt = n_rule.value   # top-face Gram: compact rule exact to degree 2p-2 (24 points at D=3, p=4; tent_fast.cu Cfg::QT)
    check(lib.ngst_compact_rule(D, 2 * order, _capi.C.byref(n_rule), None, None))
    Qs = n_rule.value   # stabilisation Gram: compact rule exact to degree 2p where the table has one
    dmma_model = (Qt * (D + 1) / 4.0 + (0 if tw.fosystem else Qs / 4.0)) * (nt_tiles * (nt_tiles + 1) / 2) * 512.0
    # algorithmic bytes of the apply variant: it also streams the stored inverse of every tent (8 N NP bytes)
    inv_bytes_step = float(len(mine)) * 8.0 * N * ((N + 7) // 8 * 8)

    def roofline_of(variant, dev_sec, nlaunch):
        """variant: 'assemble' (every tent system assembled and solved: DMMA Grams + evaluation) or 'apply' (factor cache:
        right-hand side, stored inverse, evaluation -- no tensor-path work)."""
        exe, exe_src, traffic, ncu = None, None, None, None
        fname = "executed_flops.json" if variant == "assemble" else "executed_flops_apply.json"
        try:
            with open(os.path.join(ROOT, "profiles", fname)) as fh:
                rec = json.load(fh)
            if rec.get("kernel") == tw.kernel_name() and rec.get("D") == D and rec.get("order") == order:
                exe = rec["flops_per_tent_element"]
                exe_src = f"ncu counters, profiles/{fname} ({rec.get('tag')}): {rec.get('formula')}"
                traffic = rec.get("dram_bytes_per_tent_element")
                ncu = {k: rec.get(k) for k in ("pipe_shared_pct", "dmma_pipe_pct", "fp64_pipe_pct", "dmma_flops_per_tent_element",
                                               "dfma_flops_per_tent_element", "captured_grid_sizes", "tag")}
        except (OSError, ValueError, KeyError):
            exe = None
        if exe is None and variant == "assemble":
            exe, exe_src = dmma_model, "model: DMMA tiles only (no ncu capture of this kernel variant under profiles/)"
        nlaunch = max(nlaunch, 1)
        te_per_launch = te_rank0 / nlaunch
        exe_tf = None if exe is None else exe * te_rank0 / dev_sec * 1e-12
        # the apply variant issues no tensor-path instruction: its ceiling is the measured DFMA rate, not the (higher) DMMA one
        peak = dfma.value if variant == "apply" else fp64_peak
        alg_bytes = bytes_step + (inv_bytes_step if variant == "apply" else 0.0)
        gbs = alg_bytes * args.steps / dev_sec * 1e-9
        alg_tf = flops_rank0 / dev_sec * 1e-12
        return {
            "bound": "tensor", "unit": "TFLOP/s", "achieved": exe_tf, "peak": peak,
            "frac": None if exe_tf is None else exe_tf / peak, "kernel": tw.kernel_name(), "variant": variant,
            "peak_used": "DFMA pipe" if variant == "apply" else "max(DFMA, DMMA)",
            "pipe": "FP64: DMMA m8n8k4 (tensor path) and DFMA/DMUL/DADD share one datapath per SM (sm__pipe_shared)",
            "achieved_definition": "EXECUTED FP64 flops per tent-element (ncu counters) x tent-elements of the timed steps / "
                                   "CUDA-event time",
            "executed_flops_per_tent_element": exe, "executed_source": exe_src,
            "peak_source": f"measured live on this GPU (ngst_device_fp64_peak): DFMA {dfma.value:.2f} / DMMA m8n8k4 "
                           f"{dmma.value:.2f} TFLOP/s, interleaved 4:1 {mixed.value:.2f}; MEASURED_PEAKS.json holds no FP64 figure",
            "traffic": None if traffic is None else traffic * te_per_launch,
            "algorithmic_bytes_per_launch": alg_bytes * args.steps / nlaunch,
            "ncu": ncu, "launches_per_step": nlaunch / max(args.steps, 1),
            # SURVEY 8.4's count (order-2p rule everywhere, full N x N Grams, LU) over the same time: how much faster this is
            # than executing the reference's algorithm at the FP64 peak would be
            "algorithmic": {"flops_per_tent_element": f_el, "achieved": alg_tf, "unit": "TFLOP/s",
                            "algorithmic_speedup": alg_tf / fp64_peak},
            "hbm": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                    "peak_source": hbm_src, "bytes_per_tent_element": b_el,
                    "stored_inverse_bytes_per_tent": 8 * N * ((N + 7) // 8 * 8) if variant == "apply" else 0},
        }

    roofline = roofline_of(kernel_variant if kernel_variant != "other" else "assemble", dev_sec0, launches)
    assembly_every_step = None
    if cold:
        cold_s = mx[8]
        assembly_every_step = {"value": tents_step * args.steps / cold_s, "unit": "tents/s",
                               "ms_per_step": 1e3 * cold_s / args.steps,
                               "what": "factor cache off: every tent system assembled (DMMA Grams) and solved in every "
                                       "Propagate, as the reference does",
                               "roofline": roofline_of("assemble", cold["sec"], cold["launches"])}

    line = {
        "metric": METRIC, "value": value, "unit": "tents/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "dof_updates_per_s": value * N,
        "config": {"workload": f"3+1D unit-cube brick {cells[0]}x{cells[1]}x{cells[2]} cells x 6 tets, order {order}, "
                               f"c=1, standing wave, Dirichlet, dt={args.dt}, edge pitching local_ct global_ct=2/3",
                   "tets": int(np.prod(cells) * 6), "order": order, "nbasis": N, "nip": Q,
                   "slab_tents": int(ntents_global), "levels": ts.GetNLayers(), "tent_elements": int(sm[7]),
                   "parallelism": (f"spatial partition {grid[0]}x{grid[1]}x{grid[2]} ({args.partition} vertex ownership), " +
                                   ("interface rows exchanged inside the tent kernel over NVLink peer memory, one launch per slab"
                                    if tw.peer_exchange()[0] else
                                    f"per-level NCCL halo exchange {'after' if args.no_overlap else 'overlapped with'} the interior tents"))
                                  if world > 1 else "single GPU",
                   "l2": f"state {wf0.nbytes * world / 1e6:.0f} MB per slab > 126 MB L2 (no flush needed)",
                   "pitch_s": pitch_s, "pitch_level_sync": pitch_ls,
                   "exchange": (tw.peer_exchange()[1] if tw.peer_exchange()[0] else "per-level NCCL send/recv ("
                                + tw.peer_exchange()[1] + ")") if world > 1 else None,
                   "numa_binding_rank0": numa,
                   "compact_state": bool(tw.compact_state()),
                   "factor_cache": ("on: steady state of a time loop -- the inverse of every tent matrix was formed by the first "
                                    "Propagate of this slab geometry (warm-up) and is applied by the timed ones; "
                                    "`assembly_every_step` is the same measurement with the cache off") if cache_state == 2 else "off",
                   "factor_cache_bytes": int(cache_bytes)},
        "wall_ms_per_step": 1e3 * wall_s / args.steps,
        "gpu_launches": int(sm[6]),
        "clocks": clocks,
        "roofline": roofline,
        "energy_after": energy,
    }
    if e2e:
        seq_path = (("ngst_twave_set_initial_streamed" if e2e.get("stream") else "ngst_twave_set_initial") +
                    " + set_boundary_values + propagate" +
                    (" with the result rows streamed into the host buffer during propagate (ngst_twave_set_output_mirror)"
                     if e2e.get("mirror") else " + get_wavefront") + ", every call blocking, pinned host buffers")
        seq = {"value": tents_step * args.steps / seq_s_max, "unit": "tents/s", "ms_per_step": 1e3 * seq_s_max / args.steps,
               "path": seq_path}
        line["e2e"] = {"value": tents_step * args.steps / e2e_s, "unit": "tents/s",
                       "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"],
                       "ms_per_step": 1e3 * e2e_s / args.steps, "pipelined": bool(e2e.get("pipelined")),
                       "path": ("independent steps pipelined over the copy engines through the C ABI, pinned host buffers: per step "
                                "ngst_twave_use_prefetched (input staged by the previous step's ngst_twave_prefetch_input: H2D of "
                                "state rows + boundary samples on the input stream, compact transfer) + ngst_twave_prefetch_input of the "
                                "next step + ngst_twave_propagate + ngst_twave_get_wavefront_async (device snapshot, D2H on the output "
                                "stream); ngst_twave_wait_output after the last step; the unoverlapped first input and last output are "
                                "inside the timed region; result checked bitwise against the sequential calls before timing")
                               if e2e.get("pipelined") else seq_path,
                       "sequential": seq}
    if assembly_every_step:
        line["assembly_every_step"] = assembly_every_step
    if device_data:
        line["device_data"] = device_data
    if parity is not None:
        line["parity_check"] = parity
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        sample = args.cpu_sample_tents or 4000
        r = cpu_sample(args, cells, sample, 1, 1, cores)
        line["cpu_baseline"] = cpu_baseline_entry(r, cores, "1 timed pass after 1 warm-up")
    _RESULT.append(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    # stdout carries exactly ONE JSON line: anything libraries print on fd 1 meanwhile (e.g. "NCCL version ...") goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if _RESULT:
        print(_RESULT[0], flush=True)


if __name__ == "__main__":
    main()
