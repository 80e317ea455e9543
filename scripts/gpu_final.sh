#!/bin/bash
# Short GPU-box visit (little budget left): new device-sampling tests first, then the bench line, host pitcher timing,
# and as much of the full GPU suite as fits.
# usage: scripts/gpu_final.sh [TAG]
set -u
TAG=${1:-r03a}
mkdir -p gpurun_out
nvidia-smi -L; nproc
timeout 150 python -m pytest tests/test_gpu_analytic.py -x -q > gpurun_out/${TAG}_pytest_analytic.log 2>&1; echo "analytic rc=$?"; tail -12 gpurun_out/${TAG}_pytest_analytic.log
timeout 200 python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
python - > gpurun_out/${TAG}_pitch.txt 2>&1 <<'PY'
import time, os
import ngstrefftz_b200 as ng
for n in (28, 56):
    mesh = ng.StructuredMesh([n] * 3, [0.0] * 3, [n / 28.0] * 3)
    ts = ng.TentSlab(mesh, "edge"); ts.SetMaxWavespeed(1.0)
    t0 = time.perf_counter(); assert ts.PitchTents(0.05, True, 2 / 3); t1 = time.perf_counter()
    print(f"cells {n}^3 tets {6*n**3} tents {ts.GetNTents()} levels {ts.GetNLayers()} pitch_s {t1-t0:.3f} cores {os.cpu_count()}", flush=True)
PY
cat gpurun_out/${TAG}_pitch.txt
timeout 240 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_analytic.py > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${TAG}_pytest_gpu.log
