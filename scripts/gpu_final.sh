#!/bin/bash
# Final visit of a round on one B200: GPU tests, smoke, bench (both arms), then the ncu evidence for both kernel variants
# (each ncu pass only after the same command has exited 0 without ncu; scripts/gpu_profile.sh).
# usage: scripts/gpu_final.sh [TAG]
set -u
TAG=${1:-r05}
mkdir -p gpurun_out
nvidia-smi -L; nproc
python -m pytest tests -m gpu -q -s > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/${TAG}_bench_ref.json
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
scripts/gpu_profile.sh ${TAG}_apply
scripts/gpu_profile.sh ${TAG}_assemble --no-factor-cache
