#!/bin/bash
# Short GPU-box visit: device-sampling tests, then the ncu launch list and one full capture of the tent kernel at HEAD
# (each ncu pass only after the same command has exited 0 without ncu).
# usage: scripts/gpu_ncu.sh [TAG]
set -u
TAG=${1:-r03b}
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_analytic.py -x -q > gpurun_out/${TAG}_pytest_analytic.log 2>&1; echo "analytic rc=$?"; tail -8 gpurun_out/${TAG}_pytest_analytic.log
SMALL="python bench.py --cells 12 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --level-launches"
timeout 60 $SMALL > gpurun_out/${TAG}_plain_small.log 2>&1 && \
timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $SMALL > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 100 ncu --set full --clock-control none --import-source on -k regex:tent_kernel -s 60 -c 2 -f -o gpurun_out/${TAG}_prof_tent $SMALL > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | tail -8
