#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, then ncu launch list + one full capture of the tent kernel.
# usage: scripts/gpu_round.sh [TAG]   (outputs under gpurun_out/TAG_*)
set -u
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi -L; nproc
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
if [ "${REF:-1}" = "1" ]; then
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"; cat gpurun_out/${TAG}_bench_ref.json
fi
SMALL="python bench.py --cells 12 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --level-launches"
$SMALL > gpurun_out/${TAG}_plain_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $SMALL > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$SMALL > gpurun_out/${TAG}_plain_small2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tent_kernel -s 60 -c 3 -f -o gpurun_out/${TAG}_prof_tent $SMALL > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
if [ "${QT:-0}" = "1" ]; then
QTS="python scripts/bench_configs.py --config 3 --scale 16 --slabs 1"
$QTS > gpurun_out/${TAG}_qt_small.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tent_kernel_qt_fast -s 4 -c 2 -f -o gpurun_out/${TAG}_prof_qt $QTS > gpurun_out/${TAG}_ncu_qt.log 2>&1
echo "ncu qt rc=$?"
fi
