#!/bin/bash
# GPU-box visit for the roofline evidence on the BENCHED workload (28^3 cells x 6 tets, order 4, one launch per level):
# ncu launch list of one slab, then one `--set full` capture (+ executed-FP64-op counters) of two mid-slab levels.
# Each ncu pass runs only after the same command line has exited 0 without ncu.
# usage: scripts/gpu_profile.sh [TAG] [extra bench.py flags]
set -u
TAG=${1:-r04a}
shift || true
mkdir -p gpurun_out
CMD="python bench.py --cells 28 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --level-launches $*"
FP64=sm__inst_executed_pipe_tensor_subpipe_dmma.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__sass_thread_inst_executed_op_dadd_pred_on.sum,sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg,sm__cycles_elapsed.avg.per_second
timeout 280 $CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# 3 warm-up slabs x (boundary sampling is not used here: 42 level launches each); levels 14 and 15 of the timed slab
timeout 280 $CMD > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --metrics $FP64 --clock-control none --import-source on -k regex:tent_kernel -s 140 -c 2 -f -o gpurun_out/${TAG}_prof_tent $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -3 gpurun_out/${TAG}_ncu_full.log
ls -la gpurun_out | tail -8
