#!/bin/bash
# Multi-GPU visit: parity tests for the world sizes this box has, then bench lines (A/B of partition and overlap).
# usage: scripts/gpu_multi.sh TAG N [ab]
set -u
TAG=$1; N=$2; AB=${3:-}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/${TAG}_pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest_multi.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 "$@"; }
run > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/${TAG}_bench_n${N}.json
if [ -n "$AB" ]; then
  run --no-e2e --no-overlap > gpurun_out/${TAG}_bench_n${N}_nooverlap.json 2>> gpurun_out/${TAG}_bench_n${N}.err; cut -c1-200 gpurun_out/${TAG}_bench_n${N}_nooverlap.json
  run --no-e2e --partition brick > gpurun_out/${TAG}_bench_n${N}_brick.json 2>> gpurun_out/${TAG}_bench_n${N}.err; cut -c1-200 gpurun_out/${TAG}_bench_n${N}_brick.json
  run --no-e2e --partition brick --no-overlap > gpurun_out/${TAG}_bench_n${N}_brick_nooverlap.json 2>> gpurun_out/${TAG}_bench_n${N}.err; cut -c1-200 gpurun_out/${TAG}_bench_n${N}_brick_nooverlap.json
fi
tail -3 gpurun_out/${TAG}_bench_n${N}.err
