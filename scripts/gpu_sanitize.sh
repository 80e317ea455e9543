#!/bin/bash
# One compute-sanitizer tool per gpurun call (B200_PROFILING.md), on small cases that reach every kernel variant.
# usage: scripts/gpu_sanitize.sh [memcheck|racecheck|synccheck|initcheck] [TAG]
set -u
TOOL=${1:-memcheck}
TAG=${2:-r04}
mkdir -p gpurun_out
timeout 200 python scripts/sanitize_cases.py > gpurun_out/${TAG}_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_sanitize_plain.log; exit 1; }
timeout 1200 compute-sanitizer --tool $TOOL --error-exitcode 3 python scripts/sanitize_cases.py > gpurun_out/${TAG}_sanitize_${TOOL}.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|SANITIZE_CASES_OK|RACECHECK SUMMARY" gpurun_out/${TAG}_sanitize_${TOOL}.log | tail -5
