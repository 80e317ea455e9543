#!/bin/bash
# Short GPU-box visit: output-mirror tests, bench with the mirror, and A/B lines without it / with more copy CTAs.
# usage: scripts/gpu_mirror.sh [TAG]
set -u
TAG=${1:-r03c}
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_mirror.py -x -q > gpurun_out/${TAG}_pytest_mirror.log 2>&1; echo "mirror tests rc=$?"; tail -15 gpurun_out/${TAG}_pytest_mirror.log
timeout 150 python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
AB="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-device-data"
timeout 80 $AB --no-mirror > gpurun_out/${TAG}_bench_nomirror.json 2> gpurun_out/${TAG}_bench_nomirror.err; echo "nomirror rc=$?"
NGST_MIRROR_CTAS=48 timeout 80 $AB > gpurun_out/${TAG}_bench_ctas48.json 2> gpurun_out/${TAG}_bench_ctas48.err; echo "ctas48 rc=$?"
NGST_MIRROR_CTAS=12 timeout 80 $AB > gpurun_out/${TAG}_bench_ctas12.json 2> gpurun_out/${TAG}_bench_ctas12.err; echo "ctas12 rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r03c_bench*.json")):
    try:
        d = json.loads(open(f).read())
        print(f, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "e2e_ms", round(d["e2e"]["ms_per_step"], 2), "dd", d.get("device_data", {}).get("value"))
    except Exception as ex:
        print(f, "ERR", ex)
PY
