#!/bin/bash
# round 2: GPU tests (group, tile mask) + config 3 with / without the screen-space mask
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_run6_tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/r2_run6_tests.log
for m in 1 0; do
  VR_TILE_MASK=$m timeout 300 python bench.py --workload c3 --steps 90 --no-cpu-baseline --no-partitioned > gpurun_out/r2_c3_mask$m.json 2> gpurun_out/r2_c3_mask$m.err; echo "c3 mask=$m exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_c3_mask$m.json").read().strip().splitlines()[-1]); print("mask=$m", d["ms_per_step"], d["value"], d.get("roofline",{}).get("frac"))
except Exception as e: print("parse", e); print(open("gpurun_out/r2_c3_mask$m.err").read()[-1500:])
PY
done
