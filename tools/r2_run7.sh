#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_run7_tests.log 2>&1; echo "tests exit $?"; tail -5 gpurun_out/r2_run7_tests.log
for v in "1 2" "0 2" "1 4"; do
  set -- $v
  VR_TILE_MASK=$1 VR_BATCH=$2 timeout 300 python bench.py --workload c3 --steps 90 --no-cpu-baseline --no-partitioned > gpurun_out/r2_c3_m$1_b$2.json 2> gpurun_out/r2_c3_m$1_b$2.err; echo "c3 mask=$1 batch=$2 exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_c3_m$1_b$2.json").read().strip().splitlines()[-1]); print("mask=$1 batch=$2", d["ms_per_step"], d.get("ms_per_step_in_flight_1"), d["e2e"]["ms_per_step"])
except Exception as e: print("parse", e); print(open("gpurun_out/r2_c3_m$1_b$2.err").read()[-1500:])
PY
done
python tools/exp_c3.py 2>&1 | tee gpurun_out/r2_exp_c3_v2.txt
