#!/bin/bash
mkdir -p gpurun_out
run() { # label env...
  L=$1; shift
  env "$@" timeout 300 python bench.py --workload c3 --steps 90 --no-cpu-baseline --no-partitioned > gpurun_out/r2_c3_$L.json 2> gpurun_out/r2_c3_$L.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_c3_$L.json").read().strip().splitlines()[-1]); print("$L", round(d["ms_per_step"],4), round(d.get("ms_per_step_in_flight_1") or 0,4), round(d["e2e"]["ms_per_step"],4))
except Exception as e: print("$L parse", e); print(open("gpurun_out/r2_c3_$L.err").read()[-800:])
PY
}
run new A=1
run prev VR_LIB_PATH=$PWD/tools/libvr_b200_prev.so
run new2 A=1
run prev2 VR_LIB_PATH=$PWD/tools/libvr_b200_prev.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_partition.py -x -q 2>&1 | tail -3
