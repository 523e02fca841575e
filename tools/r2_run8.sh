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
run b4g8 A=1
run b2g8 VR_BATCH_ESS=2
run b4g4 VR_ESS_GUARD=4
run b4g16 VR_ESS_GUARD=16
run b4g2 VR_ESS_GUARD=2
run b4nomask VR_TILE_MASK=0
run b4global VR_TILE_MASK=2
run b4ww4 VR_WARP_W=4
run b4ww16 VR_WARP_W=16
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -3
