#!/bin/bash
# tools/sweep2.sh: one-knob-at-a-time sweep of K1's launch knobs on C2, current kernel (run under gpurun)
run() { r=$(env "$@" python bench.py --steps 180 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), 'fetch-only', round(d['roofline_tex']['fetch_only_ms_per_step'],4))"); echo "$* : $r"; }
run VR_NONE=1
for v in 0 10 25 50; do run VR_CARVEOUT=$v; done
for v in 4 16; do run VR_WARP_W=$v; done
run VR_CTA_THREADS=256
run VR_BATCH=4
run VR_BATCH=4 VR_CTA_THREADS=256
for v in 32 64 128; do run VR_L2_GRAN=$v; done
