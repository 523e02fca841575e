#!/bin/bash
# tuning sweep of K1's knobs on C2 (run under gpurun); prints ms/step and tex fraction
for t in 256 128; do for l in 0 1; do for b in 2 4 8; do for e in "--no-ess" ""; do
  r=$(VR_CTA_THREADS=$t VR_LUT16=$l VR_BATCH=$b python bench.py --steps 120 --no-cpu-baseline $e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline_tex']['frac'],3), round(d['e2e']['ms_per_step'],3))")
  echo "threads=$t lut16=$l batch=$b ess=${e:-on} : $r"
done; done; done; done
