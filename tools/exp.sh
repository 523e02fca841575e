#!/bin/bash
# one-line result of a bench run with the given env assignments: tools/exp.sh "VR_BATCH=8 VR_WARP_W=4" [bench args]
envs="$1"; shift
r=$(env $envs python bench.py --steps 120 --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), 'tex', round(d['roofline_tex']['frac'],3), 'hbm', round(d['roofline']['frac'],3), 'e2e', round(d['e2e']['ms_per_step'],3))")
echo "$envs $@ : $r"
