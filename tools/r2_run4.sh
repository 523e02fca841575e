#!/bin/bash
# round 2, run 4: tests after the dirty clear / transactional upload; A/B of the register cap and the dirty clear
O=gpurun_out/r2_run4; mkdir -p $O
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > $O/pytest_gpu.log
python bench.py --no-cpu-baseline > $O/bench_c2_mb5.json 2> $O/bench_c2_mb5.err
VR_LIB_PATH=$PWD/tools/libvr_b200_mb1.so python bench.py --no-cpu-baseline > $O/bench_c2_mb1.json 2> $O/bench_c2_mb1.err
VR_DIRTY_CLEAR=0 python bench.py --no-cpu-baseline > $O/bench_c2_mb5_fullclear.json 2> $O/bench_c2_mb5_fullclear.err
python bench.py --no-cpu-baseline --in-flight 1 > $O/bench_c2_mb5_f1.json 2> /dev/null
VR_LIB_PATH=$PWD/tools/libvr_b200_mb1.so python bench.py --no-cpu-baseline --in-flight 1 > $O/bench_c2_mb1_f1.json 2> /dev/null
python bench.py --workload c3 --steps 90 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python bench.py --workload c1 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err
tail -4 $O/pytest_gpu.log
for f in $O/bench_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f"))
    print("$f".split('/')[-1], round(d.get('ms_per_step',0),4),'ms e2e', round(d['e2e']['ms_per_step'],4), 'fetch-only', d['roofline_tex'].get('fetch_only_ms_per_step'), 'tex', round(d.get('roofline_tex',{}).get('frac',0),3))
except Exception as e: print("$f", "ERR", e)
PY
done
