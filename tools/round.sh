#!/bin/bash
# tools/round.sh TAG: everything the profiles/ directory holds for one kernel version (run under gpurun)
TAG=${1:-vX}
O=gpurun_out/$TAG; mkdir -p $O
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > $O/pytest_gpu.log
python bench.py > $O/bench_c2.json 2> $O/bench_c2.err
python bench.py --workload c3 --steps 90 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python bench.py --workload c1 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err
python bench.py --layout 3d --no-cpu-baseline > $O/bench_c2_3d.json 2>/dev/null
python bench.py --impl reference --steps 8 --warmup 1 > $O/bench_c2_reference.json 2> $O/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march -s 12 -c 1 -f -o $O/march_c2 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march -s 12 -c 1 -f -o $O/march_c3 python bench.py --workload c3 --steps 4 --warmup 3 --no-cpu-baseline > $O/ncu_c3.log 2>&1
tail -3 $O/pytest_gpu.log
for w in c2 c3 c1 c2_3d c2_reference; do python - <<PY
import json
try:
    d=json.load(open("$O/bench_$w.json"))
    print("$w", round(d.get('ms_per_step',0),4),'ms value',f"{d['value']:.3e}", 'e2e', f"{d['e2e']['value']:.3e}", 'tex', round(d.get('roofline_tex',{}).get('frac',0),3), 'hbm', round(d.get('roofline',{}).get('frac',0),3), d.get('config',{}).get('layout'))
except Exception as e: print("$w", "ERR", e)
PY
done
