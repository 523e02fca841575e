#!/bin/bash
O=gpurun_out/dirty; mkdir -p $O
(timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -8) > $O/pytest.log; tail -3 $O/pytest.log
python bench.py --no-cpu-baseline > $O/bench_c2.json 2>$O/err_c2.txt
python bench.py --no-cpu-baseline --full-download > $O/bench_c2_full.json 2>$O/err_c2f.txt
python bench.py --no-cpu-baseline --workload c3 --steps 90 > $O/bench_c3.json 2>$O/err_c3.txt
python bench.py --no-cpu-baseline --workload c1 > $O/bench_c1.json 2>$O/err_c1.txt
python - <<PY
import json,glob
for f in sorted(glob.glob("$O/bench_*.json")):
    try:
        d=json.load(open(f)); e=d['e2e']; print(f.split('/')[-1], 'dev', round(d['ms_per_step'],4), 'e2e', round(e['ms_per_step'],4), 'd2h/step', int(e['d2h_bytes_per_step']), 'of', e['frame_bytes'])
    except Exception as ex: print(f, 'ERR', ex)
PY
cat $O/err_*.txt | tail -5
