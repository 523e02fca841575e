#!/bin/bash
O=gpurun_out/mg8b; mkdir -p $O
tools/mg.sh 8 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n8.json
tools/mg.sh 8 --workload c5 --steps 36 --warmup 3 > $O/bench_c5_n8.json
tools/mg.sh 4 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n4.json
tools/mg.sh 2 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n2.json
tools/mg.sh 1 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n1.json
VR_PAIR_POLICY=0 tools/mg.sh 8 --workload c5 --steps 36 --warmup 3 > $O/bench_c5_n8_3d.json
for f in $O/bench_*.json; do python - <<PY
import json
try:
    d=json.loads(open("$f").read().strip().splitlines()[-1])
    print("$f".split("/")[-1], d['n_gpus'], round(d['ms_per_step'],4),'ms value',f"{d['value']:.3e}", round(d.get('roofline',{}).get('frac'),3))
except Exception as e: print("$f", "ERR", e, open("$f").read()[-400:])
PY
done
