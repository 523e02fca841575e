#!/bin/bash
O=gpurun_out/mg8c; mkdir -p $O
tools/mg.sh 8 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n8.json
tools/mg.sh 8 --workload c4 --steps 36 --warmup 3 --slots 3 --tile-px 128 > $O/bench_c4_n8_s3_t128.json
for f in $O/bench_*.json; do python - <<PY
import json
try:
    d=json.loads(open("$f").read().strip().splitlines()[-1])
    print("$f".split("/")[-1], d['n_gpus'], round(d['ms_per_step'],4),'ms value',f"{d['value']:.3e}", round(d.get('roofline',{}).get('frac'),3))
except Exception as e: print("$f", "ERR", e, open("$f").read()[-400:])
PY
done
