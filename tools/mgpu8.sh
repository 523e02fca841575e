#!/bin/bash
# tools/mgpu8.sh: full-size multi-GPU lines on 8 GPUs (run under gpurun --gpus 8)
O=gpurun_out/mg8; mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -5) > $O/pytest_multi_n8.log
tools/mg.sh 8 --no-cpu-baseline > $O/bench_c2_n8.json
tools/mg.sh 8 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n8.json
tools/mg.sh 8 --workload c5 --steps 36 --warmup 3 > $O/bench_c5_n8.json
tools/mg.sh 4 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n4.json
tools/mg.sh 2 --workload c4 --steps 36 --warmup 3 > $O/bench_c4_n2.json
tail -2 $O/pytest_multi_n8.log
for f in $O/bench_*.json; do python - <<PY
import json
try:
    d=json.loads(open("$f").read().strip().splitlines()[-1])
    print("$f".split("/")[-1], d['n_gpus'], round(d['ms_per_step'],4),'ms value',f"{d['value']:.3e}", d.get('roofline',{}).get('frac'), d['config'].get('workload','')[:50])
except Exception as e: print("$f", "ERR", e, open("$f").read()[-400:])
PY
done
