#!/bin/bash
# tools/mgpu_round.sh N TAG: multi-GPU checks + bench lines on N GPUs (run under gpurun --gpus N)
N=${1:-2}; TAG=${2:-mg}; O=gpurun_out/$TAG; mkdir -p $O
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > $O/pytest_gpu_n$N.log
python __graft_entry__.py smoke > $O/smoke.log 2>&1
tools/mg.sh 1 --no-cpu-baseline > $O/bench_c2_n1.json
tools/mg.sh $N --no-cpu-baseline > $O/bench_c2_n$N.json
tools/mg.sh $N --workload c4 --size 1024 --frame 3840x2160 --steps 36 --warmup 3 > $O/bench_c4s_n$N.json
tools/mg.sh 1 --workload c4 --size 1024 --frame 3840x2160 --steps 36 --warmup 3 > $O/bench_c4s_n1.json
tools/mg.sh $N --workload c5 --size 1024 --frame 1920x1080 --steps 36 --warmup 3 > $O/bench_c5s_n$N.json
tools/mg.sh $N --workload c5 --size 1024 --frame 1920x1080 --steps 36 --warmup 3 --merge nccl > $O/bench_c5s_nccl_n$N.json
tail -2 $O/pytest_gpu_n$N.log; tail -1 $O/smoke.log
for f in $O/bench_*.json; do python - <<PY
import json
try:
    d=json.loads(open("$f").read().strip().splitlines()[-1])
    print("$f".split("/")[-1], d['n_gpus'], round(d['ms_per_step'],4),'ms value',f"{d['value']:.3e}", 'e2e', f"{d['e2e']['value']:.3e}" if d.get('e2e') else None, d['config'].get('workload','')[:40])
except Exception as e: print("$f", "ERR", e, open("$f").read()[-300:])
PY
done
