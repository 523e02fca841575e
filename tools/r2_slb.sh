#!/bin/bash
# tools/r2_slb.sh N TAG: sort-last on 2048^3 (fits one GPU) with fixed and with camera-following split planes, + the multi-process check
N=${1:-2}; TAG=${2:-r2_slb}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tests/mgpu_check.py > $O/mgpu_check_n$N.log 2>&1; tail -2 $O/mgpu_check_n$N.log
for b in 0 0.08; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --workload c5 --size 2048 --dense --balance $b --steps 36 --warmup 3 > $O/sl_n${N}_balance$b.json 2> $O/sl_n${N}_balance$b.err
  python -c "
import json; d=json.loads(open('$O/sl_n${N}_balance$b.json').read().strip().splitlines()[-1]); print('N=$N balance=$b', round(d['ms_per_step'],4), 'ms', d['frame_checksum'])" || tail -5 $O/sl_n${N}_balance$b.err
done
