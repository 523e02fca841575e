#!/bin/bash
# tools/r2_sl.sh N TAG: sort-last on 2048^3 (fits one GPU) with 1..4 slots
N=${1:-2}; TAG=${2:-r2_sl}; O=gpurun_out/$TAG; mkdir -p $O
for s in 1 2 3 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --workload c5 --size 2048 --dense --slots-last $s --steps 36 --warmup 3 > $O/sl_n${N}_slots$s.json 2> $O/sl_n${N}_slots$s.err
  python -c "
import json; d=json.loads(open('$O/sl_n${N}_slots$s.json').read().strip().splitlines()[-1]); print('N=$N slots=$s', round(d['ms_per_step'],4), 'ms', d['frame_checksum'])"
done
