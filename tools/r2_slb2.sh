#!/bin/bash
N=${1:-2}; TAG=${2:-r2_slb}; O=gpurun_out/$TAG; mkdir -p $O
run() { L=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --workload c5 --size 2048 --dense --steps 36 --warmup 3 $ARGS > $O/sl_n${N}_$L.json 2> $O/sl_n${N}_$L.err
  python -c "
import json; d=json.loads(open('$O/sl_n${N}_$L.json').read().strip().splitlines()[-1]); print('N=$N $L', round(d['ms_per_step'],4), 'ms', d['frame_checksum'], d['config'].get('layout'))" || tail -5 $O/sl_n${N}_$L.err
}
ARGS="--balance 0" run fixed A=1
ARGS="--balance 0.08" run frozen VR_BALANCE_FREEZE=1
ARGS="--balance 0.08" run moving A=1
ARGS="--balance 0.08" run moving_3d VR_PAIR_POLICY=0
ARGS="--balance 0" run fixed_3d VR_PAIR_POLICY=0
