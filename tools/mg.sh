#!/bin/bash
# tools/mg.sh N <bench args>: bench.py on N GPUs the way the driver launches it
N=$1; shift
if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@" 2>&1 | tail -1
else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" 2>&1 | grep -E '^\{|Error|error' | tail -2; fi
