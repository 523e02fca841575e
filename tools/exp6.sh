#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
tools/exp.sh "X=1"
tools/exp.sh "VR_PAIR_AXIS=2"
tools/exp.sh "VR_PAIR_AXIS=1"
tools/exp.sh "VR_PAIR_AXIS=0"
tools/exp.sh "VR_BATCH=4"
tools/exp.sh "VR_BENCH_N=256"
tools/exp.sh "X=1" --workload c3 --steps 90
tools/exp.sh "VR_PAIR_POLICY=1" --workload c3 --steps 90
tools/exp.sh "X=1" --workload c1
tools/exp.sh "VR_PAIR_POLICY=1" --workload c1
} > gpurun_out/exp6.txt 2>&1
cat gpurun_out/exp6.txt
