#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for l in 3d paired; do tools/exp.sh "X=1" --layout $l; tools/exp.sh "VR_BENCH_N=256" --layout $l; tools/exp.sh "X=1" --layout $l --workload c3 --steps 90; tools/exp.sh "X=1" --layout $l --workload c1; done
tools/exp.sh "VR_CTA_THREADS=256" --layout paired
} > gpurun_out/exp4.txt 2>&1
cat gpurun_out/exp4.txt
