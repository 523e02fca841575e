#!/bin/bash
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for l in 3d paired; do for b in 2 4; do tools/exp.sh "VR_BATCH=$b" --layout $l; done; done
for l in 3d paired; do tools/exp.sh "VR_BENCH_N=256" --layout $l;  tools/exp.sh "X=1" --layout $l --workload c3 --steps 90 --no-ess; tools/exp.sh "X=1" --layout $l --workload c1; done
tools/exp.sh "VR_CTA_THREADS=256" --layout paired
tools/exp.sh "X=1" --layout paired --no-ert
tools/exp.sh "X=1" --layout paired --in-flight 2
tools/exp.sh "X=1" --layout paired --in-flight 1
} > gpurun_out/exp3.txt 2>&1
cat gpurun_out/exp3.txt
