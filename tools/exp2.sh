#!/bin/bash
# experiment bundle 2: paired layout vs 3-D array on c2 / c3 / c1
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for l in 3d paired; do for b in 2 4; do tools/exp.sh "VR_BATCH=$b" --layout $l; done; done
tools/exp.sh "VR_CTA_THREADS=256" --layout paired
tools/exp.sh "VR_WARP_W=4" --layout paired
tools/exp.sh "VR_WARP_W=16" --layout paired
for l in 3d paired; do tools/exp.sh "X=1" --layout $l --workload c3 --steps 90; tools/exp.sh "X=1" --layout $l --workload c3 --steps 90 --no-ess; tools/exp.sh "X=1" --layout $l --workload c1; done
for n in 128 256; do tools/exp.sh "VR_BENCH_N=$n" --layout paired; done
} > gpurun_out/exp2.txt 2>&1
cat gpurun_out/exp2.txt
