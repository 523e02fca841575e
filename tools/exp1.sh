#!/bin/bash
# experiment bundle 1 (run under gpurun): fetch rates of candidate layouts, K1 vs volume size, K1 vs CTA shape
mkdir -p gpurun_out
{
./tools/exp_texrate.bin
for n in 32 64 128 256 384 512; do tools/exp.sh "VR_BENCH_N=$n" --no-ess; done
for t in 128 256; do for w in 4 8 16; do for b in 2 4; do tools/exp.sh "VR_CTA_THREADS=$t VR_WARP_W=$w VR_BATCH=$b" --no-ess; done; done; done
tools/exp.sh "VR_BBOX=0" --no-ess
tools/exp.sh "X=1" --no-ess --in-flight 1
tools/exp.sh "X=1" --no-ess --in-flight 3
tools/exp.sh "X=1" --no-ess --in-flight 4
tools/exp.sh "X=1" --no-ess --no-ert
} > gpurun_out/exp1.txt 2>&1
cat gpurun_out/exp1.txt
