#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
tools/exp.sh "X=1" --workload c3 --steps 90
tools/exp.sh "VR_PAIR_POLICY=1" --workload c3 --steps 90
tools/exp.sh "VR_BATCH=4" --workload c3 --steps 90
tools/exp.sh "VR_CTA_THREADS=256" --workload c3 --steps 90
tools/exp.sh "VR_ESS_MIN_EMPTY=0" 
tools/exp.sh "X=1"
} > gpurun_out/exp5.txt 2>&1
cat gpurun_out/exp5.txt
