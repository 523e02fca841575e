#!/bin/bash
mkdir -p gpurun_out
{
for pol in 0 2; do for ax in -1 2; do
  r=$(VR_PAIR_POLICY=$pol VR_PAIR_AXIS=$ax tools/mg.sh 1 --workload c4 --size 1024 --frame 7680x4320 --steps 36 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],3))")
  echo "c4 1024^3 8K policy=$pol axis=$ax : $r"
done; done
} > gpurun_out/exp7.txt 2>&1
cat gpurun_out/exp7.txt
