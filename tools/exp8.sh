#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
tools/exp.sh "X=1"
tools/exp.sh "X=1" --workload c3 --steps 90
tools/exp.sh "X=1" --workload c1
python tools/exp_axis.py 512 2>&1 | tail -1
for pol in 0 2; do
  r=$(VR_PAIR_POLICY=$pol tools/mg.sh 1 --workload c4 --steps 12 --warmup 2 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],3), d['value'])")
  echo "c4 2048^3 8K policy=$pol : $r"
done
nvidia-smi --query-gpu=memory.used --format=csv
} > gpurun_out/exp8.txt 2>&1
cat gpurun_out/exp8.txt
