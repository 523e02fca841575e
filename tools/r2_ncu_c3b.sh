#!/bin/bash
O=gpurun_out/r2_ncu_c3; mkdir -p $O
CMD="python bench.py --workload c3 --steps 4 --warmup 3 --no-cpu-baseline --no-partitioned"
timeout 300 $CMD > $O/bench_plain.json 2> $O/bench_plain.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_march -s 14 -c 1 -o $O/c3_march -f $CMD > $O/ncu_march.log 2>&1
echo "exit $?"
