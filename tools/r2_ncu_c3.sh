#!/bin/bash
# ncu --set full of one k_march launch of config 3 (and of the mask kernel), after the same command exited 0 without ncu
O=gpurun_out/r2_ncu_c3; mkdir -p $O
CMD="python bench.py --workload c3 --steps 4 --warmup 3 --no-cpu-baseline --no-partitioned"
timeout 300 $CMD > $O/bench_plain.json 2> $O/bench_plain.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_march -s 6 -c 1 -o $O/c3_march -f $CMD > $O/ncu_march.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tile_mask -s 6 -c 1 -o $O/c3_mask -f $CMD > $O/ncu_mask.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches.csv $CMD > $O/ncu_list.log 2>&1
echo "exit $?"; ls -la $O
