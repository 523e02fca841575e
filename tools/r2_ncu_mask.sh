#!/bin/bash
O=gpurun_out/r2_ncu_c3; mkdir -p $O
CMD="python bench.py --workload c3 --steps 4 --warmup 3 --no-cpu-baseline --no-partitioned"
timeout 300 $CMD > $O/bench_plain.json 2> $O/bench_plain.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tile_mask -s 6 -c 1 -o $O/c3_mask2 -f $CMD > $O/ncu_mask.log 2>&1
echo "exit $?"
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "tile_mask or non_finite" 2>&1 | tail -3
