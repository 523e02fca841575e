#!/bin/bash
O=gpurun_out/r2_exp2; mkdir -p $O
timeout 120 tools/exp_alias.bin 512 > $O/exp_alias_512.txt 2>&1; echo "exit $?" >> $O/exp_alias_512.txt
timeout 300 tools/exp_depth.bin 512 > $O/exp_depth_512.txt 2>&1; echo "exit $?" >> $O/exp_depth_512.txt
head -20 $O/exp_alias_512.txt; cat $O/exp_depth_512.txt
