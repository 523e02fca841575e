#!/bin/bash
O=gpurun_out/r2_exp3; mkdir -p $O
M=l1tex__t_sectors_pipe_tex_mem_texture_lookup_miss.sum,l1tex__t_sectors_pipe_tex_mem_texture.sum,l1tex__t_requests_pipe_tex_mem_texture.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum,l1tex__m_xbar2l1tex_read_bytes.sum,l1tex__data_pipe_tex_wavefronts.sum.pct_of_peak_sustained_elapsed,l1tex__f_wavefronts.sum.pct_of_peak_sustained_elapsed,l1tex__t_requests_pipe_tex_mem_texture.sum.pct_of_peak_sustained_elapsed
timeout 120 tools/exp_region.bin 512 > $O/exp_region.txt 2>&1 &&
ncu --metrics $M --clock-control none -k regex:k_dealt -s 2 -c 1 --csv --log-file $O/ncu_dealt.csv tools/exp_region.bin 512 0 > $O/ncu_dealt.log 2>&1
ncu --metrics $M --clock-control none -k regex:k_region -s 2 -c 1 --csv --log-file $O/ncu_region.csv tools/exp_region.bin 512 1 > $O/ncu_region.log 2>&1
cat $O/exp_region.txt; cut -d, -f 5,13- $O/ncu_dealt.csv | tail -12; cut -d, -f 5,13- $O/ncu_region.csv | tail -12
