#!/bin/bash
# round 2, experiment 1: linear alias of array memory + L2 prefetch; DRAM-fed ceiling of the texture path
O=gpurun_out/r2_exp1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
timeout 300 tools/exp_alias.bin 512 > $O/exp_alias_512.txt 2>&1; echo "exit $?" >> $O/exp_alias_512.txt
timeout 300 python - > $O/tex_dram.txt 2>&1 <<'PY'
import volumetric_renderer_b200 as vr
for dt, name in ((vr.U16, "u16"), (vr.U8, "u8")):
    for mode, what in ((1, "L1-resident 3-D trilinear"), (2, "L1-resident paired bilinear"), (3, "DRAM-streamed paired bilinear, 1 fetch/texel"), (4, "DRAM-streamed paired bilinear, 1 fetch / 2x2 texels")):
        r = vr.bench_tex_peak(0, dt, mode)
        bpf = {3: 2 * (2 if dt == vr.U16 else 1), 4: 8 * (2 if dt == vr.U16 else 1)}.get(mode)
        print(f"{name} mode {mode} {what}: {r/1e9:.1f} Gfetch/s" + (f" = {r*bpf/1e9:.0f} GB/s of new texels" if bpf else ""))
PY
timeout 600 python bench.py --no-cpu-baseline > $O/bench_c2_r1head.json 2> $O/bench_c2_r1head.err
tail -30 $O/exp_alias_512.txt; cat $O/tex_dram.txt; python -c "
import json; d=json.load(open('$O/bench_c2_r1head.json')); print('c2', d['ms_per_step'], d['roofline_tex']['frac'], d['e2e']['ms_per_step'])"
