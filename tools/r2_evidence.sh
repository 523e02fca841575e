#!/bin/bash
# tools/r2_evidence.sh: the round-2 single-GPU experiments whose outputs DESIGN.md cites (run under gpurun; tools/*.bin built
# beforehand, see the Build line in each .cu).  Output: gpurun_out/r2_evidence/, copied to profiles/ by tools/save_profiles_r2.py.
O=gpurun_out/r2_evidence; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
timeout 300 python - > $O/exp_tex_dram.txt 2>&1 <<'PY'
import volumetric_renderer_b200 as vr
print("# vr_bench_tex_peak (csrc/texbench.cu): texture-path ceilings, L1-resident and fed from DRAM (1 GiB array streamed once)")
for dt, name in ((vr.U16, "u16"), (vr.U8, "u8")):
    for mode, what in ((1, "L1-resident 3-D trilinear"), (2, "L1-resident paired bilinear"), (3, "DRAM-streamed paired bilinear, 1 fetch/texel"), (4, "DRAM-streamed paired bilinear, 1 fetch / 2x2 texels")):
        r = vr.bench_tex_peak(0, dt, mode)
        bpf = {3: 2 * (2 if dt == vr.U16 else 1), 4: 8 * (2 if dt == vr.U16 else 1)}.get(mode)
        print(f"{name} mode {mode} {what}: {r/1e9:.1f} Gfetch/s" + (f" = {r*bpf/1e9:.0f} GB/s of new texels" if bpf else ""))
PY
( echo "# tools/exp_alias.cu: can CUDA-array memory be aliased linearly (VMM) so that prefetch.global.L2 could feed the texture path?"; timeout 120 tools/exp_alias.bin 512; echo "exit $?" ) > $O/exp_alias.txt 2>&1
( echo "# tools/exp_depth.cu: does the march-shaped fetch stream go faster with more fetches in flight per thread / other occupancies?"; timeout 300 tools/exp_depth.bin 512; echo "exit $?" ) > $O/exp_depth.txt 2>&1
( echo "# tools/exp_region.cu: tiles dealt to CTAs as they come (A) against one compact screen region per SM (B)"; timeout 120 tools/exp_region.bin 512; echo "exit $?" ) > $O/exp_region.txt 2>&1
timeout 600 python tools/exp_sl_balance.py > $O/exp_sl_balance.txt 2>&1
timeout 300 python tools/exp_c3.py > $O/exp_c3.txt 2>&1
timeout 300 python tools/exp_c3_gran.py > $O/exp_c3_gran.txt 2>&1
# dirty clear A/B and the register cap A/B at C2
python bench.py --no-cpu-baseline --no-partitioned > $O/bench_c2.json 2> $O/bench_c2.err
VR_DIRTY_CLEAR=0 python bench.py --no-cpu-baseline --no-partitioned > $O/bench_c2_fullclear.json 2> /dev/null
for f in exp_tex_dram exp_alias exp_depth exp_region exp_sl_balance exp_c3 exp_c3_gran; do echo "== $f"; tail -6 $O/$f.txt; done
for f in bench_c2 bench_c2_fullclear; do python -c "
import json; d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1]); print('$f', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4))"; done
