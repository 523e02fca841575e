#!/bin/bash
# tools/r2_final.sh [part]: everything profiles/ holds for the final kernel version of round 2 (run under gpurun, one GPU).
#   part a: tests, smoke, bench lines;  part b: launch lists and the ncu --set full captures (each after the same command
#   exited 0 without ncu).  Output: gpurun_out/r2_final/ -> tools/save_profiles_r2.py
O=gpurun_out/r2_final; mkdir -p $O
PART=${1:-ab}
python -c "import volumetric_renderer_b200 as vr; print(vr.KERNEL_VERSION)" > $O/kernel_version.txt
if [[ $PART == *a* ]]; then
(timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 2>&1 | tail -20) > $O/pytest_gpu.log
python __graft_entry__.py smoke > $O/smoke.log 2>&1
( time python bench.py ) > $O/bench_c2.json 2> $O/bench_c2.err
( time python bench.py --impl reference ) > $O/bench_c2_reference.json 2> $O/bench_ref.err
python bench.py --workload c3 --steps 90 --no-partitioned > $O/bench_c3.json 2> $O/bench_c3.err
python bench.py --workload c1 --no-partitioned > $O/bench_c1.json 2> $O/bench_c1.err
tail -4 $O/pytest_gpu.log; tail -2 $O/smoke.log
for w in c2 c3 c1 c2_reference; do python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$w.json").read().strip().splitlines()[-1])
    print("$w", round(d.get('ms_per_step',0),4),'ms value',f"{d['value']:.3e}", 'e2e', round(d['e2e'].get('ms_per_step',0),4), 'tex', round((d.get('roofline_tex') or {}).get('frac',0),3), 'texdram', round((d.get('roofline_tex_dram') or {}).get('frac',0),3), 'hbm', round((d.get('roofline') or {}).get('frac',0),3), (d.get('cpu_baseline') or {}).get('value'))
except Exception as e: print("$w", "ERR", e)
PY
done
fi
if [[ $PART == *b* ]]; then
C2="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-partitioned"
C3="python bench.py --workload c3 --steps 4 --warmup 3 --no-cpu-baseline --no-partitioned"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base mangled -c 1 -f"
timeout 300 $C2 > $O/plain_c2.json 2> $O/plain_c2.err && {
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv $C2 > $O/ncu_list_c2.log 2>&1
  timeout 600 $NCU -k regex:k_marchILb0ELb0ELb0ELb0ELi1 -s 6 -o $O/march_c2 $C2 > $O/ncu_c2.log 2>&1
  timeout 600 $NCU -k regex:k_texstreamILi1 -s 1 -o $O/texstream $C2 > $O/ncu_texstream.log 2>&1
}
timeout 300 $C3 > $O/plain_c3.json 2> $O/plain_c3.err && {
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv $C3 > $O/ncu_list_c3.log 2>&1
  timeout 600 $NCU -k regex:k_marchILb0ELb1ELb0ELb0ELi0 -s 6 -o $O/march_c3 $C3 > $O/ncu_c3.log 2>&1
  timeout 600 $NCU -k regex:k_tile_mask -s 6 -o $O/tile_mask_c3 $C3 > $O/ncu_mask_c3.log 2>&1
}
ls -la $O/*.ncu-rep; for f in ncu_c2 ncu_c3 ncu_texstream ncu_mask_c3; do tail -n 2 $O/$f.log; done
fi
