#!/bin/bash
# round 2, run 5: full GPU test-suite with the full-size parity tests, the new bench line at N=1 (incl. the C4 N=1 denominator)
O=gpurun_out/r2_run5; mkdir -p $O
(timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 2>&1 | tail -25) > $O/pytest_gpu.log
python __graft_entry__.py smoke > $O/smoke.log 2>&1
( time python bench.py --steps 20 --warmup 3 ) > $O/bench_c2_k20.json 2> $O/bench_c2_k20.err
( time python bench.py --impl reference --steps 20 --warmup 3 ) > $O/bench_ref_k20.json 2> $O/bench_ref_k20.err
tail -6 $O/pytest_gpu.log; tail -2 $O/smoke.log; tail -4 $O/bench_c2_k20.err; tail -4 $O/bench_ref_k20.err
python - <<PY
import json
d=json.load(open("$O/bench_c2_k20.json"))
for k in ("value","ms_per_step","ms_per_step_in_flight_1","ms_per_step_early_termination_off","reps","timed_region_s"): print(k, d.get(k))
print("tex", d["roofline_tex"]["frac"], "texdram", d["roofline_tex_dram"], "hbm", d["roofline"]["frac"])
print("e2e", d["e2e"]["ms_per_step"], d["e2e"]["ms_per_step_full_frame_every_step"])
print("parity", json.dumps(d.get("cpu_baseline",{}).get("parity_vs_gpu_frames"), indent=0)[:1500])
print("partitioned", json.dumps(d.get("partitioned"))[:1500])
print(d["config"]["layout_per_frame"])
PY
