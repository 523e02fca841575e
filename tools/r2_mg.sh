#!/bin/bash
# tools/r2_mg.sh N TAG: the driver's N-GPU bench command + the multi-process test (run under gpurun --gpus N)
N=${1:-2}; TAG=${2:-r2_mg}; O=gpurun_out/$TAG; mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -8 ) > $O/pytest_multi_n$N.log 2>&1
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 ) > $O/bench_n$N.json 2> $O/bench_n$N.err
tail -3 $O/pytest_multi_n$N.log; tail -5 $O/bench_n$N.err
python - <<PY
import json
d=json.loads(open("$O/bench_n$N.json").read().strip().splitlines()[-1])
print("c2", d["n_gpus"], d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["ms_per_step_full_frame_every_step"])
print(json.dumps(d.get("partitioned"), indent=1)[:3000])
PY
# host-side picture for the full-frame D2H figure at N ranks
( nvidia-smi topo -m; lscpu | grep -i "numa\|model name\|socket" ) > $O/topo_n$N.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 tools/exp_d2h.py > $O/exp_d2h_n$N.txt 2> $O/exp_d2h_n$N.err; tail -12 $O/exp_d2h_n$N.txt; tail -3 $O/exp_d2h_n$N.err
