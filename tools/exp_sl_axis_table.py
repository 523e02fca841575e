"""tools/exp_sl_axis_table.py: K1 time of every sort-last rank's share, for every frame of the 36-frame orbit and every pair
axis (and the 3-D array) -- the table the pair-axis choice of a sort-last rank is judged against (one GPU)."""
import json, os, subprocess, sys
code = r'''
import os, sys, json
import numpy as np, torch
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
from volumetric_renderer_b200 import multigpu as mg
import bench
n, N = 2048, int(sys.argv[1])
W, H = 3840, 2160
lut = bench.product_lut(vr, "B")
frames = bench.partitioned_frames(vr, W, H)
full = (n, n, n)
grid = mg.grid_for_world(N)
st = torch.cuda.Stream(); sp = st.cuda_stream
out = {}
for r in range(N):
    org, dims = vr.partition_box(grid, full, r)
    op = vr.OffscreenPass(W, H)
    op.transfer_function_changed(lut)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1)
    op.set_partition_sort_last(full, org, dims)
    op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
    ts = []
    for k in range(36):
        op.set_uniforms(*frames[k])
        for _ in range(2): op.record(sp)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        op.sync(); e0.record(st)
        for _ in range(4): op.record(sp)
        e1.record(st); op.sync()
        fi = op.last_frame_info()
        ts.append((e0.elapsed_time(e1) / 4, int(fi.pair_axis) if fi.layout == vr.LAYOUT_PAIRED else -1))
    out[r] = ts
    op.close()
print(json.dumps(out))
'''
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
table = {}
for label, env in (("x", {"VR_PAIR_AXIS": "0"}), ("y", {"VR_PAIR_AXIS": "1"}), ("z", {"VR_PAIR_AXIS": "2"}), ("auto", {})):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code, str(N)], env=e, capture_output=True, text=True)
    try:
        table[label] = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        print(label, "failed", r.stderr[-500:]); continue
t = {k: {int(r): v for r, v in d.items()} for k, d in table.items()}
best = [max(min(t[a][r][k][0] for a in "xyz") for r in range(N)) for k in range(36)]
auto = [max(t["auto"][r][k][0] for r in range(N)) for k in range(36)]
print(f"N={N}: lock-step bound with the context's own choice {sum(auto)/36:.4f} ms; with the best axis per rank and frame {sum(best)/36:.4f} ms")
print(json.dumps(table))
