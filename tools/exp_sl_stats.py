"""tools/exp_sl_stats.py: per-rank step / sample counters and K1 time of single frames of the sort-last split (one GPU)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumetric_renderer_b200 as vr
from volumetric_renderer_b200 import multigpu as mg
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
W, H = 3840, 2160
lut = bench.product_lut(vr, "B")
frames = bench.partitioned_frames(vr, W, H)
full = (n, n, n)
N = int(os.environ.get("RANKS", "8"))
grid = mg.grid_for_world(N)
st = torch.cuda.Stream(); sp = st.cuda_stream
for r in range(N):
    org, dims = vr.partition_box(grid, full, r)
    op = vr.OffscreenPass(W, H)
    op.transfer_function_changed(lut)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1)
    op.set_partition_sort_last(full, org, dims)
    op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
    for k in [int(x) for x in os.environ.get("FRAMES", "0,4,9").split(",")]:
        op.set_uniforms(*frames[k])
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1, collect_stats=1)
        op.record(sp); op.sync(); s = op.stats(); fi = op.last_frame_info()
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1, collect_stats=0)
        for _ in range(3): op.record(sp)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        op.sync(); e0.record(st)
        for _ in range(10): op.record(sp)
        e1.record(st); op.sync()
        t = e0.elapsed_time(e1) / 10
        print(f"rank {r} box {org} frame {k}: {t:.3f} ms  covered {s.covered_pixels} visited {s.n_ref} inslice {s.n_inslice} fetched {s.n_fetch}  "
              f"bbox {tuple(fi.bbox)} blocks {fi.march_blocks} layout {fi.layout} axis {fi.pair_axis} ns/fetch {t*1e6/max(s.n_fetch,1):.4f}")
    op.close()
