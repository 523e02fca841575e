"""tools/exp_sl_planes.py [n] -- experiment: does moving the sort-last split planes with the camera (vr_partition_balance)
even out the ranks' K1 times?  On ONE GPU each rank's share of the frame (2048^3, 3840x2160, the 36-frame orbit of
bench.py) is marched in turn and timed per frame, with the fixed regular split and with planes that follow the camera
within 0.08 of the edge.  Reports the lock-step bound (mean over frames of the slowest rank) against one GPU / N."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumetric_renderer_b200 as vr  # noqa: E402
from volumetric_renderer_b200 import multigpu as mg  # noqa: E402
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
W, H = 3840, 2160
lut = bench.product_lut(vr, "B")
frames = bench.partitioned_frames(vr, W, H)
stream = torch.cuda.Stream()
sp = stream.cuda_stream
full = (n, n, n)


def time_frames(op, per_frame=None):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(37)]
    for k in range(3):
        if per_frame:
            per_frame(k)
        op.set_uniforms(*frames[k])
        op.record(sp)
    op.sync()
    ev[0].record(stream)
    for k in range(36):
        if per_frame:
            per_frame(k)
        op.set_uniforms(*frames[k])
        op.record(sp)
        ev[k + 1].record(stream)
    op.sync()
    return np.array([ev[k].elapsed_time(ev[k + 1]) for k in range(36)])


op = vr.OffscreenPass(W, H)
op.transfer_function_changed(lut)
op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0)
op.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
_, _, gmn, gmx = op.volume_info()
t1 = time_frames(op)
op.close()
print(f"one GPU: {t1.mean():.4f} ms/frame")
out = {"n1_ms": float(t1.mean())}
for N in (2, 4, 8):
    grid = mg.grid_for_world(N)
    for balance in (0.0, 0.08):
        margin = tuple(min(int(balance * n), max((n // g - 1) // 2, 0)) if g > 1 else 0 for g in grid)
        t = np.zeros((N, 36))
        samples = np.zeros((N, 36))
        for r in range(N):
            org, dims = vr.partition_stored_box(grid, full, r, margin)
            op = vr.OffscreenPass(W, H)
            op.transfer_function_changed(lut)
            op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1)
            op.set_partition_sort_last(full, org, dims)
            op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
            op.set_density_range(gmn, gmx)

            def per_frame(k, op=op, r=r):
                if balance:
                    planes = vr.partition_balance(grid, full, frames[k][2], margin)
                    op.set_owned_box(*vr.partition_box_planes(grid, full, planes, r))
            t[r] = time_frames(op, per_frame)
            op.close()
        lock, pipe, ideal = t.max(axis=0).mean(), t.mean(axis=1).max(), t1.mean() / N
        print(f"N={N} grid {grid} balance {balance}: ideal {ideal:.4f}  lock-step bound {lock:.4f} (eff {ideal / lock:.3f})  pipelined bound {pipe:.4f} "
              f"(eff {ideal / pipe:.3f})  sum of rank means / one GPU = {t.mean(axis=1).sum() / t1.mean():.3f}")
        for k in (0, 4, 9, 13):
            print(f"    frame {k}: rank times " + " ".join(f"{x:.3f}" for x in t[:, k]))
        out[f"n{N}_balance{balance}"] = {"grid": list(grid), "lock_step_ms": float(lock), "pipelined_ms": float(pipe), "ideal_ms": float(ideal)}
print(json.dumps(out))
