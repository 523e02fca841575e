"""tools/exp_sl_balance.py -- experiment: where does sort-last lose its efficiency?  On ONE GPU, each rank's share of
the frame (its sub-box of the 2048^3 volume, 3840x2160, the 36-frame orbit of bench.py) is marched in turn and
timed per frame.  From t[rank][frame]:
    lock-step bound   = mean over frames of the slowest rank   (every frame waits for its slowest rank: today)
    pipelined bound   = the slowest rank's mean                (ranks may run ahead of one another)
against the single-GPU time / N.  No merge, no hand-off: what K1 alone allows."""
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


def time_frames(op):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(37)]
    for k in range(3):
        op.set_uniforms(*frames[k])
        op.record(sp)
    op.sync()
    ev[0].record(stream)
    for k in range(36):
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
print(f"one GPU: {t1.mean():.4f} ms/frame (min {t1.min():.4f}, max {t1.max():.4f})")
out = {"n1_ms": float(t1.mean())}
for N in (2, 4, 8):
    grid = mg.grid_for_world(N)
    t = np.zeros((N, 36))
    for r in range(N):
        org, dims = vr.partition_box(grid, (n, n, n), r)
        op = vr.OffscreenPass(W, H)
        op.transfer_function_changed(lut)
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1)
        op.set_partition_sort_last((n, n, n), org, dims)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
        op.set_density_range(gmn, gmx)
        t[r] = time_frames(op)
        op.close()
    lock, pipe, ideal = t.max(axis=0).mean(), t.mean(axis=1).max(), t1.mean() / N
    print(f"N={N} grid {grid}: ideal {ideal:.4f}  lock-step bound {lock:.4f} (eff {ideal / lock:.3f})  pipelined bound {pipe:.4f} (eff {ideal / pipe:.3f})"
          f"  sum of rank means / one GPU = {t.mean(axis=1).sum() / t1.mean():.3f}")
    out[f"n{N}"] = {"grid": list(grid), "lock_step_ms": float(lock), "pipelined_ms": float(pipe), "ideal_ms": float(ideal),
                    "rank_means": [float(x) for x in t.mean(axis=1)]}
# two bricks per rank, a brick and its point mirror (the mirror of a near brick is a far brick: the perspective
# imbalance of a view cancels inside every rank)
for N in (2, 4, 8):
    grid = mg.grid_for_world(2 * N)
    nb = 2 * N
    tb = np.zeros((nb, 36))
    for b in range(nb):
        org, dims = vr.partition_box(grid, (n, n, n), b)
        op = vr.OffscreenPass(W, H)
        op.transfer_function_changed(lut)
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, write_raw=1)
        op.set_partition_sort_last((n, n, n), org, dims)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
        op.set_density_range(gmn, gmx)
        tb[b] = time_frames(op)
        op.close()
    t = np.array([tb[r] + tb[nb - 1 - r] for r in range(N)])
    lock, pipe, ideal = t.max(axis=0).mean(), t.mean(axis=1).max(), t1.mean() / N
    print(f"N={N} mirror pairs of grid {grid}: ideal {ideal:.4f}  lock-step bound {lock:.4f} (eff {ideal / lock:.3f})  pipelined bound {pipe:.4f} (eff {ideal / pipe:.3f})"
          f"  sum of rank means / one GPU = {t.mean(axis=1).sum() / t1.mean():.3f}")
    out[f"n{N}_pairs"] = {"grid": list(grid), "lock_step_ms": float(lock), "pipelined_ms": float(pipe), "ideal_ms": float(ideal)}
print(json.dumps(out))
