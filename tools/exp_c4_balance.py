#!/usr/bin/env python
"""tools/exp_c4_balance.py: C4 (2048^3 u16, 7680x4320, sort-first) on ONE GPU -- each rank's share of every frame is
rendered in turn and device-timed, so the kernel-side cost of the partition (load imbalance between ranks, per-rank
tail) is separated from the hand-off between GPUs.  Prints, per N and tile size: sum over frames of the slowest
rank's time against full-frame time / N."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import volumetric_renderer_b200 as vr
from bench import product_lut

n = int(os.environ.get("N", "2048"))
W, H = 7680, 4320
op = vr.OffscreenPass(W, H, device=0)
op.transfer_function_changed(product_lut(vr, "B"))
op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, collect_stats=0)
op.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
cam, proj, frames = vr.Camera(), vr.projection(W, H), []
for _ in range(36):
    frames.append((cam.get_view(), proj, cam.get_position()))
    cam.rotate((40.0, 0.0))
frames = frames[::3]
s = torch.cuda.Stream()

def timed(reps=3):
    out = []
    for u in frames:
        op.set_uniforms(*u)
        op.record(s.cuda_stream)          # warm
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            op.record(s.cuda_stream)
        e1.record(s)
        s.synchronize()
        out.append(e0.elapsed_time(e1) / reps)
    return out

full = timed()
print("full frame ms: mean %.3f  min %.3f max %.3f" % (sum(full) / len(full), min(full), max(full)))
for N in (2, 4, 8):
    for tile in (32, 64, 128):
        per = []
        for r in range(N):
            op.set_partition_sort_first(r, N, tile)
            per.append(timed())
        op.clear_partition()
        worst = [max(per[r][k] for r in range(N)) for k in range(len(frames))]
        mean = [sum(per[r][k] for r in range(N)) / N for k in range(len(frames))]
        print("N=%d tile=%3d: slowest-rank %.3f ms  mean-rank %.3f ms  ideal %.3f ms  -> efficiency %.2f (imbalance %.2f, overhead %.2f)" % (
            N, tile, sum(worst) / len(worst), sum(mean) / len(mean), sum(full) / len(full) / N,
            (sum(full) / N) / sum(worst), sum(mean) / sum(worst), (sum(full) / N) / sum(mean)))
