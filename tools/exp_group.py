"""tools/exp_group.py N: the one-process host (vr_create_group, csrc/group.cu) timed on N GPUs of one box -- the same two
partitioned workloads bench.py's `partitioned` record times with one process per GPU: sort-first C4 (2048^3 u16 fBm,
7680x4320, TF-B) and sort-last on the same volume at 3840x2160; 36-frame orbit, wall clock around K frames enqueued back
to back + one sync."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumetric_renderer_b200 as vr
import bench

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
lut = bench.product_lut(vr, "B")
for name, part, (W, H), slots in (("sort-first", vr.PARTITION_SORT_FIRST, (7680, 4320), 2), ("sort-last", vr.PARTITION_SORT_LAST, (3840, 2160), 4)):
    g = vr.PassGroup(list(range(N)), part, W, H, slots)
    try:
        g.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0)
        g.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
        g.transfer_function_changed(lut)
        frames = bench.partitioned_frames(vr, W, H)
        for k in range(6):
            g.set_uniforms(*frames[k])
            g.record()
        g.sync()
        K = 72
        t0 = time.perf_counter()
        for k in range(K):
            g.set_uniforms(*frames[k % 36])
            g.record()
        g.sync()
        dt = (time.perf_counter() - t0) / K
        img = g.download()
        print(f"vr_group {name} on {N} GPU(s): {dt * 1e3:.4f} ms/frame, checksum of the last frame {int(img.astype(np.int64).sum())}", flush=True)
    finally:
        g.close()
