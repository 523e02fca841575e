"""tools/exp_c3.py: where config 3's frame time goes (1024^3 u16 blobs, 3840x2160, TF-C, skipping on).
Variants per orbit frame: mask on / off, all-transparent LUT (pure set-up + traversal), statistics.  Run under gpurun."""
import os, sys, subprocess
code = r'''
import os, sys, torch, numpy as np
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
import bench
op = vr.OffscreenPass(3840, 2160)
op.synth_volume(vr.SYNTH_BLOBS, 0x5EED0003, vr.U16, 1024)
lut = bench.product_lut(vr, "C")
st = torch.cuda.Stream(); sp = st.cuda_stream
def timed(n=60):
    for _ in range(10): op.record(sp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(st)
    for _ in range(n): op.record(sp)
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
rows = []
for k in (0, 11, 23, 45, 68, 90):
    cam = vr.Camera()
    for _ in range(k): cam.rotate((4.0, 0.0))
    op.set_camera(cam)
    op.transfer_function_changed(lut)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1/1024, collect_stats=0)
    t = timed()
    m = op.tile_mask(0)
    fi = op.last_frame_info()
    info = ""
    if m is not None:
        bits, ign, th = m
        x0, y0, x1, y1 = fi.bbox
        inside = bits[y0 // th:(y1 + th - 1) // th, x0 // 16:(x1 + 15) // 16]
        info = f"tiles {inside.size} marked {int(inside.sum())} th {th} blocks {fi.march_blocks}"
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1/1024, collect_stats=1)
    op.record(sp); s = op.stats()
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1/1024, collect_stats=0)
    z = lut.copy(); z &= 0x00FFFFFF
    op.transfer_function_changed(z)
    t0 = timed()
    rows.append(f"frame {k}: t={t:.4f} transparent-lut={t0:.4f} cov={s.covered_pixels} n_ref={s.n_ref} n_fetch={s.n_fetch} n_inslice={s.n_inslice} {info}")
print("\n".join(rows))
'''
for label, env in (("mask", {}), ("nomask", {"VR_TILE_MASK": "0"})):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True)
    print(f"== {label}\n{r.stdout.strip()}\n{r.stderr[-600:] if r.returncode else ''}", flush=True)
