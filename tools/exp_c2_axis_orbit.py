"""tools/exp_c2_axis_orbit.py: K1 time of 36 orbit frames of config 2 (512^3 u16, 1920x1080, TF-B) with the pairs forced along
x, y, z, on the 3-D array, and with the context's own choice: is the per-frame choice the best one?  (one GPU)"""
import json, os, subprocess, sys
code = r'''
import os, sys, json
import torch
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
import bench
W, H, n = 1920, 1080, 512
op = vr.OffscreenPass(W, H)
op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, n)
op.transfer_function_changed(bench.product_lut(vr, "B"))
op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1/1024)
st = torch.cuda.Stream(); sp = st.cuda_stream
out = []
for k in range(36):
    cam = vr.Camera()
    for _ in range(k * 10): cam.rotate((4.0, 0.0))
    op.set_camera(cam)
    for _ in range(5): op.record(sp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(st)
    for _ in range(30): op.record(sp)
    e1.record(st); torch.cuda.synchronize()
    fi = op.last_frame_info()
    out.append((e0.elapsed_time(e1) / 30, int(fi.pair_axis) if fi.layout == vr.LAYOUT_PAIRED else -1))
print(json.dumps(out))
'''
tab = {}
for label, env in (("x", {"VR_PAIR_AXIS": "0"}), ("y", {"VR_PAIR_AXIS": "1"}), ("z", {"VR_PAIR_AXIS": "2"}), ("3d", {"VR_PAIR_POLICY": "0"}), ("auto", {})):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True)
    tab[label] = json.loads(r.stdout.strip().splitlines()[-1])
print("frame  x      y      z      3d     auto(axis)")
for k in range(36):
    print(f"{k:3d}  " + "  ".join(f"{tab[a][k][0]:.4f}" for a in ("x", "y", "z", "3d")) + f"  {tab['auto'][k][0]:.4f} ({'xyz3'[tab['auto'][k][1]]})")
best = sum(min(tab[a][k][0] for a in ("x", "y", "z", "3d")) for k in range(36)) / 36
print(f"mean: auto {sum(v[0] for v in tab['auto'])/36:.4f} ms, best per frame {best:.4f} ms (one frame in flight)")
