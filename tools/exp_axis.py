"""tools/exp_axis.py [N]: time of one c2-like frame (1920x1080, N^3 u16 fBm, TF-B) per camera direction and
per sampling layout (3-D array, pairs along x / y / z).  Experiment, run under gpurun."""
import os, sys, subprocess, json
N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
code = r'''
import os, sys, torch
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
import bench
N = int(sys.argv[1])
op = vr.OffscreenPass(1920, 1080)
op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, N)
op.transfer_function_changed(bench.product_lut(vr, "B"))
op.set_options(sampler=vr.SAMPLER_HW, empty_skip=0, early_term=1, early_term_eps=1/1024)
st = torch.cuda.Stream(); sp = st.cuda_stream
out = []
for name, moves in (("along -y", []), ("yaw 22", [(90, 0)]), ("yaw 45", [(180, 0)]), ("along x", [(360, 0)]), ("pitch 45", [(0, 180)]), ("along z", [(0, 350)]), ("oblique", [(120, 140)])):
    cam = vr.Camera()
    for m in moves: cam.rotate(m)
    op.set_camera(cam)
    for _ in range(20): op.record(sp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(st)
    for _ in range(200): op.record(sp)
    e1.record(st); torch.cuda.synchronize()
    out.append((name, e0.elapsed_time(e1) / 200, op.last_frame_pair_axis()))
print(" ".join(f"{n}={t:.4f}" for n, t, a in out))
'''
for label, env in (("3d", {"VR_PAIR_POLICY": "0"}), ("pairs x", {"VR_PAIR_AXIS": "0"}), ("pairs y", {"VR_PAIR_AXIS": "1"}), ("pairs z", {"VR_PAIR_AXIS": "2"}), ("auto", {})):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code, str(N)], env=e, capture_output=True, text=True)
    print(f"N={N} {label:8s}: {r.stdout.strip() or r.stderr[-300:]}", flush=True)
