"""tools/exp_load.py: time of the volume load path (vr_load_nrrd, fast path) for a 512^3 u16 raw NRRD against
vr_set_volume from pageable host memory.  Experiment, run under gpurun."""
import os, sys, time, tempfile
import numpy as np
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
arr = (np.arange(n ** 3, dtype=np.uint32) * 2654435761 >> 16).astype(np.uint16).reshape(n, n, n)
d = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
path = os.path.join(d, "v.nrrd")
with open(path, "wb") as f:
    f.write(f"NRRD0004\ntype: ushort\ndimension: 3\nsizes: {n} {n} {n}\nencoding: raw\nendian: little\n\n".encode())
    f.write(arr.tobytes())
op = vr.OffscreenPass(64, 64)
for layout, name in ((vr.LAYOUT_ARRAY3D, "3-D array only"), (vr.LAYOUT_AUTO, "3-D array + paired copies")):
    op.set_volume_layout(layout)
    for label, fn in (("vr_load_nrrd (file -> device)", lambda: op.import_nrrd(path)),
                      ("vr_set_volume (pageable host array)", lambda: op.volume_dataset_changed(vr.Dataset(arr)))):
        fn()
        dt = 1e9
        for _ in range(5):
            t0 = time.perf_counter()
            fn()
            dt = min(dt, time.perf_counter() - t0)
        print(f"{n}^3 u16 ({arr.nbytes / 2**20:.0f} MiB), {name}: {label}: {dt * 1e3:.1f} ms = {arr.nbytes / dt / 1e9:.1f} GB/s", flush=True)
op.close()
os.remove(path)
