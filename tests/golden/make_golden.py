#!/usr/bin/env python
"""Generate tests/golden/*.npz from oracle/_ref -- the reference's OWN sources compiled as CPU code
(res/shaders/volume.frag body + src/scene/camera.cpp + src/ui/components/gradient.cpp +
src/data/nrrd_file_parser.cpp/NrrdIO, see oracle/Makefile).  Needs /root/reference; run in the build
container:   python tests/golden/make_golden.py

The reference ships no fixtures of its own, so these vectors are the pin:
the plain-C port (oracle/vr_oracle.c) and the CUDA path are both checked against them.
Each case stores the inputs (volume, LUT, uniform block) and the outputs (float4 shader output,
RGBA8 image, sample counts), so a checker needs nothing but the file.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
W, H = 96, 72

CASES = {
    # name: (synth kind, dtype, n, TF, camera moves, zoom, (min_slice, max_slice))
    "ml_u8_tfA": (0, np.uint8, 24, "A", [(-120, 80)], 0.0, None),
    "ml_u8_tf0": (0, np.uint8, 24, "0", [(37, -21)], 0.0, None),
    "fbm_u16_tfB": (1, np.uint16, 32, "B", [(-120, 80)], 0.0, None),
    "fbm_u16_tfB_sliced_zoom": (1, np.uint16, 32, "B", [(200, 10)], 1.2, ((0.2, 0.1, 0.3), (0.9, 0.8, 0.7))),
    "blobs_u16_tfC": (2, np.uint16, 32, "C", [(45, 30)], 0.0, None),
    "fbm_f32_nonpow2": (1, np.float32, (21, 12, 7), "B", [(10, 55)], 0.5, None),
    "near_plane_clip": (0, np.uint8, 24, "B", [(15, 15)], 2.05, None),
    "camera_inside": (0, np.uint8, 24, "A", [(15, 15)], 2.6, None),
}
MARKERS = {"A": pyoracle.TF_A_MARKERS, "B": pyoracle.TF_B_MARKERS, "C": pyoracle.TF_C_MARKERS}


def main():
    ref = pyoracle.Oracle("reference")
    port = pyoracle.Oracle("port")  # synthetic volumes only (data, not the path)
    for name, (kind, dt, n, tfn, moves, zoom, slices) in CASES.items():
        vol = port.synth(kind, 0x5EED0000 + kind, dt, n)
        lut = pyoracle.TF_0.copy() if tfn == "0" else ref.discretize(*MARKERS[tfn])
        cam = ref.camera()
        for m in moves:
            cam.rotate(*m)
        if zoom:
            cam.zoom(zoom)
        mn, mx = float(vol.min()), float(vol.max())
        lo, hi = slices if slices else ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0))
        f = ref.frame(W, H, cam, mn, mx, lo, hi)
        r = ref.render(f, vol, lut, want_raw=True, want_iters=True)
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"), volume=vol, lut=lut, width=W, height=H,
            view=np.array(f.view, np.float32), proj=np.array(f.proj, np.float32),
            camera_position=np.array(f.camera_position, np.float32), min_density=np.float32(mn),
            max_density=np.float32(mx), min_slice=np.array(lo, np.float32), max_slice=np.array(hi, np.float32),
            moves=np.array(moves, np.float32).reshape(-1, 2), zoom=np.float32(zoom),
            raw=r["raw"], rgba8=r["rgba8"], iters=r["iters"],
            stats=np.array([r["stats"].covered_pixels, r["stats"].n_ref, r["stats"].n_inslice], np.uint64))
        print(name, "covered", r["stats"].covered_pixels, "n_ref", r["stats"].n_ref)
    # host-side surfaces: default gradient, TF-B/C, camera after a few moves, projections
    cam = ref.camera()
    cams = []
    for m in [(0, 0), (4, 0), (-120, 80), (33.3, -17.7), (400, 5)]:
        cam.rotate(*m)
        cams.append(np.concatenate([cam.view, cam.position]))
    np.savez_compressed(
        os.path.join(OUT, "host_surfaces.npz"),
        lut_A=ref.discretize(*MARKERS["A"]), lut_B=ref.discretize(*MARKERS["B"]), lut_C=ref.discretize(*MARKERS["C"]),
        lut_B_7=ref.discretize(*MARKERS["B"], count=7), cameras=np.array(cams, np.float32),
        proj_512=ref.projection(512, 512), proj_1080=ref.projection(1920, 1080), proj_100=ref.projection(100, 100))


if __name__ == "__main__":
    main()
