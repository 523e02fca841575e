"""Parity at the sizes that are benchmarked (BASELINE.json configs 1-3).

The other GPU tests pin res/shaders/volume.frag:21-52 on small volumes and frames; these run the CUDA path at the
full size of each single-GPU config against the reference's own shader compiled as CPU code (oracle/_ref; the
plain-C port when that build is absent -- the two are bit-identical, tests/test_oracle.py):

  C1  256^3 u8 Marschner-Lobb, 512x512, TF-0 / TF-A / TF-B x two cameras: the exact sampler BIT-IDENTICAL
      (float4 shader output, RGBA8, sample count), the hardware sampler within the north-star tolerance;
  C2  512^3 u16 fBm, 1920x1080, TF-B, orbit frames 0 / 120 / 240 (+ 45 and an oblique pitch) with the AUTO layout:
      within tolerance, and the pair axis the context chose is the one the camera looks down;
  C3  1024^3 u16 blobs, 3840x2160, TF-C, empty-space skipping active: one orbit frame within tolerance.
"""
import numpy as np
import pytest

from common import assert_within_tolerance, tf

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle(port):
    from oracle import pyoracle
    try:
        return pyoracle.Oracle("reference")
    except Exception:
        return port


def cameras(vr, o, yaw_steps=0, extra=None):
    a, b = vr.Camera(), o.camera()
    for _ in range(yaw_steps):
        a.rotate((4.0, 0.0))
        b.rotate(4.0, 0.0)
    if extra:
        a.rotate(extra)
        b.rotate(*extra)
    return a, b


@pytest.fixture(scope="module")
def c1(vr):
    op = vr.OffscreenPass(512, 512)
    op.synth_volume(vr.SYNTH_MARSCHNER_LOBB, 0x5EED0001, vr.U8, 256)
    vol = op.download_volume()
    yield op, vol
    op.close()


@pytest.mark.parametrize("tfn", ["0", "A", "B"])
@pytest.mark.parametrize("move", [None, (-120.0, 80.0)], ids=["default-camera", "yaw30-pitch20"])
def test_c1_full_size(vr, oracle, c1, tfn, move):
    op, vol = c1
    _, _, mn, mx = op.volume_info()
    lut = tf(oracle, tfn)
    op.transfer_function_changed(lut)
    cam, ocam = cameras(vr, oracle, 0, move)
    op.set_camera(cam)
    want = oracle.render(oracle.frame(512, 512, ocam, mn, mx), vol, lut, want_raw=True)
    op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1, collect_stats=1, empty_skip=1, early_term=0)
    op.record()
    raw, img, st = op.download_raw(), op.download(), op.stats()
    assert np.array_equal(raw.view(np.uint32), want["raw"].view(np.uint32)), "exact sampler: float4 output differs"
    assert np.array_equal(img, want["rgba8"])
    assert st.n_ref == want["stats"].n_ref
    op.set_options(sampler=vr.SAMPLER_HW, write_raw=0, collect_stats=0)
    op.record()
    assert_within_tolerance(op.download(), want["rgba8"], f"C1 TF-{tfn} hw sampler")


@pytest.fixture(scope="module")
def c2(vr):
    op = vr.OffscreenPass(1920, 1080)
    op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 512)
    vol = op.download_volume()
    yield op, vol
    op.close()


# (orbit frame, extra rotate, model axis the camera looks down: y at frame 0 -- the default camera sits at (0,-3,0))
C2_VIEWS = [(0, None, 1), (120, None, 0), (240, None, 0), (45, None, None), (30, (0.0, 100.0), None), (90, None, 0), (180, None, 1)]


@pytest.mark.parametrize("view", C2_VIEWS, ids=[f"frame{v[0]}" + ("-pitch25" if v[1] else "") for v in C2_VIEWS])
def test_c2_full_size_orbit_frames(vr, oracle, c2, view):
    k, extra, axis = view
    op, vol = c2
    _, _, mn, mx = op.volume_info()
    lut = tf(oracle, "B")
    op.transfer_function_changed(lut)
    cam, ocam = cameras(vr, oracle, k, extra)
    op.set_camera(cam)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=0)
    op.record()
    img = op.download()
    fi = op.last_frame_info()
    want = oracle.render(oracle.frame(1920, 1080, ocam, mn, mx), vol, lut)
    e, s = assert_within_tolerance(img, want["rgba8"], f"C2 orbit frame {k}")
    assert fi.layout == vr.LAYOUT_PAIRED, "a 1920x1080 frame of a 512^3 u16 volume samples the paired copy"
    if axis is not None:
        assert fi.pair_axis == axis, f"frame {k}: pairs along {'xyz'[fi.pair_axis]}, the camera looks down {'xyz'[axis]}"
    # the bench's setting (early termination at 1/1024) stays inside the tolerance too
    op.set_options(early_term=1, early_term_eps=1.0 / 1024.0)
    op.record()
    assert_within_tolerance(op.download(), want["rgba8"], f"C2 orbit frame {k}, early termination")
    # and the 3-D array copy of the same frame
    op2 = None
    if k == 0:
        op.set_options(early_term=0)
        import os
        os.environ["VR_PAIR_POLICY"] = "0"
        try:
            op2 = vr.OffscreenPass(1920, 1080)
            op2.set_volume_layout(vr.LAYOUT_ARRAY3D)
            op2.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 512)
            op2.transfer_function_changed(lut)
            op2.set_camera(cam)
            op2.record()
            assert_within_tolerance(op2.download(), want["rgba8"], "C2 orbit frame 0, 3-D array")
            assert op2.last_frame_info().layout == vr.LAYOUT_ARRAY3D
        finally:
            del os.environ["VR_PAIR_POLICY"]
            if op2:
                op2.close()


def test_c3_full_size_with_empty_space_skipping(vr, oracle):
    op = vr.OffscreenPass(3840, 2160)
    try:
        op.synth_volume(vr.SYNTH_BLOBS, 0x5EED0003, vr.U16, 1024)
        vol = op.download_volume()
        _, _, mn, mx = op.volume_info()
        lut = tf(oracle, "C")
        op.transfer_function_changed(lut)
        cam, ocam = cameras(vr, oracle, 37)
        op.set_camera(cam)
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=0, collect_stats=1)
        op.record()
        img, st, fi = op.download(), op.stats(), op.last_frame_info()
        want = oracle.render(oracle.frame(3840, 2160, ocam, mn, mx), vol, lut)
        assert fi.empty_skip_active == 1
        assert st.n_fetch < 0.2 * st.n_ref, "the sparse volume must be mostly skipped"
        assert abs(int(st.n_ref) - int(want["stats"].n_ref)) <= 1e-3 * want["stats"].n_ref
        assert_within_tolerance(img, want["rgba8"], "C3 hw sampler + skipping")
        op.set_options(early_term=1, early_term_eps=1.0 / 1024.0, collect_stats=0)
        op.record()
        assert_within_tolerance(op.download(), want["rgba8"], "C3 hw sampler + skipping + early termination")
    finally:
        op.close()
