"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle.

Bars (BASELINE.json north_star):
  * VR_SAMPLER_EXACT: the shader's float4 output is BIT-IDENTICAL to the oracle, and so are
    the RGBA8 image and the sample counts;
  * VR_SAMPLER_HW (hardware trilinear, 9-bit weights): <= 2/255 per channel at the
    99.9th-percentile pixel and >= 45 dB PSNR (tests/common.py).
Empty-space skipping must not change a single bit; early ray termination at eps changes a
channel by less than eps.
"""
import numpy as np
import pytest

from common import assert_within_tolerance, bits, oracle_frame, p999, psnr, tf, twin_cameras

pytestmark = pytest.mark.gpu

W, H = 160, 120


def small_volume(port, kind, dtype, n):
    return port.synth(kind, 0x5EED0000 + kind, dtype, n)


def render_both(vr, port, vol, lut, moves=(), zoom=0.0, w=W, h=H, min_slice=(0, 0, 0), max_slice=(1, 1, 1),
                minmax=None, layout=None, **opts):
    cam, ocam = twin_cameras(vr, port, moves, zoom)
    mn, mx = minmax if minmax is not None else (float(vol.min()), float(vol.max()))
    op = vr.OffscreenPass(w, h)
    try:
        op.set_options(write_raw=1, collect_stats=1, **opts)
        if layout is not None:
            op.set_volume_layout(layout)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(lut)
        op.slicing_changed(min_slice, max_slice)
        op.set_camera(cam)
        op.record()
        img, raw, st = op.download(), op.download_raw(), op.stats()
        if layout is not None and opts.get("sampler", vr.SAMPLER_HW) == vr.SAMPLER_HW and mx > mn:
            assert op.last_frame_paired() == (layout == vr.LAYOUT_PAIRED)
    finally:
        op.close()
    o = port.render(oracle_frame(port, ocam, w, h, mn, mx, min_slice, max_slice), vol, lut, want_raw=True)
    return img, raw, st, o


CASES = [
    # (volume kind, dtype, size, TF, camera moves, zoom, slices)
    (0, np.uint8, 32, "A", [(-120, 80)], 0.0, None),
    (0, np.uint8, 32, "B", [], 0.0, None),
    (0, np.uint8, 32, "0", [(37, -21)], 0.0, None),
    (1, np.uint16, 48, "B", [(-120, 80)], 0.0, None),
    (1, np.uint16, 48, "B", [(200, 10)], 1.2, ((0.2, 0.1, 0.3), (0.9, 0.8, 0.7))),
    (2, np.uint16, 40, "C", [(45, 30)], 0.0, None),
    (1, np.float32, (33, 20, 7), "B", [(10, 55)], 0.5, None),
    (0, np.uint8, 32, "A", [(15, 15)], 2.6, None),    # radius 0.4: camera inside the cube -> nothing drawn
    (0, np.uint8, 32, "B", [(15, 15)], 2.05, None),   # radius 0.95: near plane clips part of the front face
    (1, np.uint16, 48, "B", [(0, 0)], 0.0, ((0.0, 0.0, 0.0), (1.5, 1.5, 1.5))),  # max_slice > 1: pos == 1 samples
]


@pytest.mark.parametrize("case", range(len(CASES)))
@pytest.mark.parametrize("ess", [0, 1])
def test_exact_sampler_is_bit_identical(vr, port, case, ess):
    kind, dtype, n, tfn, moves, zoom, slices = CASES[case]
    vol = small_volume(port, kind, dtype, n)
    lut = tf(port, tfn)
    kw = dict(min_slice=slices[0], max_slice=slices[1]) if slices else {}
    img, raw, st, o = render_both(vr, port, vol, lut, moves, zoom, sampler=vr.SAMPLER_EXACT, empty_skip=ess, **kw)
    assert st.covered_pixels == o["stats"].covered_pixels
    assert st.n_ref == o["stats"].n_ref
    assert st.n_inslice == o["stats"].n_inslice
    assert st.n_fetch <= st.n_inslice
    if not ess:
        assert st.n_fetch == st.n_inslice
    assert np.array_equal(bits(raw), bits(o["raw"])), "shader output differs from the oracle"
    assert np.array_equal(img, o["rgba8"])


@pytest.mark.parametrize("case", range(len(CASES)))
@pytest.mark.parametrize("layout", ["array3d", "paired"])
def test_hw_sampler_within_tolerance(vr, port, case, layout):
    """Both volume layouts of the hardware sampler (vr_layout): the 3-D array (one trilinear fetch) and the
    paired 2-D layered array (one bilinear fetch + fp32 z-lerp) meet the same bar against the oracle."""
    kind, dtype, n, tfn, moves, zoom, slices = CASES[case]
    if layout == "paired" and dtype == np.float32:
        pytest.skip("f32 volumes have no paired layout")
    vol = small_volume(port, kind, dtype, n)
    lut = tf(port, tfn)
    kw = dict(min_slice=slices[0], max_slice=slices[1]) if slices else {}
    lay = vr.LAYOUT_PAIRED if layout == "paired" else vr.LAYOUT_ARRAY3D
    img, raw, st, o = render_both(vr, port, vol, lut, moves, zoom, sampler=vr.SAMPLER_HW, empty_skip=1, layout=lay, **kw)
    assert st.covered_pixels == o["stats"].covered_pixels
    # positions are bit-exact in both modes except after an empty-space jump (see skip_empty in march.cu)
    assert abs(st.n_ref - o["stats"].n_ref) <= 1e-3 * o["stats"].n_ref
    assert_within_tolerance(img, o["rgba8"], f"case {case}")


# one camera per sign octant of the view direction (the distance maps are per octant), plus axis-aligned views
ESS_VIEWS = [(45, 30), (135, 30), (225, 30), (315, 30), (45, -30), (135, -30), (225, -30), (315, -30), (0, 0), (90, 0), (0, 89)]


@pytest.mark.parametrize("layout", ["array3d", "paired"])
@pytest.mark.parametrize("view", ESS_VIEWS)
def test_empty_space_skipping_skips_and_changes_nothing(vr, port, layout, view):
    """Hardware sampler: a jump of k steps is one fused multiply-add per axis where the shader makes k
    additions, so the positions after a jump differ from the reference's by the rounding those additions
    accumulate (<= 2e-5 texture units): the frame changes by less than one 8-bit level, the visited-step
    count by at most the rays whose last step sits within that distance of the exit face."""
    vol = small_volume(port, 2, np.uint16, 64)
    lut = tf(port, "C")
    lay = vr.LAYOUT_PAIRED if layout == "paired" else vr.LAYOUT_ARRAY3D
    a = render_both(vr, port, vol, lut, [view], sampler=vr.SAMPLER_HW, empty_skip=0, layout=lay)
    b = render_both(vr, port, vol, lut, [view], sampler=vr.SAMPLER_HW, empty_skip=1, layout=lay)
    assert np.abs(a[1] - b[1]).max() < 1.0 / 255.0
    assert np.abs(a[0].astype(int) - b[0].astype(int)).max() <= 1
    assert (a[0] != b[0]).mean() < 0.01
    assert a[2].covered_pixels == b[2].covered_pixels
    assert abs(a[2].n_ref - b[2].n_ref) <= 1e-3 * a[2].n_ref
    assert b[2].n_fetch < a[2].n_fetch * 0.7, (a[2].n_fetch, b[2].n_fetch)
    assert_within_tolerance(b[0], b[3]["rgba8"], f"view {view}")


@pytest.mark.parametrize("view", ESS_VIEWS[:8])
def test_empty_space_skipping_exact_sampler_changes_nothing(vr, port, view):
    """The exact sampler consults the same per-octant maps step by step and never jumps: bit-identical."""
    vol = small_volume(port, 2, np.uint16, 64)
    lut = tf(port, "C")
    a = render_both(vr, port, vol, lut, [view], sampler=vr.SAMPLER_EXACT, empty_skip=0)
    b = render_both(vr, port, vol, lut, [view], sampler=vr.SAMPLER_EXACT, empty_skip=1)
    assert np.array_equal(bits(a[1]), bits(b[1]))
    assert np.array_equal(a[0], b[0])
    assert a[2].n_ref == b[2].n_ref and a[2].n_inslice == b[2].n_inslice
    assert b[2].n_fetch < a[2].n_fetch * 0.7, (a[2].n_fetch, b[2].n_fetch)


def test_early_termination_error_is_bounded(vr, port):
    vol = small_volume(port, 2, np.uint16, 64)
    lut = tf(port, "C")
    eps = 1.0 / 1024.0
    a = render_both(vr, port, vol, lut, [(45, 30)], sampler=vr.SAMPLER_EXACT, empty_skip=0)
    b = render_both(vr, port, vol, lut, [(45, 30)], sampler=vr.SAMPLER_EXACT, empty_skip=0, early_term=1, early_term_eps=eps)
    assert np.abs(a[1] - b[1]).max() < eps
    assert np.abs(a[0].astype(int) - b[0].astype(int)).max() <= 1
    assert b[2].n_fetch < a[2].n_fetch


@pytest.mark.parametrize("layout", ["array3d", "paired"])
def test_hw_early_termination_error_is_bounded(vr, port, layout):
    """The hardware path tests the threshold once per batch (and once per double-batch in the unchecked
    phase): it composites a few more real samples than the per-sample test would, never fewer."""
    vol = small_volume(port, 1, np.uint16, 64)
    lut = tf(port, "B")
    eps = 1.0 / 256.0
    lay = vr.LAYOUT_PAIRED if layout == "paired" else vr.LAYOUT_ARRAY3D
    a = render_both(vr, port, vol, lut, [(45, 30)], sampler=vr.SAMPLER_HW, empty_skip=0, layout=lay)
    b = render_both(vr, port, vol, lut, [(45, 30)], sampler=vr.SAMPLER_HW, empty_skip=0, early_term=1, early_term_eps=eps, layout=lay)
    assert np.abs(a[1] - b[1]).max() < eps
    assert b[2].n_fetch <= a[2].n_fetch


def test_layouts_agree_and_paired_needs_integer_texels(vr, port):
    """The two layouts sample the same voxels with different fixed-point weights (x, y, z: 8 fractional
    bits in hardware for the 3-D array; x, y in hardware and z in fp32 for the pairs)."""
    vol = small_volume(port, 1, np.uint16, 48)
    lut = tf(port, "B")
    a = render_both(vr, port, vol, lut, [(-120, 80)], sampler=vr.SAMPLER_HW, layout=vr.LAYOUT_ARRAY3D)
    b = render_both(vr, port, vol, lut, [(-120, 80)], sampler=vr.SAMPLER_HW, layout=vr.LAYOUT_PAIRED)
    assert np.abs(a[0].astype(int) - b[0].astype(int)).max() <= 2
    assert a[2].n_ref == b[2].n_ref and a[2].n_fetch == b[2].n_fetch
    op = vr.OffscreenPass(32, 32)
    try:
        op.set_volume_layout(vr.LAYOUT_PAIRED)
        with pytest.raises(vr.VrError):
            op.volume_dataset_changed(vr.Dataset(np.zeros((4, 4, 4), np.float32), 0.0, 1.0))
        op.set_volume_layout(vr.LAYOUT_AUTO)
        op.volume_dataset_changed(vr.Dataset(np.zeros((4, 4, 4), np.float32), 0.0, 1.0))
        assert op.volume_layout() == vr.LAYOUT_ARRAY3D
        op.volume_dataset_changed(vr.Dataset(np.zeros((4, 4, 4), np.uint16), 0.0, 1.0))
        assert op.volume_layout() == vr.LAYOUT_PAIRED
        op.set_volume_layout(vr.LAYOUT_ARRAY3D)
        op.volume_dataset_changed(vr.Dataset(np.zeros((4, 4, 4), np.uint16), 0.0, 1.0))
        assert op.volume_layout() == vr.LAYOUT_ARRAY3D
    finally:
        op.close()


@pytest.mark.parametrize("moves", [[], [(360, 0)], [(40, 320)], [(-300, -60)]])
@pytest.mark.parametrize("axis", [0, 1, 2])
@pytest.mark.parametrize("ess", [0, 1])
def test_every_pair_axis_with_every_camera(vr, port, moves, axis, ess, monkeypatch):
    """Pairs along x, y or z: the march works in a permuted sampling space (slices, occupancy grid and
    bounds follow).  Non-cubic volume, asymmetric slices; the axis is forced (VR_PAIR_AXIS is read when
    the context is created), the policy is tested below."""
    monkeypatch.setenv("VR_PAIR_AXIS", str(axis))
    vol = port.synth(2 if ess else 1, 0x5EED0042, np.uint16, (56, 24, 40))  # (nx, ny, nz)
    lut = tf(port, "C" if ess else "B")
    cam, ocam = twin_cameras(vr, port, moves, 0.0)
    mn, mx = float(vol.min()), float(vol.max())
    lo, hi = (0.1, 0.0, 0.2), (0.95, 0.8, 1.0)
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(write_raw=1, collect_stats=1, sampler=vr.SAMPLER_HW, empty_skip=ess)
        op.set_volume_layout(vr.LAYOUT_PAIRED)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(lut)
        op.slicing_changed(lo, hi)
        op.set_camera(cam)
        op.record()
        img, st = op.download(), op.stats()
        assert op.last_frame_pair_axis() == axis
    finally:
        op.close()
    o = port.render(oracle_frame(port, ocam, W, H, mn, mx, lo, hi), vol, lut)
    assert st.covered_pixels == o["stats"].covered_pixels
    if ess:  # positions after an empty-space jump differ from the shader's by a few ulp (skip_empty, march.cu)
        assert abs(st.n_ref - o["stats"].n_ref) <= 1e-3 * o["stats"].n_ref
        assert abs(st.n_inslice - o["stats"].n_inslice) <= 1e-3 * o["stats"].n_ref
    else:
        assert st.n_ref == o["stats"].n_ref and st.n_inslice == o["stats"].n_inslice
    assert_within_tolerance(img, o["rgba8"], f"pair axis {axis}")


@pytest.mark.parametrize("split", [25, 30])
def test_pair_copies_split_over_two_layered_arrays(vr, port, split, monkeypatch):
    """Axes longer than 2047 voxels need two layered arrays per copy (2048-layer limit); VR_PAIR_SPLIT forces
    the split on a small volume.  The frames must not change by a bit."""
    vol = small_volume(port, 1, np.uint16, 48)
    lut = tf(port, "B")
    frames = {}
    for s in (2048, split):
        monkeypatch.setenv("VR_PAIR_SPLIT", str(s))
        for moves in ([], [(360, 0)], [(40, 320)]):
            frames[(s, str(moves))] = render_both(vr, port, vol, lut, moves, sampler=vr.SAMPLER_HW, layout=vr.LAYOUT_PAIRED)[1]
    for moves in ([], [(360, 0)], [(40, 320)]):
        assert np.array_equal(bits(frames[(2048, str(moves))]), bits(frames[(split, str(moves))])), moves


def test_pair_axis_policy_prefers_the_axis_the_camera_looks_down(vr, port):
    """vr_render_frame's cost estimate: for a view along a volume axis the pairs run along that axis (one
    layer per warp fetch, fewest layers read)."""
    vol = small_volume(port, 1, np.uint16, 64)
    op = vr.OffscreenPass(W, H)
    try:
        op.set_volume_layout(vr.LAYOUT_PAIRED)
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "B"))
        for moves, axis in (([], 1), ([(360, 0)], 0), ([(0, 350)], 2)):
            cam, _ = twin_cameras(vr, port, moves, 0.0)
            op.set_camera(cam)
            op.record()
            op.sync()
            assert op.last_frame_pair_axis() == axis, moves
    finally:
        op.close()


def test_auto_layout_is_chosen_per_frame(vr, port):
    """AUTO keeps both layouts resident and samples the pairs only when the cube's screen bounding box
    fills the GPU (vr_render_frame); the result meets the bar either way."""
    vol = small_volume(port, 1, np.uint16, 64)
    lut = tf(port, "B")
    cam, ocam = twin_cameras(vr, port, [(30, 20)], 0.0)
    for (w, h), want in (((160, 120), False), ((1280, 960), True)):
        op = vr.OffscreenPass(w, h)
        try:
            op.set_options(empty_skip=0)
            op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
            op.transfer_function_changed(lut)
            op.set_camera(cam)
            op.record()
            img = op.download()
            assert op.volume_layout() == vr.LAYOUT_PAIRED
            assert op.last_frame_paired() == want, (w, h)
        finally:
            op.close()
        o = port.render(oracle_frame(port, ocam, w, h, float(vol.min()), float(vol.max()), (0, 0, 0), (1, 1, 1)), vol, lut)
        assert_within_tolerance(img, o["rgba8"], f"auto {w}x{h}")


def _write_nrrd(path, arr, endian="little", detached=None, byte_skip=None, prefix=b""):
    tname = {np.dtype(np.uint8): "uchar", np.dtype(np.uint16): "ushort", np.dtype(np.float32): "float"}[arr.dtype]
    nz, ny, nx = arr.shape
    hdr = f"NRRD0004\n# test\ntype: {tname}\ndimension: 3\nsizes: {nx} {ny} {nz}\nencoding: raw\n"
    if arr.dtype.itemsize > 1:
        hdr += f"endian: {endian}\n"
    if byte_skip is not None:
        hdr += f"byte skip: {byte_skip}\n"
    data = arr.astype(arr.dtype.newbyteorder(">" if endian == "big" else "<")).tobytes()
    if detached:
        hdr += f"data file: {detached}\n"
        (path.parent / detached).write_bytes(prefix + data)
        path.write_bytes(hdr.encode())
    else:
        path.write_bytes(hdr.encode() + b"\n" + prefix + data)


def test_nrrd_loader_fast_path_and_general_path_agree(vr, port, tmp_path):
    """vr_load_nrrd streams raw u8/u16/f32 payloads in host byte order through pinned memory straight to the
    device; everything else goes through the general reader.  Either way the device holds the file's
    voxels and the data range (nrrd_file_parser.cpp:39-40)."""
    rng = np.random.default_rng(7)
    cases = [
        ("a.nrrd", rng.integers(0, 65535, (9, 20, 33), dtype=np.uint16), dict()),
        ("b.nhdr", rng.integers(0, 255, (17, 5, 12), dtype=np.uint8), dict(detached="b.raw", byte_skip=-1, prefix=b"junkjunk")),
        ("c.nrrd", rng.random((6, 7, 8), dtype=np.float32), dict(byte_skip=5, prefix=b"12345")),
        ("d.nrrd", rng.integers(0, 65535, (4, 6, 10), dtype=np.uint16), dict(endian="big")),   # general path (byte swap)
    ]
    op = vr.OffscreenPass(32, 32)
    try:
        for name, arr, kw in cases:
            _write_nrrd(tmp_path / name, arr, **kw)
            op.import_nrrd(str(tmp_path / name))
            dt, dims, mn, mx = op.volume_info()
            assert dims == (arr.shape[2], arr.shape[1], arr.shape[0])
            assert np.array_equal(op.download_volume(), arr), name
            assert (mn, mx) == (float(arr.min()), float(arr.max())), name
            assert np.array_equal(vr.read_nrrd(str(tmp_path / name)), arr), name
        (tmp_path / "t.nrrd").write_bytes((tmp_path / "a.nrrd").read_bytes()[:-10])
        with pytest.raises(vr.VrError, match="truncated"):
            op.import_nrrd(str(tmp_path / "t.nrrd"))
    finally:
        op.close()


def test_write_image_round_trips(vr, port, tmp_path):
    op = vr.OffscreenPass(37, 21)
    try:
        op.synth_volume(vr.SYNTH_MARSCHNER_LOBB, 1, vr.U8, 16)
        op.transfer_function_changed(tf(port, "A"))
        op.set_camera(vr.Camera().rotate((-120.0, 80.0)))
        op.record()
        img = op.download()
        op.write_image(tmp_path / "f.ppm")
        op.write_image(tmp_path / "f.pam")
        with pytest.raises(vr.VrError):
            op.write_image(tmp_path / "no_such_dir" / "f.ppm")
    finally:
        op.close()
    raw = (tmp_path / "f.ppm").read_bytes()
    assert raw.startswith(b"P6\n37 21\n255\n")
    assert np.array_equal(np.frombuffer(raw[len(b"P6\n37 21\n255\n"):], np.uint8).reshape(21, 37, 3), img[..., :3])
    raw = (tmp_path / "f.pam").read_bytes()
    body = raw[raw.index(b"ENDHDR\n") + 7:]
    assert np.array_equal(np.frombuffer(body, np.uint8).reshape(21, 37, 4), img)


def test_startup_state_matches_reference_defaults(vr, port):
    """1x1x1 zero volume + one white opaque texel (offscreen_pass.cpp:118-119): every covered pixel
    is white; min == max divides by zero in volume.frag:42 (NaN -> texel 0 by the oracle's definition)."""
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(write_raw=1)
        op.record()
        img = op.download()
    finally:
        op.close()
    cam = port.camera()
    o = port.render(oracle_frame(port, cam, W, H, 0.0, 0.0), np.zeros((1, 1, 1), np.float32), tf(port, "0"), want_raw=True)
    assert np.array_equal(img, o["rgba8"])
    covered = o["raw"][..., 3] > 0
    assert covered.any() and (img[covered] == 255).all() and (img[~covered] == (28, 28, 28, 255)).all()


def test_resize_and_errors(vr, port):
    op = vr.OffscreenPass(64, 48)
    try:
        op.framebuffer_size_changed(0, 10)  # ignored like offscreen_pass.cpp:237-239
        assert op.size == (64, 48)
        op.framebuffer_size_changed(96, 80)
        assert op.size == (96, 80)
        op.set_camera(vr.Camera())
        op.record()
        assert op.download().shape == (80, 96, 4)
        with pytest.raises(vr.VrError):
            op.transfer_function_changed(np.zeros(0, np.uint32))
        with pytest.raises(vr.VrError):
            op.set_march(1.8, 0.0)
        with pytest.raises(vr.VrError):
            op.set_uniforms(np.zeros(16), np.zeros(16), np.zeros(3))  # singular proj*view
        with pytest.raises(vr.VrError):
            op.import_nrrd("/nonexistent/file.nrrd")
        with pytest.raises(vr.VrError):
            op.download_raw()  # write_raw is off
    finally:
        op.close()


def test_step_size_parameter(vr, port):
    vol = small_volume(port, 1, np.uint16, 48)
    lut = tf(port, "B")
    cam, ocam = twin_cameras(vr, port, [(30, 20)])
    mn, mx = float(vol.min()), float(vol.max())
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(write_raw=1, sampler=vr.SAMPLER_EXACT)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(lut)
        op.set_march(1.0, 0.0125)
        op.set_camera(cam)
        op.record()
        raw = op.download_raw()
    finally:
        op.close()
    f = oracle_frame(port, ocam, W, H, mn, mx)
    f.ray_dist, f.step_size = 1.0, 0.0125
    o = port.render(f, vol, lut, want_raw=True)
    assert np.array_equal(bits(raw), bits(o["raw"]))


def test_device_minmax_and_synth_match_cpu(vr, port):
    op = vr.OffscreenPass(32, 32)
    try:
        for kind, dt, code in [(1, np.uint16, vr.U16), (0, np.uint8, vr.U8), (1, np.float32, vr.F32)]:
            op.synth_volume(kind, 0x5EED0000 + kind, code, 48)
            g = op.download_volume()
            c = port.synth(kind, 0x5EED0000 + kind, dt, 48)
            _, dims, mn, mx = op.volume_info()
            assert dims == (48, 48, 48)
            assert mn == float(g.min()) and mx == float(g.max())
            if kind == 1:
                assert np.array_equal(g, c), "fBm generator is pure fp32 arithmetic: must be bit-identical"
            else:
                assert np.abs(g.astype(np.int64) - c.astype(np.int64)).max() <= 1  # libm vs device sin/cos
    finally:
        op.close()


# ---- the committed golden vectors (generated from the reference's own sources) -------------------

import glob
import os

GOLDEN = sorted(g for g in glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz"))
                if not g.endswith("host_surfaces.npz"))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_path_reproduces_golden(vr, path):
    g = np.load(path)
    w, h = int(g["width"]), int(g["height"])
    op = vr.OffscreenPass(w, h)
    try:
        op.volume_dataset_changed(vr.Dataset(np.ascontiguousarray(g["volume"]), float(g["min_density"]), float(g["max_density"])))
        op.transfer_function_changed(g["lut"])
        op.slicing_changed(g["min_slice"], g["max_slice"])
        op.set_uniforms(g["view"], g["proj"], g["camera_position"])
        op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1, collect_stats=1)
        op.record()
        raw, img, st = op.download_raw(), op.download(), op.stats()
        assert np.array_equal(bits(raw), bits(g["raw"]))
        assert np.array_equal(img, g["rgba8"])
        assert [st.covered_pixels, st.n_ref, st.n_inslice] == g["stats"].tolist()
        op.set_options(sampler=vr.SAMPLER_HW)
        op.record()
        assert_within_tolerance(op.download(), g["rgba8"], os.path.basename(path))
    finally:
        op.close()


def test_frames_in_flight_do_not_interfere(vr, port):
    """Two frames recorded on two streams into two targets equal the same frames rendered one by one."""
    import torch
    lut = tf(port, "B")
    op = vr.OffscreenPass(W, H)
    try:
        op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 64)
        op.transfer_function_changed(lut)
        cams = [vr.Camera().rotate((40.0 * k, 10.0 * k)) for k in range(4)]
        want = []
        for c in cams:
            op.set_camera(c)
            op.record()
            want.append(op.download())
        streams = [torch.cuda.Stream() for _ in range(4)]
        hosts = [torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True) for _ in range(4)]
        for k, c in enumerate(cams):
            op.set_camera(c)
            op.record(streams[k].cuda_stream, k)
            op.download_async(hosts[k].data_ptr(), streams[k].cuda_stream, k)
        torch.cuda.synchronize()
        for k in range(4):
            assert np.array_equal(hosts[k].numpy(), want[k])
    finally:
        op.close()


def test_csv_import_renders_like_the_same_dataset_set_directly(vr, port, tmp_path):
    """vr_load_csv = CsvFileParser::parse + volume_dataset_changed (importer.cpp:30-46): float voxels, the
    reference's min <= 0 <= max as the density range."""
    vol = small_volume(port, 1, np.uint16, 24).astype(np.float32) / 7.0 + 3.0
    paths = []
    for z, sl in enumerate(vol):
        p = tmp_path / f"z{z:02d}.csv"
        p.write_text("\n".join(",".join("%.9g" % v for v in row) for row in sl) + "\n")
        paths.append(str(p))
    lut = tf(port, "B")
    cam, ocam = twin_cameras(vr, port, [(30, 20)])
    imgs = []
    for how in ("csv", "direct"):
        op = vr.OffscreenPass(W, H)
        try:
            op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1)
            if how == "csv":
                op.load_csv(paths)
                dtype, dims, mn, mx = op.volume_info()
                assert dtype == vr.F32 and tuple(dims) == (24, 24, 24)
                assert mn == 0.0 and mx == float(vol.max())
            else:
                op.volume_dataset_changed(vr.Dataset(vol, 0.0, float(vol.max())))
            op.transfer_function_changed(lut)
            op.set_camera(cam)
            op.record()
            imgs.append((op.download(), op.download_raw()))
        finally:
            op.close()
    assert np.array_equal(imgs[0][0], imgs[1][0]) and np.array_equal(bits(imgs[0][1]), bits(imgs[1][1]))
    o = port.render(oracle_frame(port, ocam, W, H, 0.0, float(vol.max()), (0, 0, 0), (1, 1, 1)), vol, lut, want_raw=True)
    assert np.array_equal(bits(imgs[0][1]), bits(o["raw"]))


@pytest.mark.parametrize("size", [(W, H), (2 * W, 2 * H), (W // 2 + 3, H // 3), (1, 1), (W + 1, H), (3 * W - 1, 7)])
def test_present_resamples_like_the_viewport_sampler(vr, port, size):
    """K3 (present.cu): LINEAR / CLAMP_TO_EDGE sampling of the RGBA8 frame (offscreen_pass.cpp:361-375),
    bit-exact against the oracle's restatement; same size in and out is the identity."""
    import torch
    vol = small_volume(port, 1, np.uint16, 32)
    cam, _ = twin_cameras(vr, port, [(25, 40)])
    op = vr.OffscreenPass(W, H)
    try:
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "B"))
        op.set_camera(cam)
        op.record()
        frame = op.download()
        dw, dh = size
        dst = torch.zeros((dh, dw, 4), dtype=torch.uint8, device="cuda")
        op.present(dst.data_ptr(), dw, dh)
        torch.cuda.synchronize()
        got = dst.cpu().numpy()
        want = port.present_bilinear(frame, dw, dh)
        assert np.array_equal(got, want)
        if size == (W, H):
            assert np.array_equal(got, frame)
        # pitched destination: rows 16 pixels apart wider than the image
        pitched = torch.zeros((dh, dw + 16, 4), dtype=torch.uint8, device="cuda")
        op.present(pitched.data_ptr(), dw, dh, pitch_bytes=(dw + 16) * 4)
        torch.cuda.synchronize()
        assert np.array_equal(pitched[:, :dw].cpu().numpy(), want) and int(pitched[:, dw:].max()) == 0
        if dw > 1:  # a pitch of 0 means tightly packed
            with pytest.raises(vr.VrError):
                op.present(dst.data_ptr(), dw, dh, pitch_bytes=4 * dw - 4)
        op.present(0, 0, 0)  # zero size: no-op, like framebuffer_size_changed(0, 0)
    finally:
        op.close()


def test_dirty_rectangle_download_keeps_the_host_frame_complete(vr, port):
    """vr_download_frame_dirty_async: after one full delivery into a host buffer only the union of the cube's screen
    boxes (this frame, the frame delivered there before) is copied; the host image always equals a full download."""
    import torch
    vol = small_volume(port, 1, np.uint16, 32)
    op = vr.OffscreenPass(320, 200)
    try:
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "B"))
        host = [torch.zeros((200, 320, 4), dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        streams = [torch.cuda.Stream() for _ in range(2)]
        # orbit, zoom far out (small box), zoom in until the cube leaves the frame (box = frame), back out
        moves = [((30 * i, 10 * i), z) for i, z in enumerate([0.0, 0.0, -3.0, -3.0, -3.0, 0.0, 2.0, 2.3, 2.3, 0.0, -1.0, -1.0])]
        sizes = []
        for j, (mv, zoom) in enumerate(moves):
            f = j % 2
            cam, _ = twin_cameras(vr, port, [mv], zoom)
            op.set_camera(cam)
            op.record(streams[f].cuda_stream, f)
            n = op.download_dirty_async(host[f].data_ptr(), streams[f].cuda_stream, f)
            streams[f].synchronize()
            want = op.download_frame(f)
            assert np.array_equal(host[f].numpy(), want), f"frame {j}"
            sizes.append(n)
        full = 320 * 200 * 4
        assert sizes[0] == full and sizes[1] == full          # first delivery into each buffer
        assert 0 < sizes[2] < full and 0 < sizes[4] < sizes[2]  # far away: a small box; two small boxes: smaller still
        assert max(sizes) <= full
        # a resize forgets every buffer
        op.framebuffer_size_changed(160, 100)
        small = torch.zeros((100, 160, 4), dtype=torch.uint8, pin_memory=True)
        op.record(streams[0].cuda_stream, 0)
        assert op.download_dirty_async(small.data_ptr(), streams[0].cuda_stream, 0) == 160 * 100 * 4
        streams[0].synchronize()
        assert np.array_equal(small.numpy(), op.download_frame(0))
    finally:
        op.close()


def test_failed_upload_leaves_the_context_usable(vr, port):
    """an upload that is refused (paired layout asked for a float volume) must not leave a half-installed
    volume behind: the old volume keeps rendering, bit for bit."""
    vol = small_volume(port, 1, np.uint16, 32)
    op = vr.OffscreenPass(W, H)
    try:
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "C"))   # TF-C: empty-space skipping is active on this volume or not -- both paths below
        cam, _ = twin_cameras(vr, port, [(-120, 80)])
        op.set_camera(cam)
        op.record()
        before = op.download()
        op.set_volume_layout(vr.LAYOUT_PAIRED)
        with pytest.raises(vr.VrError):
            op.volume_dataset_changed(vr.Dataset(np.zeros((8, 8, 8), np.float32), 0.0, 1.0))
        for ess in (1, 0):
            op.set_options(empty_skip=ess)
            op.record()
            assert np.array_equal(op.download(), before)
        assert op.volume_info()[1] == (32, 32, 32)
        # and a later good upload works
        op.set_volume_layout(vr.LAYOUT_AUTO)
        op.volume_dataset_changed(vr.Dataset(np.zeros((8, 8, 8), np.float32), 0.0, 1.0))
        op.record()
        op.download()
    finally:
        op.close()


def test_dirty_clear_leaves_the_frames_unchanged(vr, port, monkeypatch):
    """A frame restores the clear colour only where the previous frame into the same target drew outside its own
    tiles (offscreen_pass.cpp:170-173 clears everything every frame): every frame -- RGBA8 and the float4 output --
    must equal the one a context that clears the whole target gives, across orbit, zoom, slot reuse and resize."""
    vol = small_volume(port, 1, np.uint16, 32)
    moves = [((30 * i, 10 * i), z) for i, z in enumerate([0.0, 0.0, -3.0, -3.0, -3.0, 0.0, 2.0, 2.3, 2.3, 0.0, -1.0, -1.0, 2.6, 0.0])]

    def run(dirty):
        monkeypatch.setenv("VR_DIRTY_CLEAR", "1" if dirty else "0")
        op = vr.OffscreenPass(320, 200)
        frames, blocks = [], []
        try:
            op.set_options(write_raw=1)
            op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
            op.transfer_function_changed(tf(port, "B"))
            for j, (mv, zoom) in enumerate(moves):
                f = j % 2
                cam, _ = twin_cameras(vr, port, [mv], zoom)
                op.set_camera(cam)
                op.record(None, f)
                blocks.append(op.last_frame_info().clear_blocks)
                frames.append(op.download_frame(f))
                if f == 0:
                    frames.append(op.download_raw())
            op.framebuffer_size_changed(160, 100)
            for f in (0, 1, 0):
                op.record(None, f)
                blocks.append(op.last_frame_info().clear_blocks)
                frames.append(op.download_frame(f))
        finally:
            op.close()
        return frames, blocks

    want, full_blocks = run(False)
    got, blocks = run(True)
    assert len(want) == len(got)
    for k, (a, b) in enumerate(zip(want, got)):
        assert np.array_equal(a, b), f"output {k}"
    assert all(b > 0 for b in full_blocks)
    assert blocks[0] > 0 and blocks[1] > 0            # first frame into each target: everything
    assert blocks[4] == 0                             # same small box again in the same slot: nothing to restore
    assert blocks[5] == 0                             # far -> near: the new tiles cover the old ones
    assert 0 < blocks[11] < full_blocks[11]           # near -> far restores the rim between the two boxes only
    assert blocks[-3] > 0 and blocks[-2] > 0 and blocks[-1] == 0   # new targets after the resize


@pytest.mark.parametrize("sampler", ["exact", "hw"])
def test_non_finite_voxels_are_not_skipped(vr, port, sampler):
    """a NaN density samples LUT texel 0, +-Inf the end texels (volume.frag:42-43 through the CLAMP_TO_EDGE
    sampler).  fminf / fmaxf drop NaNs, so a cell of NaN voxels used to get an empty range and was skipped even when
    texel 0 is visible: frames with and without empty-space skipping must agree."""
    rng = np.random.default_rng(5)
    vol = np.full((32, 32, 32), 0.5, np.float32)   # background: a transparent texel, so that most cells are empty
    vol[4:12, 4:12, 4:12] = np.nan
    vol[20:28, 6:10, 18:30] = np.inf
    vol[14:18, 20:30, 2:9] = -np.inf
    vol[22, 22, 22] = 0.7
    lut = tf(port, "B").copy()
    lut[0] = 0x80FF4020      # texel 0 visible: alpha 0x80
    lut[255] = 0x60102030
    lut[1:200] &= 0x00FFFFFF  # everything between is transparent
    frames = []
    for ess in (0, 1):
        op = vr.OffscreenPass(W, H)
        try:
            op.set_options(sampler=vr.SAMPLER_EXACT if sampler == "exact" else vr.SAMPLER_HW, empty_skip=ess, write_raw=1)
            op.volume_dataset_changed(vr.Dataset(vol, 0.0, 1.0))
            op.transfer_function_changed(lut)
            cam, _ = twin_cameras(vr, port, [(-120, 80)])
            op.set_camera(cam)
            op.record()
            frames.append((op.download(), op.download_raw()))
            if ess:
                assert op.last_frame_info().empty_skip_active == 1
        finally:
            op.close()
    assert frames[0][0].astype(int).max() > 40, "the non-finite regions must be visible at all"
    if sampler == "exact":
        assert np.array_equal(frames[0][1].view(np.uint32), frames[1][1].view(np.uint32))
    assert np.abs(frames[0][0].astype(int) - frames[1][0].astype(int)).max() <= (0 if sampler == "exact" else 1)


def test_full_downloads_keep_the_dirty_rectangle_bookkeeping_right(vr, port):
    """a full download into a buffer the dirty path knows must update what is remembered about it, and a
    buffer can be forgotten (an allocator may hand the same address out again with other contents)."""
    import torch
    vol = small_volume(port, 1, np.uint16, 32)
    op = vr.OffscreenPass(320, 200)
    try:
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "B"))
        host = torch.zeros((200, 320, 4), dtype=torch.uint8, pin_memory=True)
        s = torch.cuda.Stream()

        def frame(zoom, how):
            cam, _ = twin_cameras(vr, port, [(30, 10)], zoom)
            op.set_camera(cam)
            op.record(s.cuda_stream, 0)
            n = how(host.data_ptr(), s.cuda_stream, 0)
            s.synchronize()
            assert np.array_equal(host.numpy(), op.download_frame(0))
            return n

        frame(0.0, op.download_dirty_async)                 # first delivery: full
        frame(2.3, op.download_async)                       # a full download of a frame that fills the target ...
        assert frame(-3.0, op.download_dirty_async) > 0.5 * 320 * 200 * 4   # ... so the next dirty copy must cover it
        host.fill_(7)                                       # someone else scribbles over the buffer
        op.forget_host_buffer(host.data_ptr())
        assert frame(-3.0, op.download_dirty_async) == 320 * 200 * 4
    finally:
        op.close()


MASK_VIEWS = [([(45 * 4, 30 * 4)], 0.0), ([(135 * 4, -30 * 4)], 0.0), ([(0, 0)], 0.0), ([(0, 89 * 4)], 0.0), ([(225 * 4, 30 * 4)], 2.0),
              ([(315 * 4, -30 * 4)], 2.7), ([(90 * 4, 0)], -3.0)]


@pytest.mark.parametrize("layout,mode", [("array3d", "1"), ("paired", "1"), ("array3d", "2")])
def test_screen_space_tile_mask_changes_nothing(vr, port, layout, mode, monkeypatch):
    """with skipping on, tiles no occupied cell projects into get the clear colour without a ray being
    set up.  The frames (RGBA8 and the float4 before the blend) must be BIT-identical to the ones of a context without
    the mask, the mask must cover every pixel that drew anything, and on a sparse volume it must drop tiles."""
    vol = small_volume(port, 2, np.uint16, 64)
    lut = tf(port, "C")
    lay = vr.LAYOUT_PAIRED if layout == "paired" else vr.LAYOUT_ARRAY3D
    w, h = 480, 270

    def run(mask):
        monkeypatch.setenv("VR_TILE_MASK", mode if mask else "0")   # 2 = the path of frames too large to stage in shared memory
        op = vr.OffscreenPass(w, h)
        out = []
        try:
            op.set_volume_layout(lay)
            op.set_options(sampler=vr.SAMPLER_HW, write_raw=1, empty_skip=1, early_term=0)
            op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
            op.transfer_function_changed(lut)
            for j, (mv, zoom) in enumerate(MASK_VIEWS):
                cam, _ = twin_cameras(vr, port, mv, zoom)
                op.set_camera(cam)
                if j == 3:
                    op.slicing_changed((0.0, 0.2, 0.0), (0.7, 1.0, 1.0))
                op.record(None, j % 2)
                out.append((op.download_frame(j % 2), op.download_raw() if j % 2 == 0 else None, op.tile_mask(j % 2), op.last_frame_info()))
                if j == 3:
                    op.slicing_changed((0.0, 0.0, 0.0), (1.0, 1.0, 1.0))
            # statistics need every ray: no mask for that frame
            op.set_options(sampler=vr.SAMPLER_HW, write_raw=1, empty_skip=1, early_term=0, collect_stats=1)
            op.record()
            assert op.tile_mask(0) is None
        finally:
            op.close()
        return out

    want, got = run(False), run(True)
    dropped = 0
    for j, (a, b) in enumerate(zip(want, got)):
        assert a[2] is None and b[2] is not None
        assert np.array_equal(a[0], b[0]), f"view {j}: RGBA8 differs"
        if a[1] is not None:
            assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), f"view {j}: float4 differs"
        bits, ignored, th = b[2]
        if MASK_VIEWS[j][1] < 2.5:
            assert not ignored, f"view {j}"
        elif ignored:
            continue   # a cell behind the eye: the march ignored the mask (the frames above are identical all the same)
        clear = a[0][0, 0]
        drew = (a[0] != clear).any(axis=2)
        ys, xs = np.nonzero(drew)
        assert bits[ys // th, xs // 16].all(), f"view {j}: a pixel that drew lies in an unmarked tile"
        x0, y0, x1, y1 = b[3].bbox
        inside = bits[y0 // th:(y1 + th - 1) // th, x0 // 16:(x1 + 15) // 16]
        dropped += inside.size - int(inside.sum())
    assert dropped > 0, "the sparse volume must leave tiles of the cube's screen box unmarked"


def test_tile_mask_is_ignored_when_a_cell_is_behind_the_eye(vr, port):
    """Camera inside the near range of the cube: cells with w <= 0 cannot be projected -- the mask is built, flagged, and
    the frame equals the oracle's as without it."""
    vol = small_volume(port, 2, np.uint16, 32)
    lut = tf(port, "C")
    op = vr.OffscreenPass(320, 200)
    try:
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=0)
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(lut)
        cam, ocam = twin_cameras(vr, port, [(100, 60)], 2.9)
        for _ in range(40):
            cam.zoom(0.5)
            ocam.zoom(0.5)
        op.set_camera(cam)
        op.record()
        img = op.download()
        m = op.tile_mask(0)
        want = port.render(oracle_frame(port, ocam, 320, 200, float(vol.min()), float(vol.max())), vol, lut)
        assert_within_tolerance(img, want["rgba8"], "closest zoom")
        assert m is not None
    finally:
        op.close()


def test_gradient_discretised_on_the_device_equals_the_host_one(vr, port):
    """vr_set_gradient runs Gradient::discretize (gradient.cpp:90-108) on the device.  The texels must be
    bit-identical to vr_gradient_discretize's (itself checked against the reference's own gradient.cpp in test_host.py),
    for random marker sets incl. coincident locations and out-of-range values, and a frame rendered from them equal to
    the frame rendered from the host-made table."""
    rng = np.random.default_rng(11)
    vol = small_volume(port, 1, np.uint16, 32)
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1)
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        for trial in range(12):
            nc, na = int(rng.integers(1, 7)), int(rng.integers(1, 7))
            cl = np.sort(rng.random(nc)).astype(np.float32)
            al = np.sort(rng.random(na)).astype(np.float32)
            if trial % 4 == 3 and nc > 1:
                cl[1] = cl[0]                     # two markers at one location
            g = vr.Gradient([(float(l), tuple(float(x) for x in rng.uniform(-0.2, 1.2, 3))) for l in cl],
                            [(float(l), float(rng.uniform(-0.2, 1.2))) for l in al])
            count = int(rng.choice([1, 2, 7, 256, 1000]))
            with np.errstate(all="ignore"):
                want = g.discretize(count)
            op.gradient_changed(g, count)
            got = op.get_transfer_function()
            assert np.array_equal(got, want), f"trial {trial}: {(got != want).sum()} of {count} texels differ"
        g = vr.Gradient([(0.0, (0, 0, 0.5)), (0.5, (1, 0.5, 0)), (1.0, (1, 1, 1))], [(0.0, 0.0), (0.3, 0.0), (1.0, 0.08)])
        op.gradient_changed(g, 256)
        op.record()
        a = op.download_raw()
        op.transfer_function_changed(g.discretize(256))
        op.record()
        assert np.array_equal(a.view(np.uint32), op.download_raw().view(np.uint32))
        # texels handed over in device memory
        import torch
        t = torch.from_numpy(g.discretize(256).view(np.int32)).cuda()
        op.transfer_function_changed_device(t.data_ptr(), 256)
        op.record()
        assert np.array_equal(a.view(np.uint32), op.download_raw().view(np.uint32))
    finally:
        op.close()


# ---- pre-integrated classification (optional quality mode, no reference counterpart; oracle/preint.c) -----------------

PREINT_CASES = [
    # (volume kind, dtype, size, TF, texels, moves, zoom, slices, layout)
    (1, np.uint16, 48, "B", 256, [(-120, 80)], 0.0, None, "array3d"),
    (1, np.uint16, 48, "B", 256, [(-120, 80)], 0.0, None, "paired"),
    (2, np.uint16, 40, "C", 256, [(45, 30)], 0.0, None, "array3d"),
    (0, np.uint8, 32, "A", 64, [(37, -21)], 1.0, None, "paired"),
    (1, np.uint16, 48, "B", 256, [(200, 10)], 1.2, ((0.2, 0.1, 0.3), (0.9, 0.8, 0.7)), "array3d"),
    (1, np.float32, (33, 20, 7), "B", 100, [(10, 55)], 0.5, None, "array3d"),
]


@pytest.mark.parametrize("case", PREINT_CASES, ids=[f"{c[3]}-{c[1].__name__}-{c[8]}-{i}" for i, c in enumerate(PREINT_CASES)])
def test_preintegrated_mode_matches_its_oracle(vr, port, case):
    """vr_set_preintegration: the table built on the device equals the oracle's (double on the CPU) to 1e-5, and the frame
    is within the hardware-sampler tolerance of the oracle's pre-integrated frame (same positions and tests as the
    reference loop, segments classified from the table)."""
    kind, dt, size, tfn, texels, moves, zoom, slices, layout = case
    vol = small_volume(port, kind, dt, size)
    lut = tf(port, tfn, texels)
    mn, mx = float(vol.min()), float(vol.max())
    cam, ocam = twin_cameras(vr, port, moves, zoom)
    lo, hi = slices if slices else ((0, 0, 0), (1, 1, 1))
    op = vr.OffscreenPass(W, H)
    try:
        op.set_volume_layout(vr.LAYOUT_PAIRED if layout == "paired" and dt != np.float32 else vr.LAYOUT_ARRAY3D)
        op.set_options(sampler=vr.SAMPLER_HW, write_raw=1, empty_skip=1, early_term=0)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(lut)
        op.slicing_changed(lo, hi)
        op.set_camera(cam)
        op.record()
        plain = op.download()
        op.set_preintegration(8)
        tab = op.preintegration_table()
        want_tab = port.preint_table(lut, 8)
        assert tab.shape == want_tab.shape and np.abs(tab - want_tab).max() < 1e-5
        op.record()
        img, raw = op.download(), op.download_raw()
        want = port.render_preint(oracle_frame(port, ocam, W, H, mn, mx, lo, hi), vol, lut, 8)
        assert_within_tolerance(img, want["rgba8"], "pre-integrated frame")
        assert np.abs(raw - want["raw"]).max() < 2.5e-2
        # another sub-sample count rebuilds the table; switching the mode off gives the plain frame back, bit for bit
        op.set_preintegration(2)
        assert np.abs(op.preintegration_table() - port.preint_table(lut, 2)).max() < 1e-5
        op.set_preintegration(0)
        op.record()
        assert np.array_equal(op.download(), plain)
    finally:
        op.close()


def test_preintegration_sees_a_feature_the_samples_step_over(vr, port):
    """What the mode is for: a transfer function with a thin opaque feature and a coarse step (0.02: 90 steps).  Per-sample
    classification only sees the feature where a sample happens to fall on it; segments always cross it.  The
    pre-integrated frame carries clearly more opacity than the per-sample frame (the oracle says how much), and the mode
    follows a new transfer function (the table is rebuilt)."""
    vol = small_volume(port, 1, np.uint16, 48)
    mn, mx = float(vol.min()), float(vol.max())
    spike = np.zeros(256, np.uint32)
    spike[127], spike[128], spike[129] = 0x4020C0FF, 0x9920C0FF, 0x4020C0FF     # alpha 0.25 / 0.6 / 0.25, three texels wide
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(sampler=vr.SAMPLER_HW, write_raw=1, early_term=0)
        op.set_march(1.8, 0.02)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(spike)
        cam, ocam = twin_cameras(vr, port, [(-120, 80)])
        op.set_camera(cam)
        f = oracle_frame(port, ocam, W, H, mn, mx)
        f.step_size = 0.02
        want_plain = port.render(f, vol, spike, want_raw=True)["raw"][..., 3].sum()
        want = port.render_preint(f, vol, spike, 16)
        want_pre = want["raw"][..., 3].sum()
        assert want_pre > 1.1 * want_plain
        op.record()
        a = op.download_raw()[..., 3].sum()
        op.set_preintegration(16)
        op.record()
        got = op.download_raw()
        b = got[..., 3].sum()
        assert abs(a / want_plain - 1.0) < 1e-2 and abs(b / want_pre - 1.0) < 1e-2, (a, want_plain, b, want_pre)
        err = np.abs(got - want["raw"])
        assert np.percentile(err, 99.9) < 2e-2, (float(err.max()), float(np.percentile(err, 99.9)))
        op.transfer_function_changed(tf(port, "B"))
        op.record()
        want = port.render_preint(f, vol, tf(port, "B"), 16)
        assert_within_tolerance(op.download(), want["rgba8"], "after a transfer-function change")
    finally:
        op.close()


def test_preintegration_refuses_what_it_does_not_support(vr, port):
    vol = small_volume(port, 1, np.uint16, 32)
    op = vr.OffscreenPass(64, 48)
    try:
        op.volume_dataset_changed(vr.Dataset(vol, float(vol.min()), float(vol.max())))
        op.transfer_function_changed(tf(port, "B"))
        with pytest.raises(vr.VrError):
            op.set_preintegration(65)
        with pytest.raises(vr.VrError, match="off"):
            op.preintegration_table()
        op.set_preintegration(4)
        op.set_options(sampler=vr.SAMPLER_EXACT)
        with pytest.raises(vr.VrError, match="pre-integration needs"):
            op.record()
        op.set_options(sampler=vr.SAMPLER_HW, collect_stats=1)
        with pytest.raises(vr.VrError, match="pre-integration needs"):
            op.record()
        op.set_options(sampler=vr.SAMPLER_HW, collect_stats=0)
        op.transfer_function_changed(np.zeros(2048, np.uint32))
        with pytest.raises(vr.VrError, match="1024"):
            op.record()
        op.transfer_function_changed(tf(port, "B"))
        op.record()
        op.download()
    finally:
        op.close()


# ---- non-cubic voxels (vr_set_volume_spacing; the reference ignores a file's spacings; oracle/spacing.c) ---------------

SPACING_CASES = [
    # (dtype, dims (nx, ny, nz), spacing, moves, zoom, slices)
    (np.uint16, 32, (1.0, 2.0, 4.0), [(-120, 80)], 0.0, None),
    (np.uint16, 32, (3.0, 1.0, 1.5), [(200, 10)], 1.2, ((0.2, 0.1, 0.3), (0.9, 0.8, 0.7))),
    (np.float32, (33, 20, 7), (1.0, 1.0, 1.0), [(10, 55)], 0.5, None),       # non-cubic grid of cubic voxels
    (np.uint8, 32, (0.7, 0.7, 0.7), [(37, -21)], 0.0, None),                 # cubic: extents (1, 1, 1)
]


@pytest.mark.parametrize("case", SPACING_CASES, ids=[f"{c[0].__name__}-{i}" for i, c in enumerate(SPACING_CASES)])
def test_honoured_spacings_match_the_oracle(vr, port, case):
    """Exact sampler: float4 output, RGBA8 and counts BIT-identical to the CPU restatement (same fp32 operations); hardware
    sampler with and without skipping, both layouts: within tolerance.  Switching the spacing off gives the unit-cube
    frame back."""
    dt, size, spacing, moves, zoom, slices = case
    vol = small_volume(port, 1, dt, size)
    nz, ny, nx = vol.shape
    phys = np.array([nx * spacing[0], ny * spacing[1], nz * spacing[2]], np.float64)
    ext = (phys / phys.max()).astype(np.float32)
    lut = tf(port, "B")
    mn, mx = float(vol.min()), float(vol.max())
    cam, ocam = twin_cameras(vr, port, moves, zoom)
    lo, hi = slices if slices else ((0, 0, 0), (1, 1, 1))
    f = oracle_frame(port, ocam, W, H, mn, mx, lo, hi)
    want = port.render_spacing(f, ext, vol, lut)
    cube = port.render(f, vol, lut, want_raw=True)
    op = vr.OffscreenPass(W, H)
    try:
        op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1, collect_stats=1, early_term=0)
        op.volume_dataset_changed(vr.Dataset(vol, mn, mx))
        op.transfer_function_changed(lut)
        op.slicing_changed(lo, hi)
        op.set_camera(cam)
        op.set_volume_spacing(spacing)
        _, e = op.volume_spacing()
        assert np.array_equal(e, ext)
        op.record()
        raw, img, st = op.download_raw(), op.download(), op.stats()
        assert np.array_equal(raw.view(np.uint32), want["raw"].view(np.uint32)), "exact sampler: float4 output differs"
        assert np.array_equal(img, want["rgba8"])
        assert (st.n_ref, st.n_inslice, st.covered_pixels) == (want["stats"].n_ref, want["stats"].n_inslice, want["stats"].covered_pixels)
        if (ext == 1.0).all():
            assert np.array_equal(raw.view(np.uint32), cube["raw"].view(np.uint32)), "unit extents must be the reference's frame"
        for ess in (0, 1):
            op.set_options(sampler=vr.SAMPLER_HW, write_raw=1, collect_stats=0, empty_skip=ess, early_term=0)
            op.record()
            assert_within_tolerance(op.download(), want["rgba8"], f"hw sampler, skipping {ess}")
        op.set_volume_spacing(None)
        op.set_options(sampler=vr.SAMPLER_EXACT, write_raw=1, collect_stats=0, early_term=0)
        op.record()
        assert np.array_equal(op.download_raw().view(np.uint32), cube["raw"].view(np.uint32)), "spacing off: the unit cube again"
    finally:
        op.close()


def test_spacings_of_a_nrrd_are_reported_and_applied_on_request(vr, port, tmp_path):
    """vr_load_nrrd ignores the file's spacings like the reference (unit cube); vr_get_volume_spacing reports them and
    vr_set_volume_spacing applies them -- also with the paired layout and the screen-space tile mask of the skipping path."""
    vol = small_volume(port, 2, np.uint16, 40)
    path = tmp_path / "v.nrrd"
    hdr = "NRRD0004\ntype: ushort\ndimension: 3\nsizes: 40 40 40\nspacings: 1 1 2.5\nendian: little\nencoding: raw\n\n"
    path.write_bytes(hdr.encode() + vol.astype("<u2").tobytes())
    lut = tf(port, "C")
    cam, ocam = twin_cameras(vr, port, [(45, 30)])
    op = vr.OffscreenPass(W, H)
    try:
        op.set_volume_layout(vr.LAYOUT_PAIRED)
        op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=0)
        op.import_nrrd(str(path))
        _, _, mn, mx = op.volume_info()
        fs, e = op.volume_spacing()
        assert list(fs) == [1.0, 1.0, 2.5] and list(e) == [1.0, 1.0, 1.0]
        op.transfer_function_changed(lut)
        op.set_camera(cam)
        f = oracle_frame(port, ocam, W, H, mn, mx)
        op.record()
        assert_within_tolerance(op.download(), port.render(f, vol, lut)["rgba8"], "spacings ignored by default")
        op.set_volume_spacing(fs)
        _, e = op.volume_spacing()
        assert np.allclose(e, [0.4, 0.4, 1.0])
        op.record()
        assert op.last_frame_paired() and op.tile_mask(0) is not None
        assert_within_tolerance(op.download(), port.render_spacing(f, e, vol, lut)["rgba8"], "spacings honoured")
        with pytest.raises(vr.VrError):
            op.set_volume_spacing((1.0, 0.0, 1.0))
    finally:
        op.close()
