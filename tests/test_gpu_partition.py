"""GPU tests of the two partitioned modes, with the N ranks emulated as N contexts on ONE
device (no kernel waits on another, so this is safe on one GPU): sort-first tiles must
reassemble to the single-GPU image bit for bit; sort-last partials merged by K2 in view order
must match the single-GPU image up to fp32 re-association.
"""
import ctypes as C

import numpy as np
import pytest

from common import bits, oracle_frame, p999, tf, twin_cameras

pytestmark = pytest.mark.gpu

W, H = 200, 136  # not a multiple of the tile size on purpose


def setup_pass(vr, w, h, lut, cam, layout=None, **opts):
    op = vr.OffscreenPass(w, h)
    if layout is not None:
        op.set_volume_layout(layout)
    op.set_options(write_raw=1, collect_stats=1, **opts)
    op.transfer_function_changed(lut)
    op.set_camera(cam)
    return op


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_sort_first_tiles_reassemble(vr, port, nranks):
    lut = tf(port, "B")
    cam, _ = twin_cameras(vr, port, [(-120, 80)])
    full = setup_pass(vr, W, H, lut, cam)
    full.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 64)
    full.record()
    want, want_stats = full.download(), full.stats()
    full.close()
    tile = 32
    got = np.zeros_like(want)
    owner = np.full((H, W), -1)
    n_ref = 0
    for r in range(nranks):
        op = setup_pass(vr, W, H, lut, cam)
        op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 64)
        op.set_partition_sort_first(r, nranks, tile)
        op.record()
        img = op.download()
        n_ref += op.stats().n_ref
        op.close()
        ty, tx = np.meshgrid(np.arange(H) // tile, np.arange(W) // tile, indexing="ij")
        mine = vr.sort_first_owner(tx, ty, nranks) == r
        assert (img[~mine] == 0).all(), "a rank wrote outside its tiles"
        got[mine] = img[mine]
        owner[mine] = r
    assert (owner >= 0).all()
    assert np.array_equal(got, want)
    assert n_ref == want_stats.n_ref


# camera moves: one per sign octant of the view direction, two of them close in (other bounding rectangles)
OCTANT_MOVES = [([(-120, 80)], 0.0), ([(60, 80)], 0.0), ([(240, 80)], 1.2), ([(420, 80)], 0.0),
                ([(-120, -140)], 0.0), ([(60, -140)], 1.6), ([(240, -140)], 0.0), ([(420, -140)], 0.0)]


@pytest.mark.parametrize("grid,view", [((2, 1, 1), 0), ((2, 2, 1), 0), ((2, 2, 2), 0), ((3, 2, 1), 0)] + [((2, 2, 2), v) for v in range(1, 8)]
                         + [((3, 2, 1), 5), ((1, 2, 2), 2)])
@pytest.mark.parametrize("sampler", ["exact", "hw", "hw-pairs"])
def test_sort_last_partials_compose_to_single_gpu_image(vr, port, grid, view, sampler):
    import torch  # device memory plumbing only

    n = 48
    lut = tf(port, "B")
    moves, zoom = OCTANT_MOVES[view]
    cam, ocam = twin_cameras(vr, port, moves, zoom)
    smp = vr.SAMPLER_EXACT if sampler == "exact" else vr.SAMPLER_HW
    lay = vr.LAYOUT_PAIRED if sampler == "hw-pairs" else vr.LAYOUT_ARRAY3D   # pairs: every rank's sub-box in the paired layout
    single = setup_pass(vr, W, H, lut, cam, layout=lay, sampler=smp)
    single.synth_volume(vr.SYNTH_FBM, 0x5EED0005, vr.U16, n)
    _, _, mn, mx = single.volume_info()
    single.record()
    want, want_raw, want_stats = single.download(), single.download_raw(), single.stats()

    nbox = grid[0] * grid[1] * grid[2]
    parts, passes, n_inslice, rects, boxes = [], [], 0, [], []
    for b in range(nbox):
        org, dims = vr.partition_box(grid, (n, n, n), b)
        boxes.append((org, dims))
        op = setup_pass(vr, W, H, lut, cam, layout=lay, sampler=smp)
        op.set_partition_sort_last((n, n, n), org, dims)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0005, vr.U16)
        op.set_density_range(mn, mx)
        op.record()
        op.sync()
        assert op.last_frame_paired() == (sampler == "hw-pairs")
        n_inslice += op.stats().n_inslice
        parts.append(op.get_raw())
        passes.append(op)
        # the rank stores nothing outside the screen box of its sub-box: there its partial is the identity
        r = op.partition_screen_rect((n, n, n), org, dims)
        assert r == tuple(op.last_frame_info().bbox)
        rects.append(r)
        raw = op.download_raw()
        outside = np.ones((H, W), bool)
        outside[r[1]:r[3], r[0]:r[2]] = False
        assert not raw[outside].any()
    if sampler == "exact":
        assert n_inslice == want_stats.n_inslice, "every sample must be owned by exactly one rank"
    else:
        # hardware sampler: the ranks step pos_i = entry + i * d (one FMA, the same on every rank: each sample still has
        # exactly one owner), one GPU accumulates -- a ray may gain or lose its last sample at the far face
        assert abs(int(n_inslice) - int(want_stats.n_inslice)) <= 2e-4 * want_stats.n_inslice

    order = vr.partition_order(grid, (n, n, n), cam.get_position())
    single.composite([parts[i] for i in order], 0, W * H)
    got_full = single.download()
    single.composite([parts[i] for i in order], 0, W * H, rects=[rects[i] for i in order])
    got = single.download()
    assert np.array_equal(got, got_full), "reading only the bounding rectangles must change nothing"
    # the same merge over the region that can differ from the clear colour, in bands of 8 rows dealt to three callers
    # (vr_composite_region: what a sort-last rank runs per frame): outside the region the target keeps the clear colour
    slot_region = np.zeros(4, np.uint32)
    ordered = [rects[i] for i in order]
    assert vr.sort_last_region(ordered, W, H, True, slot_region) == (0, 0, W, H)
    region = vr.sort_last_region(ordered, W, H, False, slot_region)
    if zoom == 0.0:
        assert region[2] - region[0] < W and region[3] - region[1] < H     # (close-ups fill the frame)
    single.composite([parts[i] for i in order], 0, W * H, rects=[(0, 0, 0, 0)] * nbox)      # the whole target: clear colour
    for band in range(3):
        single.composite_region([parts[i] for i in order], ordered, region, band, 3)
    assert np.array_equal(single.download(), got), "the region merge in bands must give the same frame"
    for op in passes:
        op.close()
    single.close()
    d = np.abs(got.astype(int) - want.astype(int))
    # paired layout: a rank takes the pair axis of the cube face its rays enter through (api.cu), so the ranks of one frame
    # may filter along different axes than the single GPU did -- still one 8-bit level at most, on a few more pixels
    assert d.max() <= 1 and (d > 0).mean() < (0.03 if sampler == "hw-pairs" else 0.01), (d.max(), (d > 0).mean())
    if sampler == "exact":
        # the wrong order would be visibly wrong; check against the oracle's frame
        o = port.render(oracle_frame(port, ocam, W, H, mn, mx), port.synth(1, 0x5EED0005, np.uint16, n), lut)
        assert p999(got, o["rgba8"]) <= 1


def test_partition_order_is_front_to_back(vr):
    grid, full = (2, 2, 2), (64, 64, 64)
    order = vr.partition_order(grid, full, (0.0, -3.0, 0.0)).tolist()
    pos = {b: k for k, b in enumerate(order)}
    for b in range(8):
        x, y, z = b & 1, (b >> 1) & 1, b >> 2
        if y == 0:  # camera looks along +y from y = -3: the y=0 half is in front of the y=1 half
            assert pos[b] < pos[b | 2]
    assert sorted(order) == list(range(8))


@pytest.mark.parametrize("grid,view", [((2, 2, 2), v) for v in range(8)] + [((2, 1, 1), 0), ((3, 2, 1), 5), ((2, 2, 1), 2)])
@pytest.mark.parametrize("sampler", ["exact", "hw"])
def test_sort_last_with_planes_that_follow_the_camera(vr, port, grid, view, sampler):
    """Sort-last load balance (include/vr.h: vr_partition_balance): every rank HOLDS its regular sub-box grown by a margin
    and OWNS, per frame, the box between split planes computed from the camera.  The planes must stay within the margin,
    the composed frame must equal the single-GPU frame within the sort-last bar, and -- exact sampler -- every sample must
    still be owned by exactly one rank."""
    n = 48
    margin = tuple(5 if g > 1 else 0 for g in grid)
    lut = tf(port, "B")
    moves, zoom = OCTANT_MOVES[view]
    cam, ocam = twin_cameras(vr, port, moves, zoom)
    smp = vr.SAMPLER_EXACT if sampler == "exact" else vr.SAMPLER_HW
    single = setup_pass(vr, W, H, lut, cam, layout=vr.LAYOUT_ARRAY3D, sampler=smp)
    single.synth_volume(vr.SYNTH_FBM, 0x5EED0005, vr.U16, n)
    _, _, mn, mx = single.volume_info()
    single.record()
    want, want_stats = single.download(), single.stats()

    full = (n, n, n)
    nbox = grid[0] * grid[1] * grid[2]
    planes = vr.partition_balance(grid, full, cam.get_position(), margin)
    regular = [k * n // g for a, g in enumerate(grid) for k in range(1, g)]
    assert len(planes) == len(regular) and all(abs(int(p) - r) <= 5 for p, r in zip(planes, regular))
    assert any(int(p) != r for p, r in zip(planes, regular)), "an off-centre camera must move some plane"
    parts, passes, rects, n_inslice, covered = [], [], [], 0, np.zeros((n, n, n), np.int32)
    for b in range(nbox):
        s_org, s_dims = vr.partition_stored_box(grid, full, b, margin)
        r_org, r_dims = vr.partition_box(grid, full, b)
        assert all(so <= ro and so + sd >= ro + rd for so, sd, ro, rd in zip(s_org, s_dims, r_org, r_dims))
        o_org, o_dims = vr.partition_box_planes(grid, full, planes, b)
        covered[o_org[2]:o_org[2] + o_dims[2], o_org[1]:o_org[1] + o_dims[1], o_org[0]:o_org[0] + o_dims[0]] += 1
        op = setup_pass(vr, W, H, lut, cam, layout=vr.LAYOUT_ARRAY3D, sampler=smp)
        op.set_partition_sort_last(full, s_org, s_dims)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0005, vr.U16)
        op.set_density_range(mn, mx)
        op.set_owned_box(o_org, o_dims)
        op.record()
        op.sync()
        n_inslice += op.stats().n_inslice
        parts.append(op.get_raw())
        rects.append(op.partition_screen_rect(full, o_org, o_dims))
        assert rects[-1] == tuple(op.last_frame_info().bbox)
        passes.append(op)
    assert (covered == 1).all(), "the owned boxes must tile the volume"
    with pytest.raises(vr.VrError):
        passes[0].set_owned_box((0, 0, 0), full)      # more than the rank holds
    if sampler == "exact":
        assert n_inslice == want_stats.n_inslice, "every sample must be owned by exactly one rank"
    else:
        assert abs(int(n_inslice) - int(want_stats.n_inslice)) <= 2e-4 * want_stats.n_inslice
    order = vr.partition_order(grid, full, cam.get_position(), planes)
    single.composite([parts[i] for i in order], 0, W * H, rects=[rects[i] for i in order])
    got = single.download()
    for op in passes:
        op.close()
    single.close()
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01, (d.max(), (d > 0).mean())
