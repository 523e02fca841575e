"""GPU tests of the peer-memory plumbing.  With one GPU: the redirected image target and
vr_composite_to inside one process.  With two or more GPUs visible: tests/mgpu_check.py under
torchrun (one process per GPU, NCCL + CUDA IPC over NVLink)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from common import tf

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sort_first_ranks_write_into_one_image(vr, port):
    """Rank contexts on one device, all targeting rank 0's image: the assembled frame is the 1-GPU frame."""
    W, H, nranks = 200, 136, 4
    lut = tf(port, "B")
    cam = vr.Camera().rotate((-120.0, 80.0))
    ops = []
    for r in range(nranks):
        op = vr.OffscreenPass(W, H)
        op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 64)
        op.transfer_function_changed(lut)
        op.set_camera(cam)
        ops.append(op)
    ops[0].record()
    want = ops[0].download()
    target, _ = ops[0].get_image()
    for r, op in enumerate(ops):
        op.set_partition_sort_first(r, nranks, 32)
        if r:
            op.set_image_target(target)
    # wipe the display image, then let every rank store its tiles
    ops[0].framebuffer_size_changed(W, H)
    target, _ = ops[0].get_image()
    for r, op in enumerate(ops):
        if r:
            op.set_image_target(target)
        op.record()
        op.sync()
    got = ops[0].download()
    for op in ops:
        op.close()
    assert np.array_equal(got, want)


def test_sort_first_moving_camera_restores_the_clear_colour(vr, port):
    """Sort-first over several frames into ONE display image (rank contexts on one device, run in turn): every rank
    restores the clear colour only on its own tiles and only where the previous frame drew outside this frame's
    tiles; the assembled frame must stay bit-identical to one GPU while the cube moves, shrinks and grows."""
    W, H, nranks = 328, 200, 3
    lut = tf(port, "B")
    ops = []
    for r in range(nranks + 1):          # the last one is the single-GPU truth
        op = vr.OffscreenPass(W, H)
        op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 48)
        op.transfer_function_changed(lut)
        ops.append(op)
    target, _ = ops[0].get_image()
    for r in range(nranks):
        ops[r].set_partition_sort_first(r, nranks, 32)
        if r:
            ops[r].set_image_target(target)
    try:
        stored = []
        for j, z in enumerate([0.0, 0.0, -3.0, -3.0, 0.0, 2.0, 2.3, 0.0, -1.0]):
            cam = vr.Camera().rotate((30.0 * j, 10.0 * j))
            if z:
                cam.zoom(z)
            for op in ops:
                op.set_camera(cam)
            for r in range(nranks):
                ops[r].record()
                ops[r].sync()
            stored.append(sum(ops[r].last_frame_info().target_bytes for r in range(nranks)))
            ops[nranks].record()
            assert np.array_equal(ops[0].download(), ops[nranks].download()), f"frame {j}"
        assert stored[3] < stored[0] // 4        # a small box that did not move: almost nothing is stored
    finally:
        for op in ops:
            op.close()


def _group_devices(n):
    """n ranks: on the visible GPUs when there are enough, else logical ranks on device 0 (events order them)."""
    import torch
    have = torch.cuda.device_count()
    return list(range(n)) if have >= n else [0] * n


@pytest.mark.parametrize("nranks", [2, 3])
@pytest.mark.parametrize("partition", ["sort_first", "sort_last"])
def test_group_renders_the_single_gpu_frames(vr, port, partition, nranks):
    """vr_create_group (include/vr.h, csrc/group.cu; SURVEY 8b last row): ONE process records a frame on N ranks the way
    the application records its one pass (src/rendering/main_pass.cpp:91).  Frames in flight, moving camera; sort-first
    must reproduce the single-GPU frame bit for bit, sort-last within 1/255 on < 3 % of the pixels (the ranks may sample different pair axes)."""
    W, H, n = 328, 200, 48
    lut = tf(port, "B")
    single = vr.OffscreenPass(W, H)
    single.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, n)
    single.transfer_function_changed(lut)
    vol = single.download_volume()
    g = vr.PassGroup(_group_devices(nranks), vr.PARTITION_SORT_FIRST if partition == "sort_first" else vr.PARTITION_SORT_LAST, W, H, slots=2)
    try:
        assert len(g) == nranks
        g.transfer_function_changed(lut)
        if partition == "sort_first":
            g.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, n)
        else:
            g.volume_dataset_changed(vr.Dataset(vol))      # cut into sub-boxes + ghosts on the host, global range reduced
        cams = []
        for j, z in enumerate([0.0, 0.0, -3.0, 1.5, 0.0, 2.0, -1.0]):
            cam = vr.Camera().rotate((50.0 * j - 120.0, 80.0 - 35.0 * j))
            if z:
                cam.zoom(z)
            cams.append(cam)
        for j, cam in enumerate(cams):                      # enqueue everything, no host synchronisation in between
            g.set_camera(cam)
            g.record()
            if j in (2, 4, 6):
                got = g.download()
                single.set_camera(cam)
                single.record()
                want = single.download()
                d = np.abs(got.astype(int) - want.astype(int))
                if partition == "sort_first":
                    assert d.max() == 0, f"frame {j}: {np.count_nonzero(d.max(axis=-1))} pixels differ"
                else:
                    assert d.max() <= 1 and (d > 0).mean() < 0.03, (j, d.max(), (d > 0).mean())
    finally:
        g.close()
        single.close()


def test_group_replicas_render_their_own_cameras(vr, port):
    W, H = 160, 120
    g = vr.PassGroup(_group_devices(2), vr.PARTITION_REPLICAS, W, H)
    single = vr.OffscreenPass(W, H)
    try:
        lut = tf(port, "B")
        for x in (g, single):
            x.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, 32)
            x.transfer_function_changed(lut)
        cams = [vr.Camera().rotate((40.0, 10.0)), vr.Camera().rotate((-200.0, 60.0))]
        for r, cam in enumerate(cams):
            g.set_rank_camera(r, cam)
        g.record()
        for r, cam in enumerate(cams):
            single.set_camera(cam)
            single.record()
            assert np.array_equal(g.download(r), single.download())
    finally:
        g.close()
        single.close()


def test_stream_flags_signal_wait_and_time_out(vr):
    """vr_stream_signal stores into local or foreign counters in stream order; vr_stream_wait holds a stream
    until its counters reach the value; a wait nobody answers gives up and makes the next vr_sync fail
    instead of hanging the GPU.  (Here the counter is raised from the host: two compute streams of ONE
    process may share a hardware queue, so a signal kernel could sit behind the held stream; between
    processes -- tests/mgpu_check.py -- the signals come from other GPUs.)"""
    import torch
    from volumetric_renderer_b200.multigpu import _DevBuf
    a, b = vr.OffscreenPass(64, 64), vr.OffscreenPass(64, 64)
    try:
        fa = torch.as_tensor(_DevBuf(a.flags(), (256,), "<u4"), device="cuda")
        fb = torch.as_tensor(_DevBuf(b.flags(), (256,), "<u4"), device="cuda")
        assert int(fa.sum()) == 0 and int(fb.sum()) == 0
        b.stream_signal([a.flags() + 4 * 7, b.flags() + 4 * 200], 41)   # one foreign, one own counter
        b.sync()
        assert int(fa[7]) == 41 and int(fb[200]) == 41 and int(fa.sum()) == 41
        sa = torch.cuda.Stream()
        a.stream_wait(a.flags() + 4 * 5, 2, 3, sa.cuda_stream)          # counters 5 and 6 of `a` must reach 3
        done = torch.cuda.Event()
        done.record(sa)
        fa[5:7] = torch.tensor([3, 2], dtype=fa.dtype, device="cuda")
        torch.cuda.current_stream().synchronize()
        assert not done.query()
        fa[6] = 7                                                       # counters only grow: 7 >= 3
        sa.synchronize()
        assert done.query()
        a.sync()
        b.stream_wait(b.flags() + 4 * 9, 2, 1, None, timeout_ms=50)
        with pytest.raises(vr.VrError, match="timed out"):
            b.sync()
        b.sync()                                                        # the error is reported once
    finally:
        a.close()
        b.close()


def test_multi_process_partitions_on_real_gpus():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "mgpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MGPU_CHECK OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_cpp_host_mirror_renders_the_same_frames(vr, tmp_path):
    """host/example_orbit (C++ over host/offscreen_pass.hpp) against the Python mirror: same frames."""
    exe = os.path.join(ROOT, "volumetric-renderer_b200", "host", "example_orbit")
    if not os.path.exists(exe):
        pytest.skip("example_orbit not built")
    W, H, n, frames = 160, 120, 48, 3
    out = tmp_path / "o.ppm"
    r = subprocess.run([exe, str(out), str(W), str(H), str(n), str(frames)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    got_sum = int(r.stdout.split("checksum=")[1].split()[0])
    op = vr.OffscreenPass(W, H)
    op.synth_volume(vr.SYNTH_FBM, 0x5EED0002, vr.U16, n)
    op.transfer_function_changed(vr.Gradient([(0.0, (0, 0, 0.5)), (0.5, (1, 0.5, 0)), (1.0, (1, 1, 1))],
                                             [(0.0, 0.0), (0.3, 0.0), (1.0, 0.08)]).discretize(256))
    cam, total, img = vr.Camera(), 0, None
    for _ in range(frames):
        op.set_camera(cam)
        op.record()
        img = op.download()
        total += int(img.astype(np.int64).sum())
        cam.rotate((4.0, 0.0))
    op.close()
    assert got_sum == total
    ppm = open(out, "rb").read()
    assert ppm.startswith(b"P6\n160 120\n255\n")
    body = np.frombuffer(ppm[len(b"P6\n160 120\n255\n"):], np.uint8).reshape(H, W, 3)
    assert np.array_equal(body, img[..., :3])


@pytest.mark.parametrize("mode", ["sort_first", "sort_last"])
def test_cpp_group_mirror_matches_one_gpu(vr, mode):
    """host/example_group (C++ VolB200::Rendering::PassGroup over vr_create_group): the group's frames against a single
    pass in the same program -- identical for sort-first, within one 8-bit level for sort-last."""
    exe = os.path.join(ROOT, "volumetric-renderer_b200", "host", "example_group")
    if not os.path.exists(exe):
        pytest.skip("example_group not built")
    devs = _group_devices(3)
    r = subprocess.run([exe, mode, "200", "144", "48", "3"] + [str(d) for d in devs], capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, r.stderr
    f = dict(kv.split("=") for kv in r.stdout.split())
    assert int(f["ranks"]) == 3 and int(f["frames"]) == 3
    if mode == "sort_first":
        assert f["checksum_group"] == f["checksum_single"] and int(f["max_abs_diff"]) == 0
    else:
        assert int(f["max_abs_diff"]) <= 1
