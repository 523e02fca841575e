"""Multi-process check of the partitioned modes on real GPUs (one process per GPU, NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29517 tests/mgpu_check.py

Each mode's frame, assembled in rank 0's image, is compared with the single-GPU render of the same
scene on rank 0: sort-first must be bit-identical, sort-last within 1/255 (fp32 re-association).
Run by tests/test_gpu_multi.py when more than one GPU is visible.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_checks(vr, mg, rank, world, local):
    """The checks proper (the process group is up, the device is set).  Returns the list of failures of THIS
    rank; callers reduce over the ranks.  Also run by bench.py for its `partitioned.check` record."""
    import torch.distributed as dist

    W, H, n = 400, 304, 96
    lut = vr.Gradient([(0.0, (0, 0, 0.5)), (0.5, (1, 0.5, 0)), (1.0, (1, 1, 1))], [(0.0, 0.0), (0.3, 0.0), (1.0, 0.08)]).discretize(256)
    cam = vr.Camera().rotate((-120.0, 80.0))
    failures = []

    # single-GPU truth (every rank renders it; rank 0's is the one compared)
    ref = vr.OffscreenPass(W, H, device=local)
    ref.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
    _, _, gmn, gmx = ref.volume_info()
    ref.transfer_function_changed(lut)
    ref.set_camera(cam)
    ref.record()
    want = ref.download()
    ref.close()

    # ---- sort-first: tiles written straight into rank 0's image over NVLink ----------------------
    op = vr.OffscreenPass(W, H, device=local)
    op.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
    op.transfer_function_changed(lut)
    op.set_camera(cam)
    sf = mg.SortFirst(op, rank, world, tile_px=32)
    for _ in range(2):
        sf.render()
        sf.finish()
    if rank == 0:
        got = op.download()
        if not np.array_equal(got, want):
            failures.append(f"sort-first differs in {(got != want).any(axis=-1).sum()} pixels")
    dist.barrier()
    # pipelined: 9 more frames with a moving camera (the cube shrinks and grows: the clear colour is restored by
    # the rank that owns the tile) and no host synchronisation in between; the two image slots then hold the last
    # two frames, each complete and bit-identical to one GPU
    cams = []
    for k in range(9):
        c = vr.Camera().rotate((-120.0 + 40.0 * k, 80.0 - 10.0 * k))
        c.zoom((-2.0, 0.0, 1.5, 0.0, -3.0, 1.0, 0.0, -1.0, 0.5)[k])
        cams.append(c)
        op.set_camera(c)
        sf.render()
    sf.finish()
    if rank == 0:
        for k in (7, 8):
            ref = vr.OffscreenPass(W, H, device=local)
            ref.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
            ref.transfer_function_changed(lut)
            ref.set_camera(cams[k])
            ref.record()
            w_k = ref.download()
            ref.close()
            got = np.empty((H, W, 4), np.uint8)
            op.download_async(got.ctypes.data, None, (2 + k) % 2)
            op.sync()
            if not np.array_equal(got, w_k):
                failures.append(f"sort-first pipelined frame {k} differs in {(got != w_k).any(axis=-1).sum()} pixels")
    dist.barrier()
    sf.close()
    op.close()

    # ---- sort-last: peer-memory direct-send, then the NCCL exchange fallback; two view octants ------------------
    # "balanced": the split planes follow the camera (vr_partition_balance), every rank holds a margin around its box
    for mode in ("p2p", "nccl", "balanced"):
        op = vr.OffscreenPass(W, H, device=local)
        op.transfer_function_changed(lut)
        op.set_camera(cam)
        if mode == "balanced":
            sl = mg.SortLast(op, rank, world, (n, n, n), mode="p2p", balance=0.08)
            if world in (2, 4, 8) and not sl.balanced:
                failures.append("balanced sort-last did not get a margin")
        else:
            sl = mg.SortLast(op, rank, world, (n, n, n), mode=mode)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0004, vr.U16)
        mn, mx = sl.set_global_density_range()
        if (mn, mx) != (gmn, gmx):
            failures.append(f"global density range {mn, mx} != {gmn, gmx}")
        for _ in range(2):
            sl.render(cam.get_position())
            sl.finish()
        for _ in range(5):  # back to back: ordered by the counters only (p2p) / by the collectives (nccl)
            sl.render(cam.get_position())
        sl.finish()
        if rank == 0:
            got = op.download_frame(sl.image_slot())
            d = np.abs(got.astype(int) - want.astype(int))
            if d.max() > 1 or (d > 0).mean() > 0.03:   # the ranks of a frame may sample different pair axes (api.cu: entry-face rule)
                failures.append(f"sort-last/{mode}: max diff {d.max()}, differing {(d > 0).mean():.4f}")
        dist.barrier()
        # a second camera in another octant, closer in (other view order, other bounding rectangles)
        cam2 = vr.Camera().rotate((250.0, -140.0))
        cam2.zoom(1.0)
        op.set_camera(cam2)
        sl.render(cam2.get_position())
        sl.finish()
        if rank == 0:
            ref = vr.OffscreenPass(W, H, device=local)
            ref.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
            ref.transfer_function_changed(lut)
            ref.set_camera(cam2)
            ref.record()
            want2 = ref.download()
            ref.close()
            got = op.download_frame(sl.image_slot())
            d = np.abs(got.astype(int) - want2.astype(int))
            if d.max() > 1 or (d > 0).mean() > 0.03:   # the ranks of a frame may sample different pair axes (api.cu: entry-face rule)
                failures.append(f"sort-last/{mode} second view: max diff {d.max()}, differing {(d > 0).mean():.4f}")
        dist.barrier()
        sl.close()
        op.close()
    return failures


def main():
    import torch
    import torch.distributed as dist

    import volumetric_renderer_b200 as vr
    from volumetric_renderer_b200 import multigpu as mg

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = run_checks(vr, mg, rank, world, local)
    ok = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(ok)
    if rank == 0:
        print("MGPU_CHECK", "OK" if ok.item() == 0 else "FAIL", f"world={world}", failures, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok.item() == 0 else 1)


if __name__ == "__main__":
    main()
