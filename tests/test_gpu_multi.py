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
