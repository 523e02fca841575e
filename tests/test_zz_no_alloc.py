"""vr_minmax_static (include/vr.h): the entry point that allocates nothing -- its scratch is a static device array.
No counterpart in the reference.  Collected last on purpose.
Checked through the C ABI against the existing path that computes the same thing with a cudaMalloc'd
scratch buffer: vr_set_volume with NULL min / max (k_global_minmax + host reduction), read back with vr_get_volume_info.
min / max of the same values: the results must be EQUAL, not just close."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _existing(vr, vol):
    op = vr.OffscreenPass(16, 16)
    try:
        op.volume_dataset_changed(vr.Dataset(vol))        # min / max None: measured on the device
        _, _, mn, mx = op.volume_info()
        return mn, mx
    finally:
        op.close()


@pytest.mark.parametrize("dtype", ["u8", "u16", "f32"])
def test_static_scratch_minmax_equals_the_allocating_path_twice_in_a_row(vr, dtype):
    import torch
    np_t, code = {"u8": (np.uint8, vr.U8), "u16": (np.uint16, vr.U16), "f32": (np.float32, vr.F32)}[dtype]
    rng = np.random.default_rng(17)
    # two calls in a row on different data: the second is smaller, ragged (not a multiple of a 16-byte load) and has a
    # narrower range, so stale partials of the first call would show
    shapes = [(37, 41, 53), (5, 7, 11)]
    vols = []
    for i, sh in enumerate(shapes):
        if dtype == "f32":
            v = (rng.standard_normal(sh) * (100.0 if i == 0 else 0.01)).astype(np.float32)
        else:
            hi = np.iinfo(np_t).max
            v = rng.integers(0, hi + 1, sh).astype(np_t) if i == 0 else rng.integers(hi // 3, hi // 2, sh).astype(np_t)
        vols.append(v)
    want = [_existing(vr, v) for v in vols]
    torch_t = {"u8": torch.uint8, "u16": torch.int16, "f32": torch.float32}[dtype]
    dev = [torch.from_numpy(v.view(np.int16) if dtype == "u16" else v).to("cuda").contiguous() for v in vols]
    assert all(d.dtype == torch_t for d in dev)
    out = torch.full((2, 2), float("nan"), dtype=torch.float32, device="cuda")
    stream = torch.cuda.Stream()
    # (the first launch of a kernel makes the driver load it, which may take device memory: not what is checked below)
    warm = torch.zeros(2, dtype=torch.float32, device="cuda")
    vr.minmax_static(0, dev[1].data_ptr(), code, 1, warm.data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for i in range(2):   # back to back on one stream, no synchronisation in between: stream order serialises them
        vr.minmax_static(0, dev[i].data_ptr(), code, vols[i].size, out[i].data_ptr(), stream.cuda_stream)
    stream.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the entry point must not allocate device memory"
    got = out.cpu().numpy()
    for i in range(2):
        assert (float(got[i, 0]), float(got[i, 1])) == want[i], f"call {i}: {got[i]} != {want[i]}"
        assert want[i] == (float(vols[i].min()), float(vols[i].max()))


def test_static_scratch_minmax_refuses_what_it_cannot_hold(vr):
    import torch
    lib = vr.lib
    x = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(2, dtype=torch.float32, device="cuda")
    assert lib.vr_minmax_static_max_count() == 1 << 36
    with pytest.raises(vr.VrError, match="exceeds the maximum"):
        vr.minmax_static(0, x.data_ptr(), vr.U8, (1 << 36) + 1, out.data_ptr())
    with pytest.raises(vr.VrError, match="count must be"):
        vr.minmax_static(0, x.data_ptr(), vr.U8, 0, out.data_ptr())
    with pytest.raises(vr.VrError, match="16-byte aligned"):
        vr.minmax_static(0, x.data_ptr() + 1, vr.U8, 8, out.data_ptr())
    with pytest.raises(vr.VrError, match="dtype"):
        vr.minmax_static(0, x.data_ptr(), 7, 8, out.data_ptr())
    with pytest.raises(vr.VrError, match="null"):
        vr.minmax_static(0, 0, vr.U8, 8, out.data_ptr())
    # errbuf may be NULL
    rc = lib.vr_minmax_static(0, C.c_void_p(x.data_ptr()), vr.U8, 0, C.c_void_p(out.data_ptr()), None, None, 0)
    assert rc != 0
    # a single value, default stream
    x[0] = 9
    vr.minmax_static(0, x.data_ptr(), vr.U8, 1, out.data_ptr())
    torch.cuda.synchronize()
    assert out.tolist() == [9.0, 9.0]
