"""CPU tests (gloo, world_size 2 and 3) of the host-side logic of the N > 1 paths: pixel ranges,
sub-box grids, and the direct-send exchange schedule of sort-last compositing.  The merge itself
(K2) is CUDA; here a numpy `over` stands in as the checker so the exchange + ordering logic can be
verified without a GPU."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def over(parts):
    """Front-to-back merge of partials [k][n,4] (premultiplied rgb, alpha = 1 - T)."""
    c = np.zeros_like(parts[0][:, :3])
    t = np.ones((parts[0].shape[0], 1), np.float32)
    for p in parts:
        c = c + t * p[:, :3]
        t = t * (1.0 - p[:, 3:4])
    return np.concatenate([c, 1.0 - t], axis=1)


def make_partial(rank, npix):
    rng = np.random.default_rng(100 + rank)
    a = rng.uniform(0, 0.7, (npix, 1)).astype(np.float32)
    return np.concatenate([rng.uniform(0, 1, (npix, 3)).astype(np.float32) * a, a], axis=1)


def _worker(rank, world, port, npix, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from volumetric_renderer_b200 import multigpu as mg
        ranges = mg.pixel_ranges(npix, world)
        order = list(range(world))[::-1] if world == 2 else [1, 2, 0]  # an arbitrary view order
        mine = torch.from_numpy(make_partial(rank, npix))
        recv = mg.exchange_ranges(mine, ranges, rank, world)
        b, e = ranges[rank]
        merged = over([recv[r].numpy() for r in order]) if e > b else np.zeros((0, 4), np.float32)
        per = max(r[1] - r[0] for r in ranges)
        buf = torch.zeros((per, 4))
        buf[: e - b] = torch.from_numpy(merged)
        out = [torch.empty_like(buf) for _ in range(world)] if rank == 0 else None
        dist.gather(buf, out, dst=0)
        if rank == 0:
            full = np.concatenate([out[r][: ranges[r][1] - ranges[r][0]].numpy() for r in range(world)])
            want = over([make_partial(r, npix) for r in order])
            q.put(bool(np.array_equal(full, want)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,npix", [(2, 1000), (3, 77), (2, 31)])
def test_direct_send_exchange_schedule_gloo(world, npix):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, npix, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_pixel_ranges_cover_the_frame():
    from volumetric_renderer_b200 import multigpu as mg
    for npix, world in [(1920 * 1080, 8), (100, 3), (31, 2), (64, 1), (5, 8)]:
        r = mg.pixel_ranges(npix, world)
        assert len(r) == world and r[0][0] == 0 and r[-1][1] == npix
        assert all(a[1] == b[0] for a, b in zip(r, r[1:])) and all(b <= e for b, e in r)
        assert all(b % 32 == 0 for b, _ in r if b < npix)


def test_grid_for_world():
    from volumetric_renderer_b200 import multigpu as mg
    assert sorted(mg.grid_for_world(8)) == [2, 2, 2]
    assert sorted(mg.grid_for_world(4)) == [1, 2, 2]
    assert sorted(mg.grid_for_world(2)) == [1, 1, 2]
    assert mg.grid_for_world(1) == (1, 1, 1)
    for w in (3, 6, 12):
        g = mg.grid_for_world(w)
        assert g[0] * g[1] * g[2] == w


def test_frame_parallel_assignment_is_a_partition():
    """bench.py deals orbit frames round-robin: rank r renders frames r, r+N, ... (no frame twice per sweep)."""
    N, K = 4, 90
    seen = sorted((r + j * N) % 360 for r in range(N) for j in range(K))
    assert seen == list(range(360))


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 5, 6, 7, 8, 12, 16])
def test_sort_first_dealing_is_even_in_every_rectangle(nranks):
    """vr_sort_first_owner deals partition tiles along diagonals, (x + y*m) % n with m coprime to n: the C entry
    point and the Python helper agree, every tile has one owner, and any rectangle of tiles (the cube's screen box)
    that is at least n tiles wide gives every rank the same share to within one tile per row -- also when the
    number of tiles per row is a multiple of n (120 tiles at 8K with 8 ranks used to give vertical strips)."""
    import volumetric_renderer_b200 as vr
    ty, tx = np.meshgrid(np.arange(40), np.arange(120), indexing="ij")
    own = vr.sort_first_owner(tx, ty, nranks)
    assert own.min() == 0 and own.max() == nranks - 1
    for x, y in [(0, 0), (1, 0), (0, 1), (7, 5), (119, 39), (64, 33)]:
        assert vr.lib.vr_sort_first_owner(x, y, nranks) == own[y, x]
    rng = np.random.default_rng(nranks)
    for _ in range(20):
        w, h = int(rng.integers(nranks, 60)), int(rng.integers(1, 30))
        x0, y0 = int(rng.integers(0, 120 - w)), int(rng.integers(0, 40 - h))
        counts = np.bincount(own[y0:y0 + h, x0:x0 + w].ravel(), minlength=nranks)
        assert counts.max() - counts.min() <= h, (w, h, counts)
        if h >= nranks:  # the per-row remainders rotate with the rows: closer than one tile per row
            assert counts.max() - counts.min() <= max(2, h // 2), (w, h, counts)
    # no two horizontally or vertically adjacent tiles share a rank (n >= 2)
    if nranks >= 2:
        assert (own[:, 1:] != own[:, :-1]).all()
    if nranks >= 4:
        assert (own[1:, :] != own[:-1, :]).all()
