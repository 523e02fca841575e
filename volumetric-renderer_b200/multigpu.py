"""Multi-GPU partitioning of the ray-march pass: one process per GPU, torch.distributed for the
plumbing (rendezvous, barriers, the NCCL fallback exchange), the data path in the CUDA library.

The reference is single-GPU; these are the two partitions BASELINE.json names.

SortFirst   image tiles dealt along diagonals to ranks (vr_sort_first_owner), volume replicated.  Every rank's K1 stores its
            tiles straight into the display rank's RGBA8 image through a CUDA-IPC mapped pointer
            (NVLink peer writes): the gather is the kernel's own epilogue, there is no collective.
SortLast    volume split into a regular grid of sub-boxes (one per rank, one ghost voxel towards
            each neighbour).  Every rank marches the full frame over ITS samples only (global step
            positions, half-open ownership) into a float4 partial; then rank r merges pixel range r
            of all partials in view order with K2, reading the peers' partials directly over NVLink
            and writing into the display rank's image (direct-send; exchange + merge + display
            blend in one kernel).  mode="nccl" exchanges the ranges with send/recv instead.

Host-side logic that does not need a GPU (ranges, grids, ordering, the exchange schedule) is kept in
plain functions so it is testable with the gloo backend.
"""
from __future__ import annotations

import os

import numpy as np


def pixel_ranges(npix: int, world: int):
    """Contiguous pixel range of every rank for direct-send compositing (multiples of 32 pixels)."""
    per = -(-npix // world)
    per = -(-per // 32) * 32
    return [(min(r * per, npix), min((r + 1) * per, npix)) for r in range(world)]


def grid_for_world(world: int):
    """Sub-box grid for `world` ranks: split the axes as evenly as possible (8 -> 2x2x2, 4 -> 2x2x1)."""
    g = [1, 1, 1]
    w, a = world, 0
    f = 2
    factors = []
    while w > 1:
        while w % f == 0:
            factors.append(f)
            w //= f
        f += 1
    for f in sorted(factors, reverse=True):
        a = int(np.argmin(g))
        g[a] *= f
    return tuple(g)


def exchange_ranges(partial, ranges, rank: int, world: int, group=None):
    """Direct-send without peer memory: rank r receives pixel range r of every rank's partial.
    partial: tensor [npix, 4] float32.  Returns tensor [world, len_r, 4] (slot i = rank i's data).
    Uses batched send/recv so it runs on NCCL and on gloo alike."""
    import torch
    import torch.distributed as dist

    b, e = ranges[rank]
    recv = torch.empty((world, e - b, 4), dtype=partial.dtype, device=partial.device)
    recv[rank].copy_(partial[b:e])
    ops = []
    for peer in range(world):
        if peer == rank:
            continue
        pb, pe = ranges[peer]
        if pe > pb:
            ops.append(dist.P2POp(dist.isend, partial[pb:pe].contiguous(), peer, group))
        if e > b:
            ops.append(dist.P2POp(dist.irecv, recv[peer], peer, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return recv


class _DevBuf:
    """A device pointer as a __cuda_array_interface__ object so torch can wrap it without a copy."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2}


def _gather_handles(handle: bytes, world: int, group=None):
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, handle, group=group)
    return out


# flag block layout (uint32 index, VR_FLAG_COUNT = 1024) used by the partitions below; world <= 31, slots <= 4
_F_ARRIVE = 0        # sort-first, display rank: [slot * 32 + r] = frames rank r has stored into image slot `slot`
_F_CONSUMED = 128    # sort-first, every rank: [slot] = frames presented out of image slot `slot` (the slot is free again)
_L_READY = 256       # sort-last, every rank: [slot * 32 + r] = frames whose partial (slot) rank r has finished marching
_L_READ_DONE = 384   # sort-last, every rank: [slot * 32 + r] = frames rank r has finished merging (my partial slot is free again)
_L_RANGE_DONE = 512  # sort-last, display rank: [slot * 32 + r] = frames whose pixel range r is merged into image slot `slot`
_L_CONSUMED = 640    # sort-last, every rank: [slot] = frames presented out of image slot `slot`


class _PeerFlags:
    """Every rank's flag block mapped into this process."""

    def __init__(self, op, rank: int, world: int, group=None):
        self.op, self.rank = op, rank
        handles = _gather_handles(op.ipc_export(2), world, group)
        self.base = [op.flags() if r == rank else op.ipc_open(handles[r]) for r in range(world)]

    def ptr(self, r: int, index: int) -> int:
        return self.base[r] + 4 * index

    def close(self):
        for r, p in enumerate(self.base):
            if r != self.rank:
                self.op.ipc_close(p)
        self.base = []


def _present_priority() -> int:
    """Priority of the display rank's second stream (the one-warp wait / signal kernels of the frame hand-off).
    VR_PRESENT_PRIORITY=1 asks for the highest priority so that they are not queued behind the CTAs of a frame."""
    import os
    return -1 if os.environ.get("VR_PRESENT_PRIORITY", "0") == "1" else 0


class SortFirst:
    """Sort-first image-tile partition (BASELINE.json config 4).

    Frames are handed off in stream order, without a host synchronisation or a collective per frame
    (include/vr.h: vr_stream_signal / vr_stream_wait): every rank stores its tiles into image slot
    k % slots of the display rank and then raises its arrival counter there; the display rank, on a
    second stream, waits for all arrivals of frame k, runs `on_frame(slot, stream)` (e.g. a download)
    and raises the slot's consumed counter on every rank, which is what a rank waits for before it
    overwrites that slot with frame k + slots."""

    def __init__(self, op, rank: int, world: int, tile_px: int = 64, display_rank: int = 0, group=None, slots: int = 2,
                 on_frame=None):
        import torch.distributed as dist
        assert 1 <= slots <= 4 and world <= 31
        self.op, self.rank, self.world, self.display, self.group = op, rank, world, display_rank, group
        self.slots, self.on_frame, self._k = slots, on_frame, 0
        op.set_partition_sort_first(rank, world, tile_px)
        self._peer_imgs, self._flags, self._present = [], None, None
        if world > 1:
            for s in range(slots):
                handles = _gather_handles(op.ipc_export(16 + s) if rank == display_rank else b"", world, group)
                if rank != display_rank:
                    p = op.ipc_open(handles[display_rank])
                    self._peer_imgs.append(p)
                    op.set_image_target(p, s)
            self._flags = _PeerFlags(op, rank, world, group)
            if rank == display_rank:
                import torch
                self._present = torch.cuda.Stream(priority=_present_priority())
            dist.barrier(group)

    def render(self, stream: int | None = None) -> int:
        """Enqueue this rank's tiles of the next frame (and, on the display rank, its presentation).
        Returns the image slot the frame lands in."""
        op, k = self.op, self._k
        s = k % self.slots if self.world > 1 else 0
        self._k += 1
        if self.world == 1:
            op.record(stream)
            if self.on_frame:
                self.on_frame(0, stream)
            return 0
        f = self._flags
        if k >= self.slots:
            op.stream_wait(f.ptr(self.rank, _F_CONSUMED + s), 1, k - self.slots + 1, stream)
        op.record(stream, s)
        op.stream_signal([f.ptr(self.display, _F_ARRIVE + s * 32 + self.rank)], k + 1, stream)
        if self.rank == self.display:
            ps = self._present.cuda_stream
            op.stream_wait(f.ptr(self.rank, _F_ARRIVE + s * 32), self.world, k + 1, ps)
            if self.on_frame:
                self.on_frame(s, ps)
            op.stream_signal([f.ptr(r, _F_CONSUMED + s) for r in range(self.world)], k + 1, ps)
        return s

    def image_slot(self, k: int | None = None) -> int:
        """Image slot of the display rank that holds frame k (default: the frame enqueued last)."""
        k = self._k - 1 if k is None else k
        return k % self.slots if self.world > 1 else 0

    def finish(self):
        """Every enqueued frame is complete in the display rank's image slots when this returns."""
        import torch.distributed as dist
        self.op.sync()
        if self.world > 1:
            dist.barrier(self.group)

    def close(self):
        self.op.sync()
        for s, p in enumerate(self._peer_imgs):
            self.op.set_image_target(None, s)
            self.op.ipc_close(p)
        self._peer_imgs = []
        if self._flags:
            self._flags.close()
            self._flags = None


class SortLast:
    """Sort-last sub-box partition with view-ordered direct-send compositing (BASELINE.json config 5).

    mode "p2p": frames are pipelined over `slots` partial buffers and display-image slots, with the march (K1) and the
    merge (K2) of a rank on two streams: a rank marches frame k + 1 while frame k is still being merged, and may run
    up to `slots` - 1 frames ahead of the slowest rank -- the front-most sub-box of a view has up to 1.4x the samples
    of the rear-most (perspective), and which rank that is changes along an orbit (tools/exp_sl_balance.py)."""

    def __init__(self, op, rank: int, world: int, full_dims, grid=None, display_rank: int = 0, mode: str = "p2p",
                 group=None, on_frame=None, slots: int = 2, balance: float = 0.0):
        """balance: how far (fraction of the volume's edge) the split planes may follow the camera to even out the
        samples per rank (vr_partition_balance).  A rank then holds its regular sub-box grown by that margin on the
        interior faces -- (1/g + 2 b)^3 instead of (1/g)^3 of the volume -- and owns a different part of it per frame."""
        import torch.distributed as dist
        from . import partition_box, partition_stored_box
        self.op, self.rank, self.world, self.display, self.mode, self.group = op, rank, world, display_rank, mode, group
        self.full = tuple(int(x) for x in full_dims)
        self.grid = tuple(grid) if grid else grid_for_world(world)
        assert self.grid[0] * self.grid[1] * self.grid[2] == world
        # the margin must leave each box at least one voxel when both its planes move towards each other
        self.margin = tuple(min(int(balance * n), max((n // g - 1) // 2, 0)) if g > 1 else 0 for n, g in zip(self.full, self.grid))
        self.balanced = any(self.margin)
        self._planes = None
        self._boxes = [partition_box(self.grid, self.full, r) for r in range(world)]
        self.org, self.dims = partition_stored_box(self.grid, self.full, rank, self.margin) if self.balanced else self._boxes[rank]
        op.set_options(write_raw=1)
        op.set_partition_sort_last(self.full, self.org, self.dims)     # what this rank holds
        w, h = op.size
        self.npix = w * h
        self.ranges = pixel_ranges(self.npix, world)
        self.slots = slots if (world > 1 and mode == "p2p") else 1
        assert 1 <= self.slots <= 4 and world <= 31
        self._peers_raw, self._peer_img = None, None
        self._flags, self._present, self._merge, self._k, self.on_frame = None, None, None, 0, on_frame
        if world > 1 and mode == "p2p":
            import torch
            self._peers_raw, self._peer_img = [], []
            for s in range(self.slots):
                raws = _gather_handles(op.ipc_export(32 + s), world, group)
                imgs = _gather_handles(op.ipc_export(16 + s) if rank == display_rank else b"", world, group)
                self._peers_raw.append([op.get_raw(s) if r == rank else op.ipc_open(raws[r]) for r in range(world)])
                self._peer_img.append(None if rank == display_rank else op.ipc_open(imgs[display_rank]))
            self._own_img = [op.get_image_frame(s) for s in range(self.slots)] if rank == display_rank else None
            self._flags = _PeerFlags(op, rank, world, group)
            self._merge = torch.cuda.Stream()
            if rank == display_rank:
                self._present = torch.cuda.Stream(priority=_present_priority())
            # the whole frame as one host call (vr_sort_last_frame): the wiring it needs, kept alive here
            import ctypes as C
            from . import _lib
            w = _lib.SortLastWiring()
            w.rank, w.world, w.display, w.slots = rank, world, display_rank, self.slots
            for i in range(3):
                w.grid[i], w.full[i], w.margin[i] = self.grid[i], self.full[i], self.margin[i]
            self._w_flags = (C.c_void_p * world)(*self._flags.base)
            self._w_raw = (C.c_void_p * (self.slots * world))(*[self._peers_raw[s][r] for s in range(self.slots) for r in range(world)])
            self._w_img = (C.c_void_p * self.slots)(*[(self._own_img[s] if rank == display_rank else self._peer_img[s]) for s in range(self.slots)])
            w.flags = C.cast(self._w_flags, C.POINTER(C.c_void_p))
            w.raw = C.cast(self._w_raw, C.POINTER(C.c_void_p))
            w.image = C.cast(self._w_img, C.POINTER(C.c_void_p))
            w.pixel_begin, w.pixel_end = self.ranges[rank]
            self._wiring = w
            dist.barrier(group)

    def set_global_density_range(self):
        """Reduce the per-rank density range to the global one (ubo.min/max_density are global)."""
        import torch
        import torch.distributed as dist
        _, _, mn, mx = self.op.volume_info()
        if self.world > 1:
            t = torch.tensor([-mn, mx], dtype=torch.float64, device="cuda" if dist.get_backend(self.group) == "nccl" else "cpu")
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            mn, mx = -float(t[0]), float(t[1])
        self.op.set_density_range(mn, mx)
        return mn, mx

    def image_slot(self, k: int | None = None) -> int:
        """Image slot of the display rank that holds frame k (default: the frame enqueued last)."""
        k = self._k - 1 if k is None else k
        return k % self.slots

    def render(self, camera_position, stream: int | None = None):
        """March this rank's samples, then merge.  The frame is complete in the display rank's image
        (slot image_slot()) after finish()."""
        import torch
        import torch.distributed as dist
        from . import partition_balance, partition_box_planes, partition_order
        op, rank, world = self.op, self.rank, self.world
        if world > 1 and self.mode == "p2p" and self.on_frame is None and os.environ.get("VR_SL_HOST", "c") == "c":
            # the frame in one call into the library (the steps below, enqueued by vr_sort_last_frame)
            import ctypes as C
            import numpy as np
            from . import lib
            k = self._k
            self._k += 1
            cam = np.ascontiguousarray(camera_position, np.float32)
            ps = self._present.cuda_stream if rank == self.display else None
            op._check(lib.vr_sort_last_frame(op._h, C.byref(self._wiring), k, cam.ctypes.data_as(C.c_void_p), stream,
                                             self._merge.cuda_stream, ps))
            return
        if self.balanced:
            # the same planes on every rank (a pure function of the camera), then each rank's box of this frame
            self._planes = partition_balance(self.grid, self.full, camera_position, self.margin)
            self._boxes = [partition_box_planes(self.grid, self.full, self._planes, r) for r in range(world)]
            op.set_owned_box(*self._boxes[rank])
        order = [int(b) for b in partition_order(self.grid, self.full, camera_position, self._planes)]
        # every rank's partial is the identity outside the screen box of its sub-box: K1 stores and K2 reads only inside
        rects = [op.partition_screen_rect(self.full, *self._boxes[r]) for r in order]
        b, e = self.ranges[rank]
        if world == 1:
            op.record(stream)
            op.composite([op.get_raw()], 0, self.npix, stream, rects=rects)
            if self.on_frame:
                self.on_frame(0, stream)
            self._k += 1
            return
        if self.mode == "p2p":
            # Stream-ordered hand-off (include/vr.h: vr_stream_signal / vr_stream_wait): no host synchronisation and
            # no collective per frame.  Frame k lives in partial slot / image slot s = k % slots; counters hold k + 1
            # when frame k has passed the stage.
            f, k, disp, S = self._flags, self._k, self.display, self.slots
            s = k % S
            self._k += 1
            ms = self._merge.cuda_stream
            # -- march stream: my partial slot must have been merged by everyone (frame k - S), then K1 --
            if k >= S:
                op.stream_wait(f.ptr(rank, _L_READ_DONE + s * 32), world, k - S + 1, stream)
            op.record(stream, s)
            op.stream_signal([f.ptr(r, _L_READY + s * 32 + rank) for r in range(world)], k + 1, stream)
            # -- merge stream: every partial of frame k complete, image slot s presented (frame k - S), then K2 --
            op.stream_wait(f.ptr(rank, _L_READY + s * 32), world, k + 1, ms)
            if k >= S:
                op.stream_wait(f.ptr(rank, _L_CONSUMED + s), 1, k - S + 1, ms)
            if e > b:
                op.composite([self._peers_raw[s][r] for r in order], b, e, ms, out_ptr=self._peer_img[s] if rank != disp else self._own_img[s],
                             rects=rects)
            op.stream_signal([f.ptr(r, _L_READ_DONE + s * 32 + rank) for r in range(world)] + [f.ptr(disp, _L_RANGE_DONE + s * 32 + rank)],
                             k + 1, ms)
            if rank == disp:
                ps = self._present.cuda_stream
                op.stream_wait(f.ptr(rank, _L_RANGE_DONE + s * 32), world, k + 1, ps)  # the frame is complete in image slot s
                if self.on_frame:
                    self.on_frame(s, ps)
                op.stream_signal([f.ptr(r, _L_CONSUMED + s) for r in range(world)], k + 1, ps)
            return
        self._k += 1
        op.record(stream)
        op.sync()
        dist.barrier(self.group)  # every partial is complete before anyone reads it
        if True:
            raw = torch.as_tensor(_DevBuf(op.get_raw(), (self.npix, 4), "<f4"), device="cuda")
            recv = exchange_ranges(raw, self.ranges, rank, world, self.group)
            torch.cuda.current_stream().synchronize()  # the merge runs on the context's stream
            self._recv = recv  # keep alive until the merge has run
            base = recv.data_ptr()
            n = e - b
            # K2 indexes parts[k] + pixel, so shift each slot back by `b` pixels (16 bytes per pixel)
            ptrs = [base + r * n * 16 - b * 16 for r in order]
            img = torch.as_tensor(_DevBuf(op.get_image()[0], (self.npix, 4), "|u1"), device="cuda")
            if e > b:
                op.composite(ptrs, b, e, None)
            op.sync()
            # gather the merged ranges on the display rank
            per = max(r[1] - r[0] for r in self.ranges)
            mine = torch.zeros((per, 4), dtype=torch.uint8, device="cuda")
            mine[: e - b].copy_(img[b:e])
            out = [torch.empty_like(mine) for _ in range(world)] if rank == self.display else None
            dist.gather(mine, out, dst=self.display, group=self.group)
            if rank == self.display:
                for r, (rb, re_) in enumerate(self.ranges):
                    img[rb:re_].copy_(out[r][: re_ - rb])

    def finish(self):
        import torch.distributed as dist
        self.op.sync()
        if self.world > 1:
            dist.barrier(self.group)

    def close(self):
        if self._peers_raw:
            self.op.sync()
            for slot in self._peers_raw:
                for r, p in enumerate(slot):
                    if r != self.rank:
                        self.op.ipc_close(p)
            for p in self._peer_img:
                if p:
                    self.op.ipc_close(p)
            self._peers_raw, self._peer_img = None, None
        if self._flags:
            self._flags.close()
            self._flags = None
