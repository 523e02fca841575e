#!/usr/bin/env python
"""tools/exp_handoff_trace.py -- where the time of a sort-first frame goes on N GPUs (round-2 tool, written when
the 8-GPU budget of round 1 was spent: C4 takes 0.374 ms/frame on 8 GPUs although the slowest rank's kernel alone
takes 0.294 ms).

Run like bench.py:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
                        tools/exp_handoff_trace.py [--size 2048] [--frame 7680x4320] [--frames 36] [--slots 2] [--tile-px 64]

Every rank brackets the three things it enqueues per frame with CUDA events on its stream
    wait(consumed) | K1 (tiles stored into the display rank's image) | signal(arrive)
and the display rank the two on its present stream
    wait(all arrived) | signal(consumed to every rank).
Printed per rank: mean time spent in each, the mean gap between the end of one frame's signal and the start of
the next frame's kernel, and the frame period.  A rank whose `wait consumed` is long is waiting for the display rank;
a display rank whose `wait arrived` ends long after the slowest K1 ended is waiting on the hand-off itself."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    import volumetric_renderer_b200 as vr
    from bench import product_lut
    from volumetric_renderer_b200 import multigpu as mg

    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--frame", default="7680x4320")
    ap.add_argument("--frames", type=int, default=36)
    ap.add_argument("--slots", type=int, default=2)
    ap.add_argument("--tile-px", type=int, default=64)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H = (int(x) for x in args.frame.lower().split("x"))
    op = vr.OffscreenPass(W, H, device=local)
    op.transfer_function_changed(product_lut(vr, "B"))
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0, collect_stats=0)
    op.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, args.size)
    part = mg.SortFirst(op, rank, world, tile_px=args.tile_px, slots=args.slots)
    cam, proj, views = vr.Camera(), vr.projection(W, H), []
    for _ in range(36):
        views.append((cam.get_view(), proj, cam.get_position()))
        cam.rotate((40.0, 0.0))
    stream = torch.cuda.Stream()
    sp = stream.cuda_stream

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def traced_frame(k, rec):
        """SortFirst.render, step by step, with an event after every step."""
        op.set_uniforms(*views[k % 36])
        if world == 1:
            e0, e1 = ev(), ev()
            e0.record(stream)
            op.record(sp)
            e1.record(stream)
            rec.append(dict(k0=e0, k1=e1))
            return
        f, s = part._flags, part._k % part.slots
        kk = part._k
        part._k += 1
        e = dict(w0=ev(), k0=ev(), k1=ev(), s1=ev())
        e["w0"].record(stream)
        if kk >= part.slots:
            op.stream_wait(f.ptr(rank, mg._F_CONSUMED + s), 1, kk - part.slots + 1, sp)
        e["k0"].record(stream)
        op.record(sp, s)
        e["k1"].record(stream)
        op.stream_signal([f.ptr(part.display, mg._F_ARRIVE + s * 32 + rank)], kk + 1, sp)
        e["s1"].record(stream)
        if rank == part.display:
            ps = part._present
            e.update(p0=ev(), p1=ev(), p2=ev())
            e["p0"].record(ps)
            op.stream_wait(f.ptr(rank, mg._F_ARRIVE + s * 32), world, kk + 1, ps.cuda_stream)
            e["p1"].record(ps)
            op.stream_signal([f.ptr(r, mg._F_CONSUMED + s) for r in range(world)], kk + 1, ps.cuda_stream)
            e["p2"].record(ps)
        rec.append(e)

    warm = []
    for k in range(4):
        traced_frame(k, warm)
    part.finish()
    rec = []
    origin = ev()
    origin.record(stream)
    for k in range(args.frames):
        traced_frame(k, rec)
    part.finish()

    def mean(xs):
        xs = list(xs)
        return sum(xs) / max(len(xs), 1)

    n = len(rec)
    out = {"rank": rank, "kernel_ms": mean(r["k0"].elapsed_time(r["k1"]) for r in rec),
           "period_ms": origin.elapsed_time(rec[-1]["k1"]) / n}
    if world > 1:
        out["wait_consumed_ms"] = mean(r["w0"].elapsed_time(r["k0"]) for r in rec)
        out["signal_ms"] = mean(r["k1"].elapsed_time(r["s1"]) for r in rec)
        out["gap_signal_to_next_kernel_ms"] = mean(rec[i]["s1"].elapsed_time(rec[i + 1]["k0"]) for i in range(n - 1))
        if rank == part.display:
            out["present_wait_arrived_ms"] = mean(r["p0"].elapsed_time(r["p1"]) for r in rec)
            out["present_signal_ms"] = mean(r["p1"].elapsed_time(r["p2"]) for r in rec)
            # how long after its own kernel ended did the display rank see the whole frame?
            out["own_kernel_end_to_all_arrived_ms"] = mean(r["k1"].elapsed_time(r["p1"]) for r in rec)
    rows = [None] * world
    if world > 1:
        dist.all_gather_object(rows, out)
    else:
        rows = [out]
    if rank == 0:
        for r in rows:
            print("  ".join(f"{k}={v:.4f}" if isinstance(v, float) else f"{k}={v}" for k, v in r.items()))
    op.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
