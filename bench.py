#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

Metric: reference-equivalent ray samples/s (= iterations of res/shaders/volume.frag:33-48 that pass
the bounds test, what the reference shader would execute for the same frames) on config C2:
512^3 uint16 fractal-noise volume, 1920x1080, 360-frame orbit (frame k = default camera after k x
Camera::rotate({4,0})), transfer function TF-B.

A step = one frame = one launch of K1 (march).  The K steps of a rank are spread over the WHOLE orbit
(rank r, step j renders orbit frame ((r + j N) 360) // (K N)), so every N renders the same mix of views;
the K steps are repeated `reps` times inside the event-timed region until it lasts >= 0.3 s.
N GPUs: the orbit's frames are independent units (weak scaling, no data-path collective).  With N > 1 the same
run then measures the partitioned configs (one frame rendered by ALL ranks): sort-first C4 and sort-last on a
volume that also fits one GPU, each with its own single-GPU time -> the `partitioned` record.

  python bench.py [--gpus N] [--steps K] [--warmup W]            this backend
  python bench.py --impl reference [...]                         the reference's own CPU code
                                                                 (oracle/_ref: its shader compiled as C++)
Other workloads (not the driver's line): --workload c1|c3|c4|c5, see DESIGN.md.
"""
import argparse
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
T_START = time.time()

WORKLOADS = {
    # name: (synth kind, dtype name, n, W, H, TF, description)
    "c1": (0, "u8", 256, 512, 512, "A", "256^3 u8 Marschner-Lobb, 512x512, TF-A (reference default gradient)"),
    "c2": (1, "u16", 512, 1920, 1080, "B", "512^3 u16 fBm, 1920x1080, 360-frame orbit, TF-B"),
    "c3": (2, "u16", 1024, 3840, 2160, "C", "1024^3 u16 sparse blobs, 3840x2160, 360-frame orbit, TF-C"),
}
SEEDS = {"c1": 0x5EED0001, "c2": 0x5EED0002, "c3": 0x5EED0003}
ORBIT_FRAMES = 360
MIN_TIMED_S = 0.3


# transfer functions of SURVEY.md section 8(d): (colour markers (loc, r, g, b), alpha markers (loc, a))
TF_MARKERS = {
    "A": ([(0.0, 0, 0, 0), (1.0, 1, 1, 1)], [(0.0, 1.0), (1.0, 1.0)]),  # gradient.cpp:64-70 default
    "B": ([(0.0, 0, 0, 0.5), (0.5, 1, 0.5, 0), (1.0, 1, 1, 1)], [(0.0, 0.0), (0.3, 0.0), (1.0, 0.08)]),
    "C": ([(0.0, 0, 0, 0.5), (0.5, 1, 0.5, 0), (1.0, 1, 1, 1)], [(0.0, 0.0), (0.05, 0.0), (1.0, 0.6)]),
}


def tf_markers(name):
    return TF_MARKERS[name]


def product_lut(vr, name):
    cm, am = tf_markers(name)
    return vr.Gradient([(m[0], tuple(m[1:])) for m in cm], [tuple(m) for m in am]).discretize(256)


def orbit_uniforms(vr, w, h, frames):
    """(view, proj, camera_position) of orbit frame k, k = 0..frames-1."""
    cam, proj, out = vr.Camera(), vr.projection(w, h), []
    for _ in range(frames):
        out.append((cam.get_view(), proj, cam.get_position()))
        cam.rotate((4.0, 0.0))
    return out


def orbit_frames(count, rank, world):
    """The orbit frames of this rank's `count` steps: all ranks together cover the orbit evenly, and so does
    every rank alone -- N = 1 and N = 8 render the same distribution of views."""
    return [((rank + j * world) * ORBIT_FRAMES) // (count * world) % ORBIT_FRAMES for j in range(count)]


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU, sampled with NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz, self.ok = index, [], set(), False, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def sample(self):
        if not self.ok:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("hw_thermal_slowdown", 0x40),
                              ("sw_thermal_slowdown", 0x20), ("hw_power_brake_slowdown", 0x80)):
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self.stop_flag:
            self.sample()
            time.sleep(0.002)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def measured_traffic(workload, kernel_version):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_march launch of this workload, from the committed
    ncu --set full capture (profiles/traffic.json names the summary and the kernel version it was read from).
    Returns (bytes or None, note)."""
    try:
        e = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[workload]
    except Exception:
        return None, "no ncu capture committed for this workload"
    note = f"ncu capture {e.get('source')} of kernel version {e.get('kernel_version')}"
    if e.get("kernel_version") != kernel_version:
        note += f" -- STALE: this run is kernel version {kernel_version}"
    return e["dram_bytes_per_launch"], note


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# ------------------------------------------------------------------------------------------------
# The driver's line: orbit frames of config 2 dealt to the ranks
# ------------------------------------------------------------------------------------------------

def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import volumetric_renderer_b200 as vr  # raises if libvr_b200.so is missing: no fallback

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this backend has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    kind, dname, n, W, H, tfn, desc = WORKLOADS[args.workload]
    if os.environ.get("VR_BENCH_N"):  # tuning experiments only (changes the workload: not a bench line)
        n = int(os.environ["VR_BENCH_N"])
        desc += f" [EXPERIMENT: volume {n}^3]"
    dtype = {"u8": vr.U8, "u16": vr.U16, "f32": vr.F32}[dname]
    esz = {"u8": 1, "u16": 2, "f32": 4}[dname]
    K, Wu = args.steps, args.warmup

    op = vr.OffscreenPass(W, H, device=local)
    op.set_volume_layout({"auto": vr.LAYOUT_AUTO, "3d": vr.LAYOUT_ARRAY3D, "paired": vr.LAYOUT_PAIRED}[args.layout])
    op.synth_volume(kind, SEEDS[args.workload], dtype, n)
    lut = product_lut(vr, tfn)
    op.transfer_function_changed(lut)
    prod = dict(sampler=vr.SAMPLER_EXACT if args.exact else vr.SAMPLER_HW, empty_skip=int(not args.no_ess),
                early_term=int(not args.no_ert), early_term_eps=1.0 / 1024.0)
    uniforms = orbit_uniforms(vr, W, H, ORBIT_FRAMES)
    my_frames = orbit_frames(K, rank, world)
    warm_frames = orbit_frames(Wu, rank, world)
    F = max(1, min(args.in_flight, 4))      # frames in flight (the reference keeps MAX_FRAMES_IN_FLIGHT = 2)
    streams = [torch.cuda.Stream() for _ in range(F)]
    sp = streams[0].cuda_stream

    # ---- untimed: work per frame (reference-equivalent samples, executed fetches), what each frame launched --------
    n_ref, n_exec, info = {}, {}, {}
    for k in sorted(set(my_frames)):
        op.set_uniforms(*uniforms[k])
        op.set_options(collect_stats=1, write_raw=0, sampler=prod["sampler"], empty_skip=0, early_term=0,
                       early_term_eps=prod["early_term_eps"])
        op.record(sp)
        n_ref[k] = op.stats().n_ref
        op.set_options(collect_stats=1, **prod)
        op.record(sp)
        n_exec[k] = op.stats().n_fetch
        fi = op.last_frame_info()
        info[k] = (fi.layout, fi.pair_axis, fi.empty_skip_active)
    op.set_options(collect_stats=0, **prod)
    if args.preintegration:
        # optional quality mode (changes the image; not the headline): segments classified from a 2-D table
        op.set_preintegration(args.preintegration)
    cells, cell_vox, empty = op.occupancy_info()
    total_ref = sum(n_ref[k] for k in my_frames)
    total_exec = sum(n_exec[k] for k in my_frames)
    ess_active = any(v[2] for v in info.values())
    paired_frames = sum(1 for k in my_frames if info[k][0] == vr.LAYOUT_PAIRED)
    axis_hist = {}
    for k in my_frames:
        a = "3-D array" if info[k][0] != vr.LAYOUT_PAIRED else "pairs along " + "xyz"[info[k][1]]
        axis_hist[a] = axis_hist.get(a, 0) + 1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def agree_max(x):
        if world == 1:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def fork(ss):
        ev0.record(ss[0])
        for s in ss[1:]:
            s.wait_event(ev0)

    def join(ss):
        for s in ss[1:]:
            e = torch.cuda.Event()
            e.record(s)
            ss[0].wait_event(e)
        ev1.record(ss[0])

    def timed_frames(frames, reps, ss):
        """reps x len(frames) launches of K1 on the streams ss (one frame slot per stream), device-timed."""
        barrier()
        fork(ss)
        j = 0
        for _ in range(reps):
            for k in frames:
                f = j % len(ss)
                op.set_uniforms(*uniforms[k])
                op.record(ss[f].cuda_stream, f)
                j += 1
        join(ss)
        barrier()
        return ev0.elapsed_time(ev1)

    # ---- warm-up, then the number of repetitions that makes the timed region >= MIN_TIMED_S -------------------------
    timed_frames(warm_frames, 1, streams)
    probe_ms = agree_max(timed_frames(my_frames, 1, streams))
    reps = int(min(2000, max(1, math.ceil(MIN_TIMED_S * 1e3 / max(probe_ms, 1e-3)))))

    # ---- device-timed: inputs resident, reps x K launches of K1, F frames in flight ----------------------------------
    clocks = ClockSampler(physical_gpu_index(local))
    clocks.start()
    dev_ms = timed_frames(my_frames, reps, streams)
    clocks.sample()
    clocks.stop_flag = True
    # the same steps one frame at a time (no overlap between consecutive frames), and with early termination off
    f1_ms = timed_frames(my_frames, max(1, reps // 2), streams[:1])
    noert_ms = None
    if prod["early_term"]:
        op.set_options(collect_stats=0, **dict(prod, early_term=0))
        noert_ms = timed_frames(my_frames, max(1, reps // 2), streams)
        op.set_options(collect_stats=0, **prod)

    # ---- end to end through the public API: host uniforms in, RGBA8 frame out to pinned host memory ---
    host = [torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True) for _ in range(F)]
    # Each step leaves the complete frame in pinned host memory.  --full-download: the whole RGBA8 image is copied
    # every step; default: vr_download_frame_dirty_async copies the rectangle that can differ from what that host
    # buffer already holds (outside the cube's screen box a frame is the clear colour) -- d2h bytes are counted.
    e2e_reps = max(1, reps // 2)

    def run_e2e(fetch):
        """e2e_reps x K steps; every step: uniforms from host memory, render, the frame brought up to date in pinned host memory."""
        for j, k in enumerate(warm_frames[:F] if len(warm_frames) >= F else list(warm_frames) * F):
            op.set_uniforms(*uniforms[k])
            op.record(streams[j % F].cuda_stream, j % F)
            fetch(host[j % F].data_ptr(), streams[j % F].cuda_stream, j % F)   # the first delivery into a buffer is always full
        barrier()
        nbytes, j = 0, 0
        t0 = time.perf_counter()
        for _ in range(e2e_reps):
            for k in my_frames:
                f = j % F
                streams[f].synchronize()           # frame j-F is in host memory (its buffer may be reused now)
                op.set_uniforms(*uniforms[k])      # 192-byte uniform block: the step's input, goes down with the launch
                op.record(streams[f].cuda_stream, f)
                nbytes += fetch(host[f].data_ptr(), streams[f].cuda_stream, f) or W * H * 4
                j += 1
        barrier()                              # every frame of the steps has landed in host memory
        ms = (time.perf_counter() - t0) * 1e3
        # the host frames must be what a full download gives
        for f in range(min(F, K)):
            full = torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True)
            op.download_async(full.data_ptr(), streams[f].cuda_stream, f)
            streams[f].synchronize()
            if not torch.equal(full, host[f]):
                raise SystemExit(f"bench.py: host frame of slot {f} differs from a full download")
        return ms, nbytes

    # the same steps with the whole frame copied every step (PCIe-bound), reported beside the headline
    e2e_full_ms, _ = run_e2e(op.download_async)
    if args.full_download:
        e2e_ms, d2h_bytes = e2e_full_ms, e2e_reps * K * W * H * 4
    else:
        e2e_ms, d2h_bytes = run_e2e(op.download_dirty_async)
    checksum = int(host[(e2e_reps * K - 1) % F].view(torch.int32).sum().item())

    # ---- what the memory system alone allows: the same frames through the fetch-only build of the kernel ---
    # (vr_bench_fetch_only: ray setup, stepping and every volume fetch of the frame, no transfer function and
    # no compositing; a measurement aid, reported beside the rooflines, not part of `value`)
    fetch_only_ms = None
    if world == 1 and not args.exact:
        try:
            op.bench_fetch_only(True)
            timed_frames(warm_frames[:3], 1, streams)
            fetch_only_ms = timed_frames(my_frames, max(1, reps // 2), streams) / (max(1, reps // 2) * K)
        except vr.VrError:
            fetch_only_ms = None     # empty-space skipping is active for this data: no fetch-only build
        finally:
            op.bench_fetch_only(False)

    t = torch.tensor([dev_ms, e2e_ms, e2e_full_ms, f1_ms, float(total_ref), float(total_exec)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, e2e_ms, e2e_full_ms, f1_ms = mx[0].item(), mx[1].item(), mx[2].item(), mx[3].item()
        all_ref, all_exec = sm[4].item(), sm[5].item()
    else:
        all_ref, all_exec = float(total_ref), float(total_exec)

    # screen-space tile mask (skipping active, no statistics): one more launch of ours per frame, ahead of K1
    mask = op.tile_mask(0)
    mask_info = None
    if mask is not None:
        bits, ignored, th = mask
        x0, y0, x1, y1 = op.last_frame_info().bbox
        inside = bits[y0 // th:(y1 + th - 1) // th, x0 // 16:(x1 + 15) // 16]
        mask_info = {"tiles_in_screen_box": int(inside.size), "tiles_marched": int(inside.sum()), "ignored": bool(ignored),
                     "what": "k_tile_mask projects the occupied cells ahead of K1; unmarked tiles get the clear colour without ray set-up (last frame of the run)"}
    launches_per_step = 2 if mask is not None else 1
    line = None
    if rank == 0:
        peaks = load_peaks()
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        ncell = cells[0] * cells[1] * cells[2]
        occ_frac = (ncell - empty) / ncell if ess_active else 1.0      # cells are skipped only when the context consults the grid
        alg_bytes = n ** 3 * esz * occ_frac + W * H * 4 + 1024   # touched bricks once + RGBA8 out + LUT
        step_s = dev_ms * 1e-3 / (K * reps)                       # one launch per step
        paired = paired_frames * 2 >= K
        tex_peak = vr.bench_tex_peak(local, dtype, 2 if paired else 1)
        tex3d_peak = vr.bench_tex_peak(local, dtype, 1) if paired else tex_peak   # one 3-D trilinear fetch per sample
        tex_dram = vr.bench_tex_peak(local, dtype, 3) if dtype != vr.F32 else None
        tex_dram_bytes = vr.bench_tex_peak(local, dtype, 4) if dtype != vr.F32 else None
        exec_rate = (total_exec / K) / step_s
        traffic, traffic_note = measured_traffic(args.workload, vr.KERNEL_VERSION)
        line = {
            "metric": "ray samples/s @1920x1080, 512^3 u16" if args.workload == "c2" else f"ray samples/s ({args.workload})",
            "value": all_ref * reps / (dev_ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wu,
            "reps": reps, "timed_steps": K * reps, "timed_region_s": dev_ms * 1e-3,
            "ms_per_step": dev_ms / (K * reps), "ms_per_step_in_flight_1": f1_ms / (K * max(1, reps // 2)),
            "ms_per_step_early_termination_off": (noert_ms / (K * max(1, reps // 2))) if noert_ms else None,
            "fps": world * K * reps / (dev_ms * 1e-3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "volume_dtype": dname, "kernel_version": vr.KERNEL_VERSION,
                       "sampler": "exact(8 point fetches + fp32 lerp)" if args.exact else "hw trilinear",
                       "preintegration_substeps": args.preintegration,
                       "empty_space_skipping": {"requested": bool(prod["empty_skip"]), "active": bool(ess_active),
                                                "why": "the context consults the occupancy grid only when >= 25 % of the cells are empty",
                                                "screen_space_mask": mask_info},
                       "early_ray_termination_eps": prod["early_term_eps"] if prod["early_term"] else 0,
                       "l2": (f"inputs larger than L2 ({n ** 3 * esz / 2 ** 20:.0f} MiB volume, no flush)" if n ** 3 * esz > 126e6 else
                              f"the {n ** 3 * esz / 2 ** 20:.0f} MiB volume FITS the 126 MB L2 and stays resident across frames (not flushed): "
                              "a parity-size configuration, not a headline number"),
                       "frames_in_flight": F,
                       "frames": f"{K} steps per rank spread evenly over the 360-frame orbit (same mix of views for every N), repeated {reps}x inside the timed region",
                       "layout_per_frame": axis_hist,
                       "layout": "paired slices, 2-D layered array, pair axis chosen per frame (+ the 3-D array kept for the exact sampler)" if paired else "3-D array",
                       "parallelism": f"frames of the orbit dealt to {world} GPU(s), no collective"},
            "executed_samples_per_s": all_exec * reps / (dev_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": alg_bytes / step_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": alg_bytes / step_s / 1e9 / hbm_peak, "traffic": traffic, "traffic_source": traffic_note,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback",
                         "algorithmic_bytes_per_launch": alg_bytes},
            "roofline_tex": {"bound": "tex", "achieved": exec_rate / 1e9, "peak": tex_peak / 1e9, "unit": "Gfetch/s",
                             "frac": exec_rate / tex_peak,
                             "peak_source": "vr_bench_tex_peak: L1-resident " + ("paired-layout bilinear fetches (2-D layered, two slices per texel)" if paired else "3-D trilinear fetches") + ", measured in this run",
                             "frac_of_3d_trilinear_peak": exec_rate / tex3d_peak,
                             "fetch_only_ms_per_step": fetch_only_ms,
                             "fetch_only_share_of_step": (fetch_only_ms / (dev_ms / (K * reps))) if fetch_only_ms else None},
            "roofline_tex_dram": None if not tex_dram else {
                "bound": "tex path fed from DRAM", "achieved": exec_rate / 1e9, "peak": tex_dram / 1e9, "unit": "Gfetch/s",
                "frac": exec_rate / tex_dram,
                "peak_source": "vr_bench_tex_peak mode 3, measured in this run: a 1 GiB paired layered array (8x the L2) streamed once "
                               "through bilinear fetches with ideal quad locality, one fetch per texel "
                               f"({2 * esz} new bytes per fetch; K1 here: {alg_bytes / max(total_exec / K, 1):.1f})",
                "byte_ceiling_gbs": tex_dram_bytes * 8 * esz / 1e9,
                "byte_ceiling_source": "mode 4: the same array, one fetch per 2x2 texels"},
            "e2e": {"value": all_ref * e2e_reps / (e2e_ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": 192,
                    "d2h_bytes_per_step": d2h_bytes / (K * e2e_reps), "frame_bytes": W * H * 4,
                    "d2h": "full frame every step" if args.full_download else "dirty rectangle (vr_download_frame_dirty_async); the host frame is complete and checked against a full download",
                    "ms_per_step": e2e_ms / (K * e2e_reps), "ms_per_step_full_frame_every_step": e2e_full_ms / (K * e2e_reps),
                    "value_full_frame_every_step": all_ref * e2e_reps / (e2e_full_ms * 1e-3), "fps": world * K * e2e_reps / (e2e_ms * 1e-3),
                    "reps": e2e_reps, "frame_checksum": checksum},
            "gpu_launches": K * reps * launches_per_step,
            "clocks": clocks.result(),
            "occupancy": {"cells": list(cells), "cell_voxels": cell_vox, "empty_cells": int(empty)},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(op, vr, uniforms, lut, W, H, args.workload, sp)
    op.close()
    del op
    if args.workload == "c2" and not args.no_partitioned:
        rec = partitioned_record(vr, rank, world, local, args)
        if line is not None:
            line["partitioned"] = rec
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line:
        print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# CPU legs.  The ONLY places bench.py touches oracle/: the reported cpu_baseline and --impl reference.
# ------------------------------------------------------------------------------------------------

def load_cpu_oracle():
    """oracle/_ref (the reference's shader, camera and gradient compiled as CPU code) when it is
    there, else the plain-C port."""
    from oracle import pyoracle
    try:
        return pyoracle.Oracle("reference"), "reference"
    except Exception:
        return pyoracle.Oracle("port"), "port"


def oracle_orbit_camera(o, k, pitch=0.0):
    cam = o.camera()
    for _ in range(k):
        cam.rotate(4.0, 0.0)
    if pitch:
        cam.rotate(0.0, pitch)
    return cam


def timed_rows(o, frame, vol, lut, stride, start=0):
    t0 = time.perf_counter()
    r = o.render(frame, vol, lut, row_start=start, row_stride=stride)
    return time.perf_counter() - t0, r


# frames whose GPU image is compared with the oracle at the benchmarked size: three pair axes' worth of the orbit
# and one oblique pitch (VERDICT r1 item 1b / SURVEY 8d); (orbit frame, extra Camera::rotate({0, pitch}))
PARITY_VIEWS = [(0, 0.0), (120, 0.0), (240, 0.0), (45, 0.0), (30, 100.0)]


def cpu_baseline(op, vr, uniforms, lut, W, H, workload, sp, budget_s=12.0):
    """The same workload on the host cores, a bounded sample of ~budget_s: frames of the orbit (every
    s-th row of each when one frame alone would exceed the budget).  Then parity at full size: the GPU's frames of
    PARITY_VIEWS against the oracle's, every pixel (or every s-th row when a full frame costs more than ~4 s)."""
    import numpy as np
    o, kind = load_cpu_oracle()
    vol = op.download_volume()
    _, _, mn, mx = op.volume_info()
    k0 = 0
    frame = o.frame(W, H, oracle_orbit_camera(o, k0), mn, mx)
    t_probe, _ = timed_rows(o, frame, vol, lut, 64)
    stride = max(1, int(math.ceil(t_probe * 64 / budget_s)))
    t, r = timed_rows(o, frame, vol, lut, stride)
    tot_t, tot_n, nframes, kk = t, r["stats"].n_ref, 1, k0
    while tot_t < budget_s and nframes < ORBIT_FRAMES:
        kk = (kk + 7) % ORBIT_FRAMES
        t2, r2 = timed_rows(o, o.frame(W, H, oracle_orbit_camera(o, kk), mn, mx), vol, lut, stride)
        tot_t, tot_n, nframes = tot_t + t2, tot_n + r2["stats"].n_ref, nframes + 1
    # parity at the benchmarked size (tolerance: tests/common.py = BASELINE.json north_star)
    pstride = max(1, int(math.ceil(t_probe * 64 / 4.0)))
    views, worst_p, worst_s = [], 0.0, float("inf")
    for k, pitch in PARITY_VIEWS:
        cam = vr.Camera()
        for _ in range(k):
            cam.rotate((4.0, 0.0))
        if pitch:
            cam.rotate((0.0, pitch))
        op.set_camera(cam)
        op.record(sp)
        img = op.download()
        fi = op.last_frame_info()
        want = o.render(o.frame(W, H, oracle_orbit_camera(o, k, pitch), mn, mx), vol, lut, row_start=0, row_stride=pstride)["rgba8"]
        a, b = img[0::pstride].astype(np.int16), want[0::pstride].astype(np.int16)
        d = np.abs(a - b).max(axis=-1).ravel()
        mse = float(np.mean((a - b).astype(np.float64) ** 2))
        p, s = float(np.percentile(d, 99.9)), (float("inf") if mse == 0 else 10.0 * math.log10(255.0 ** 2 / mse))
        worst_p, worst_s = max(worst_p, p), min(worst_s, s)
        views.append({"orbit_frame": k, "pitch_rotate": pitch,
                      "sampled": "3-D array" if fi.layout != vr.LAYOUT_PAIRED else "pairs along " + "xyz"[fi.pair_axis],
                      "p99.9_abs_diff_of_255": p, "max_abs_diff_of_255": int(d.max()), "psnr_db": None if mse == 0 else round(s, 2)})
    return {"value": tot_n / tot_t, "unit": "samples/s", "cores": os.cpu_count(), "kind": kind,
            "sample": f"{nframes} frames of the orbit (frame {k0}, then every 7th), every {stride}-th row of {W}x{H} "
                      f"({tot_n} samples, {tot_t:.1f} s, OpenMP on all {os.cpu_count()} host cores)",
            "parity_vs_gpu_frames": {"tolerance": "<= 2/255 at the 99.9th-percentile pixel, >= 45 dB PSNR (north_star)",
                                     "rows_compared": f"every {pstride}-th row" if pstride > 1 else "all",
                                     "worst_p99.9_abs_diff_of_255": worst_p, "worst_psnr_db": round(worst_s, 2),
                                     "within_tolerance": bool(worst_p <= 2.0 and worst_s >= 45.0), "views": views}}


def run_reference(args):
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    import numpy as np
    from oracle import pyoracle
    o, kind_name = load_cpu_oracle()
    port = pyoracle.Oracle("port")  # synthetic input generator (data only)
    kind, dname, n, W, H, tfn, desc = WORKLOADS[args.workload]
    np_dt = {"u8": np.uint8, "u16": np.uint16, "f32": np.float32}[dname]
    vol = port.synth(kind, SEEDS[args.workload], np_dt, n)
    mn, mx = float(vol.min()), float(vol.max())
    lut = o.discretize(*tf_markers(tfn))
    K, Wu = args.steps, args.warmup
    frames, warm = orbit_frames(K, 0, world), orbit_frames(Wu, 0, world)   # the views rank 0 of our arm renders
    cams = {}

    def frame_of(k):
        if k not in cams:
            cams[k] = o.frame(W, H, oracle_orbit_camera(o, k), mn, mx)
        return cams[k]

    t_probe, _ = timed_rows(o, frame_of(0), vol, lut, 32)
    budget = 150.0
    stride = max(1, int(math.ceil(t_probe * 32 * (K + Wu) / budget)))
    for j, k in enumerate(warm):
        timed_rows(o, frame_of(k), vol, lut, stride, start=j % stride)
    total, t0 = 0, time.perf_counter()
    for j, k in enumerate(frames):
        _, r = timed_rows(o, frame_of(k), vol, lut, stride, start=j % stride)
        total += r["stats"].n_ref
    dt = time.perf_counter() - t0
    v = total / dt
    sample = f"each step = every {stride}-th row of one {W}x{H} orbit frame ({total // K} samples/step; frames spread over the orbit like our arm's), OpenMP on all cores"
    print(json.dumps({
        "impl": "reference", "metric": "ray samples/s @1920x1080, 512^3 u16" if args.workload == "c2" else f"ray samples/s ({args.workload})",
        "value": v, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wu, "ms_per_step": dt * 1e3 / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {desc}", "volume_dtype": dname,
                   "what": "reference shader body (res/shaders/volume.frag) compiled as C++ with its camera and gradient; "
                           "no graphics stack on this box, so this CPU build is the reference arm" if kind_name == "reference"
                           else "plain-C port of the reference shader (oracle/_ref missing)"},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": os.cpu_count(), "kind": kind_name, "sample": sample},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


# ------------------------------------------------------------------------------------------------
# Partitioned workloads (BASELINE.json configs 4 and 5): one frame is rendered by ALL ranks.
# ------------------------------------------------------------------------------------------------

def partitioned_frames(vr, W, H):
    """36-frame orbit, 10 degrees per frame."""
    cam, proj, frames = vr.Camera(), vr.projection(W, H), []
    for _ in range(36):
        frames.append((cam.get_view(), proj, cam.get_position()))
        cam.rotate((40.0, 0.0))
    return frames


def measure_partitioned(vr, mode, rank, world, local, n, W, H, K, Wu, synth=None, tf="B", ess=1, ert=1, tile_px=64, slots=2,
                        merge="p2p", grid=None, alone_sizes=(), balance=0.0):
    """One partitioned configuration, timed by the wall clock over K frames enqueued back to back (frames are handed
    off between the ranks by stream-ordered counters in peer memory), one device sync + barrier at the end, max over
    ranks.  mode "first": volume replicated, image tiles dealt to the ranks; "last": sub-boxes, ordered merge.
    alone_sizes: frame sizes at which rank 0 ALONE also renders the same volume and orbit afterwards (sort-first only:
    the replicated volume is the whole volume) -> the single-GPU denominators measured in the same run."""
    import torch
    import torch.distributed as dist
    from volumetric_renderer_b200 import multigpu as mg

    first = mode == "first"
    synth = synth or (vr.SYNTH_FBM, 0x5EED0004)
    lut = product_lut(vr, tf)
    op = vr.OffscreenPass(W, H, device=local)
    op.transfer_function_changed(lut)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=ess, early_term=ert, early_term_eps=1.0 / 1024.0, collect_stats=1)
    if first:
        op.synth_volume(synth[0], synth[1], vr.U16, n)
        part = mg.SortFirst(op, rank, world, tile_px=tile_px, slots=slots)
    else:
        part = mg.SortLast(op, rank, world, (n, n, n), mode=merge, grid=grid, slots=slots, balance=balance)
        op.synth_subvolume(synth[0], synth[1], vr.U16)
        part.set_global_density_range()
    frames = partitioned_frames(vr, W, H)
    stream = torch.cuda.Stream()
    sp = stream.cuda_stream

    def frame(k):
        u = frames[k % 36]
        op.set_uniforms(*u)
        if first:
            part.render(sp)
        else:
            part.render(u[2], sp)

    # untimed: samples per frame (each rank counts its own tiles / its own samples), bytes stored into the target
    counts = torch.zeros(3, dtype=torch.float64, device="cuda")
    for k in range(K):
        frame(k)
        part.finish()
        st = op.stats()
        remote = op.last_frame_info().target_bytes if (first and rank != 0) else 0
        counts += torch.tensor([float(st.n_ref if first else st.n_inslice), float(st.n_fetch), float(remote)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(counts)
    op.set_options(collect_stats=0)
    for k in range(Wu):
        frame(k)
    part.finish()
    clocks = ClockSampler(physical_gpu_index(local))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks.start()
    t0 = time.perf_counter()
    for k in range(K):
        frame(k)           # enqueue only
    part.finish()          # device sync + barrier: every frame is complete in rank 0's image (slots)
    dt = time.perf_counter() - t0
    clocks.stop_flag = True
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = tt.item()
    checksum = int(np_sum_image(op, part.image_slot())) if rank == 0 else None
    fi = op.last_frame_info()
    out = {"n_gpus": world, "ms_per_step": dt / K * 1e3, "steps": K, "samples_per_s": counts[0].item() / dt,
           "executed_samples_per_s": counts[1].item() / dt, "executed_per_frame": counts[1].item() / K,
           "layout": "3-D array" if fi.layout != vr.LAYOUT_PAIRED else "pairs along " + "xyz"[fi.pair_axis],
           "frame_checksum": checksum, "clocks": clocks.result()}
    if first:
        out["bytes_into_display_rank_per_frame"] = counts[2].item() / K
    else:
        out["grid"] = list(part.grid)
    part.close()
    # the same volume and orbit on rank 0 alone (the other ranks wait): the denominators of the efficiency
    alone = {}
    if first and alone_sizes:
        op.clear_partition()
        for (w1, h1) in alone_sizes:
            if rank == 0:
                if (w1, h1) != op.size:
                    op.framebuffer_size_changed(w1, h1)
                fr = partitioned_frames(vr, w1, h1)
                for k in range(max(Wu, 3)):
                    op.set_uniforms(*fr[k % 36])
                    op.record(sp)
                op.sync()
                t0 = time.perf_counter()
                for k in range(K):
                    op.set_uniforms(*fr[k % 36])
                    op.record(sp)
                op.sync()
                alone[f"{w1}x{h1}"] = (time.perf_counter() - t0) / K * 1e3
            if world > 1:
                dist.barrier()
    out["alone_ms_per_step"] = alone
    op.close()
    return out


def measure_group(vr, world, mode, n, W, H, K, slots):
    """The same partitioned workload through the ONE-PROCESS host (vr_create_group, csrc/group.cu: a context per device,
    frames handed between the devices by CUDA events, one worker thread per rank) -- called on rank 0 only, which drives all
    `world` devices of the box while the other ranks wait.  Wall clock around K frames enqueued back to back + one sync."""
    import numpy as np
    part = vr.PARTITION_SORT_FIRST if mode == "first" else vr.PARTITION_SORT_LAST
    g = vr.PassGroup(list(range(world)), part, W, H, slots)
    try:
        g.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1.0 / 1024.0)
        g.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
        g.transfer_function_changed(product_lut(vr, "B"))
        frames = partitioned_frames(vr, W, H)
        for k in range(6):
            g.set_uniforms(*frames[k % 36])
            g.record()
        g.sync()
        t0 = time.perf_counter()
        for k in range(K):
            g.set_uniforms(*frames[k % 36])
            g.record()
        g.sync()
        dt = (time.perf_counter() - t0) / K
        img = g.download()
        return {"n_gpus": world, "ms_per_step": dt * 1e3, "steps": K, "frame_checksum": int(img.astype(np.int64).sum()), "slots": slots}
    finally:
        g.close()


def partitioned_record(vr, rank, world, local, args):
    """The evidence for BASELINE.json's partitioned configs inside the driver's own run.
    world == 1: the single-GPU times (the denominators of the curve).  world > 1: the correctness check of
    tests/mgpu_check.py, sort-first C4 and sort-last on the same 2048^3 volume (it fits one GPU, so an efficiency
    exists), each against the time rank 0 needs alone in the same run; world == 8 also the full C5."""
    import torch
    import torch.distributed as dist

    rec = {"note": "strong scaling: one frame rendered by all ranks; efficiency = (rank 0 alone, same volume, same orbit, same run) / (N x time on N ranks)"}
    budget_s = float(os.environ.get("VR_BENCH_PART_BUDGET_S", "600"))

    def time_left():
        t = torch.tensor([budget_s - (time.time() - T_START)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return t.item()

    n = args.size or 2048
    K, Wu = 36, 3
    try:
        if world > 1:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import mgpu_check
            from volumetric_renderer_b200 import multigpu as mg
            fails = mgpu_check.run_checks(vr, mg, rank, world, local)
            nf = torch.tensor([len(fails)], device="cuda")
            dist.all_reduce(nf)
            rec["check"] = "OK" if nf.item() == 0 else f"FAIL {fails}"
            rec["check_what"] = "tests/mgpu_check.py on these GPUs: sort-first bit-identical to one GPU (also pipelined, moving camera), sort-last <= 1/255 on < 3 % of pixels (peer-memory and NCCL merge, moving split planes, two view octants)"
        if time_left() < 240:
            rec["skipped"] = "time budget"
            return rec
        c4 = measure_partitioned(vr, "first", rank, world, local, n, 7680, 4320, K, Wu, tile_px=args.tile_px, slots=args.slots,
                                 ess=int(not args.no_ess), ert=int(not args.no_ert), alone_sizes=((7680, 4320), (3840, 2160)))
        n1 = c4["alone_ms_per_step"].get("7680x4320")
        c4["n1_ms_per_step"] = n1
        c4["efficiency"] = (n1 / (world * c4["ms_per_step"])) if n1 else None
        c4["workload"] = f"c4: {n}^3 u16 fBm replicated, 7680x4320, sort-first {args.tile_px}-px tiles, peer stores into rank 0's image, 36-frame orbit"
        rec["c4"] = c4
        n1_4k = c4["alone_ms_per_step"].get("3840x2160")
        if world > 1 and time_left() > 150:
            c5s = measure_partitioned(vr, "last", rank, world, local, n, 3840, 2160, K, Wu, ess=int(not args.no_ess), ert=int(not args.no_ert),
                                      slots=args.slots_last, balance=args.balance)
            c5s["balance"] = args.balance
            c5s["n1_ms_per_step"] = n1_4k
            c5s["efficiency"] = (n1_4k / (world * c5s["ms_per_step"])) if n1_4k else None
            c5s["workload"] = f"sort-last on a volume that fits one GPU: {n}^3 u16 fBm in {world} sub-boxes, 3840x2160, direct-send merge over NVLink, 36-frame orbit"
            rec["sort_last_2048"] = c5s
        elif world == 1:
            rec["sort_last_2048"] = {"n_gpus": 1, "ms_per_step": n1_4k, "n1_ms_per_step": n1_4k, "efficiency": 1.0,
                                     "workload": f"{n}^3 u16 fBm, 3840x2160 on one GPU (the denominator of the sort-last curve)"}
        if world > 1 and time_left() > 120:
            # the one-process host on the same devices (rank 0 drives them all; the other ranks have released their volumes)
            import gc
            gc.collect()
            torch.cuda.empty_cache()
            torch.cuda.synchronize()
            # the other ranks must wait on the HOST: an NCCL barrier is a kernel spinning on their GPUs, which rank 0's
            # kernels on those GPUs would be time-sliced against
            host_group = dist.new_group(backend="gloo")
            dist.barrier(group=host_group)
            grp = {"what": "vr_create_group from rank 0 alone: one process, a context per device, CUDA events between the devices, a worker thread per rank; "
                           "same volume, orbit and frame sizes as c4 / sort_last_2048 above"}
            if rank == 0:
                try:
                    gf = measure_group(vr, world, "first", n, 7680, 4320, 72, args.slots)
                    gf["n1_ms_per_step"] = n1
                    gf["efficiency"] = (n1 / (world * gf["ms_per_step"])) if n1 else None
                    grp["c4"] = gf
                    gl = measure_group(vr, world, "last", n, 3840, 2160, 72, args.slots_last)
                    gl["n1_ms_per_step"] = n1_4k
                    gl["efficiency"] = (n1_4k / (world * gl["ms_per_step"])) if n1_4k else None
                    grp["sort_last_2048"] = gl
                except Exception as e:  # noqa: BLE001
                    grp["error"] = f"{type(e).__name__}: {e}"
            dist.barrier(group=host_group)
            rec["group"] = grp
        if world == 8 and not args.size and time_left() > 200:
            c5 = measure_partitioned(vr, "last", rank, world, local, 4096, 3840, 2160, K, Wu, synth=(vr.SYNTH_BLOBS_FBM, 0x5EED0005), tf="C",
                                     ess=int(not args.no_ess), ert=int(not args.no_ert), slots=args.slots_last, balance=args.balance)
            c5["balance"] = args.balance
            c5["workload"] = "c5: 4096^3 u16 blobs+fBm (128 GiB) in 2x2x2 sub-boxes, TF-C, 3840x2160, sort-last direct-send, 36-frame orbit"
            rec["c5"] = c5
    except Exception as e:  # the C2 line must survive a failure of the extra measurements
        rec["error"] = f"{type(e).__name__}: {e}"
    return rec


def run_partitioned(args):
    """--workload c4: 2048^3 u16 fBm replicated, 7680x4320, sort-first tiles (strong scaling).
    --workload c5: 4096^3 u16 blobs+fBm split into sub-boxes, TF-C, 3840x2160, sort-last + direct-send merge over NVLink."""
    import torch
    import torch.distributed as dist

    import volumetric_renderer_b200 as vr

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    c4 = args.workload == "c4"
    n = args.size or (2048 if c4 else 4096)
    W, H = (7680, 4320) if c4 else (3840, 2160)
    if args.frame:
        W, H = (int(x) for x in args.frame.lower().split("x"))
    K, Wu = args.steps, args.warmup
    dense = args.dense or c4
    m = measure_partitioned(vr, "first" if c4 else "last", rank, world, local, n, W, H, K, Wu,
                            synth=(vr.SYNTH_FBM, 0x5EED0004 if c4 else 0x5EED0005) if dense else (vr.SYNTH_BLOBS_FBM, 0x5EED0005),
                            tf="B" if dense else "C", ess=int(not args.no_ess), ert=int(not args.no_ert), tile_px=args.tile_px,
                            slots=args.slots if c4 else args.slots_last, merge=args.merge, alone_sizes=((W, H),) if (c4 and args.alone) else (),
                            balance=0.0 if c4 else args.balance)
    if rank == 0:
        peaks = load_peaks()
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # algorithmic bytes per GPU and frame: its share of the volume once (c4: the 1/N of the replicated volume
        # its tiles' rays cross; c5: its sub-box), but never more than the 8 texels (16 B) each of its executed
        # samples needs (steps are 10 voxels apart at 2048^3 and 20 at 4096^3: most voxels are never touched),
        # + its share of the RGBA8 image (c4) or its float4 partial and its share of the merge (c5).
        step_s = m["ms_per_step"] * 1e-3
        exec_per_gpu = m["executed_per_frame"] / world
        alg = min(n ** 3 * 2 / world, exec_per_gpu * 16) + W * H * (4 if c4 else 20) / world
        n1 = m["alone_ms_per_step"].get(f"{W}x{H}")
        print(json.dumps({
            "metric": f"ray samples/s ({args.workload})", "value": m["samples_per_s"], "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": Wu, "ms_per_step": m["ms_per_step"], "fps": 1e3 / m["ms_per_step"],
            "n1_ms_per_step": n1, "efficiency": (n1 / (world * m["ms_per_step"])) if n1 else None,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": (f"c4: {n}^3 u16 fBm replicated, {W}x{H}, sort-first {args.tile_px}-px tiles dealt along diagonals, each rank launches its own tiles, peer stores into rank 0's image ({args.slots} slots)"
                                    if c4 else f"c5: {n}^3 u16 {'fBm, TF-B' if dense else 'blobs+fBm, TF-C'} in {world} sub-boxes {m.get('grid')}, {W}x{H}, sort-last, direct-send merge ({args.merge}), split planes: {"fixed" if not args.balance else f"follow the camera within {args.balance} of the edge"}"),
                       "timing": "wall clock over K frames enqueued back to back (frames handed off between ranks by stream-ordered counters in peer memory; nccl merge: collectives), one device sync + barrier at the end, max over ranks",
                       "layout": m["layout"], "l2": "inputs larger than L2"},
            "executed_samples_per_s": m["executed_samples_per_s"],
            "bytes_into_display_rank_per_frame": m.get("bytes_into_display_rank_per_frame"),
            "roofline": {"bound": "hbm", "achieved": alg / step_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": alg / step_s / 1e9 / hbm_peak, "traffic": None, "per": "GPU"},
            "gpu_launches": K * (1 if c4 else 2), "clocks": m["clocks"], "frame_checksum": m["frame_checksum"]}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def np_sum_image(op, slot=0):
    import numpy as np
    return op.download_frame(slot).astype(np.int64).sum()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=ORBIT_FRAMES)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS) + ["c4", "c5"], default="c2")
    ap.add_argument("--size", type=int, default=0, help="c4/c5 and the partitioned record: volume edge override")
    ap.add_argument("--frame", default="", help="c4/c5: WxH override")
    ap.add_argument("--tile-px", type=int, default=64, help="c4: sort-first partition tile edge in pixels")
    ap.add_argument("--balance", type=float, default=0.0, help="sort-last: how far (fraction of the edge) the split planes may follow the camera; "
                    "each rank holds that margin around its sub-box (0 = fixed regular split, the default: measured slower, DESIGN.md section 5)")
    ap.add_argument("--slots-last", type=int, default=4, help="sort-last: raw/image slots per rank (frames a rank may run ahead of the merge)")
    ap.add_argument("--slots", type=int, default=2, help="c4: image slots on the display rank (frames a rank may run ahead)")
    ap.add_argument("--full-download", action="store_true", help="e2e: copy the whole frame to the host every step")
    ap.add_argument("--merge", choices=["p2p", "nccl"], default="p2p", help="c5: peer-memory merge or NCCL exchange")
    ap.add_argument("--dense", action="store_true", help="c5: fBm with TF-B in every sub-box (nothing to skip) instead of blobs+fBm with TF-C")
    ap.add_argument("--alone", action="store_true", help="c4: also time rank 0 alone on the same volume (efficiency in the line)")
    ap.add_argument("--exact", action="store_true", help="verification sampler instead of hardware trilinear")
    ap.add_argument("--no-ess", action="store_true")
    ap.add_argument("--no-ert", action="store_true")
    ap.add_argument("--preintegration", type=int, default=0, help="time the pre-integrated classification mode with this many sub-samples "
                    "(quality mode, changes the image; use with --no-cpu-baseline --no-partitioned)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-partitioned", action="store_true", help="c2: leave out the partitioned record (configs 4/5 evidence)")
    ap.add_argument("--in-flight", type=int, default=3, help="frames in flight (streams), 1..4")
    ap.add_argument("--layout", choices=["auto", "3d", "paired"], default="auto", help="volume layout (vr_set_volume_layout)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.workload in ("c4", "c5"):
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "the reference is single-GPU and holds the volume as float: configs 4/5 do not fit it"}))
            return
        if args.steps == ORBIT_FRAMES:
            args.steps = 36
        run_partitioned(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
