#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

Metric: reference-equivalent ray samples/s (= iterations of res/shaders/volume.frag:33-48 that pass
the bounds test, what the reference shader would execute for the same frames) on config C2:
512^3 uint16 fractal-noise volume, 1920x1080, 360-frame orbit (frame k = default camera after k x
Camera::rotate({4,0})), transfer function TF-B.

A step = one frame = one launch of K1 (march).  N GPUs: the orbit's frames are independent units;
rank r renders frames r, r+N, ... (K frames per rank, weak scaling, no data-path collective).

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

WORKLOADS = {
    # name: (synth kind, dtype name, n, W, H, TF, description)
    "c1": (0, "u8", 256, 512, 512, "A", "256^3 u8 Marschner-Lobb, 512x512, TF-A (reference default gradient)"),
    "c2": (1, "u16", 512, 1920, 1080, "B", "512^3 u16 fBm, 1920x1080, 360-frame orbit, TF-B"),
    "c3": (2, "u16", 1024, 3840, 2160, "C", "1024^3 u16 sparse blobs, 3840x2160, 360-frame orbit, TF-C"),
}
ORBIT_FRAMES = 360


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


def measured_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_march launch of this workload, from the committed
    ncu --set full capture (profiles/traffic.json names the summary it was read from); None if not captured."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[workload]["dram_bytes_per_launch"]
    except Exception:
        return None


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


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
    op.synth_volume(kind, {"c1": 0x5EED0001, "c2": 0x5EED0002, "c3": 0x5EED0003}[args.workload], dtype, n)
    lut = product_lut(vr, tfn)
    op.transfer_function_changed(lut)
    prod = dict(sampler=vr.SAMPLER_EXACT if args.exact else vr.SAMPLER_HW, empty_skip=int(not args.no_ess),
                early_term=int(not args.no_ert), early_term_eps=1.0 / 1024.0)
    uniforms = orbit_uniforms(vr, W, H, ORBIT_FRAMES)
    my_frames = [(rank + j * world) % ORBIT_FRAMES for j in range(K)]
    warm_frames = [(rank + j * world) % ORBIT_FRAMES for j in range(Wu)]
    F = max(1, min(args.in_flight, 4))      # frames in flight (the reference keeps MAX_FRAMES_IN_FLIGHT = 2)
    streams = [torch.cuda.Stream() for _ in range(F)]
    stream, sp = streams[0], streams[0].cuda_stream

    # ---- untimed: work per frame (reference-equivalent samples, executed fetches) ----------------
    n_ref, n_exec = {}, {}
    for k in sorted(set(my_frames)):
        op.set_uniforms(*uniforms[k])
        op.set_options(collect_stats=1, write_raw=0, sampler=prod["sampler"], empty_skip=0, early_term=0,
                       early_term_eps=prod["early_term_eps"])
        op.record(sp)
        n_ref[k] = op.stats().n_ref
        op.set_options(collect_stats=1, **prod)
        op.record(sp)
        n_exec[k] = op.stats().n_fetch
    op.set_options(collect_stats=0, **prod)
    cells, cell_vox, empty = op.occupancy_info()
    total_ref = sum(n_ref[k] for k in my_frames)
    total_exec = sum(n_exec[k] for k in my_frames)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed: inputs resident, K launches of K1 --------------------------------------------
    for j, k in enumerate(warm_frames):
        op.set_uniforms(*uniforms[k])
        op.record(streams[j % F].cuda_stream, j % F)
    clocks = ClockSampler(physical_gpu_index(local))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def fork():
        ev0.record(streams[0])
        for s in streams[1:]:
            s.wait_event(ev0)

    def join():
        for s in streams[1:]:
            e = torch.cuda.Event()
            e.record(s)
            streams[0].wait_event(e)
        ev1.record(streams[0])

    barrier()
    clocks.start()
    fork()
    for j, k in enumerate(my_frames):
        op.set_uniforms(*uniforms[k])
        op.record(streams[j % F].cuda_stream, j % F)
    join()
    clocks.sample()  # at least one sample while the queue is still draining
    barrier()
    clocks.stop_flag = True
    dev_ms = ev0.elapsed_time(ev1)

    # ---- end to end through the public API: host uniforms in, RGBA8 frame out to pinned host memory ---
    host = [torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True) for _ in range(F)]
    # Each step leaves the complete frame in pinned host memory.  --full-download: the whole RGBA8 image is copied
    # every step; default: vr_download_frame_dirty_async copies the rectangle that can differ from what that host
    # buffer already holds (outside the cube's screen box a frame is the clear colour) -- d2h bytes are counted.
    def run_e2e(fetch):
        """K steps; every step: uniforms from host memory, render, the frame brought up to date in pinned host memory."""
        for j, k in enumerate(warm_frames[:F] if len(warm_frames) >= F else list(warm_frames) * F):
            op.set_uniforms(*uniforms[k])
            op.record(streams[j % F].cuda_stream, j % F)
            fetch(host[j % F].data_ptr(), streams[j % F].cuda_stream, j % F)   # the first delivery into a buffer is always full
        barrier()
        nbytes = 0
        t0 = time.perf_counter()
        for j, k in enumerate(my_frames):
            f = j % F
            streams[f].synchronize()           # frame j-F is in host memory (its buffer may be reused now)
            op.set_uniforms(*uniforms[k])      # 192-byte uniform block: the step's input, goes down with the launch
            op.record(streams[f].cuda_stream, f)
            nbytes += fetch(host[f].data_ptr(), streams[f].cuda_stream, f) or W * H * 4
        barrier()                              # every frame of the K steps has landed in host memory
        ms = (time.perf_counter() - t0) * 1e3
        # the host frames must be what a full download gives
        for f in range(min(F, K)):
            full = torch.empty((H, W, 4), dtype=torch.uint8, pin_memory=True)
            op.download_async(full.data_ptr(), streams[f].cuda_stream, f)
            streams[f].synchronize()
            if not torch.equal(full, host[f]):
                raise SystemExit(f"bench.py: host frame of slot {f} differs from a full download")
        return ms, nbytes

    # the same K steps with the whole frame copied every step (PCIe-bound), reported beside the headline
    e2e_full_ms, _ = run_e2e(op.download_async)
    if args.full_download:
        e2e_ms, d2h_bytes = e2e_full_ms, K * W * H * 4
    else:
        e2e_ms, d2h_bytes = run_e2e(op.download_dirty_async)
    host = host[(K - 1) % F]
    checksum = int(host.view(torch.int32).sum().item())

    # ---- what the memory system alone allows: the same frames through the fetch-only build of the kernel ---
    # (vr_bench_fetch_only: ray setup, stepping and every volume fetch of the frame, no transfer function and
    # no compositing; a measurement aid, reported beside the rooflines, not part of `value`)
    fetch_only_ms = None
    if world == 1 and not args.exact:
        try:
            op.bench_fetch_only(True)
            for j, k in enumerate(warm_frames[:3]):
                op.set_uniforms(*uniforms[k])
                op.record(streams[j % F].cuda_stream, j % F)
            barrier()
            fork()
            for j, k in enumerate(my_frames):
                op.set_uniforms(*uniforms[k])
                op.record(streams[j % F].cuda_stream, j % F)
            join()
            barrier()
            fetch_only_ms = ev0.elapsed_time(ev1) / K
        except vr.VrError:
            fetch_only_ms = None     # empty-space skipping is active for this data: no fetch-only build
        finally:
            op.bench_fetch_only(False)

    t = torch.tensor([dev_ms, e2e_ms, float(total_ref), float(total_exec)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, e2e_ms = mx[0].item(), mx[1].item()
        all_ref, all_exec = sm[2].item(), sm[3].item()
    else:
        all_ref, all_exec = float(total_ref), float(total_exec)

    line = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        ncell = cells[0] * cells[1] * cells[2]
        occ_frac = (ncell - empty) / ncell if prod["empty_skip"] else 1.0
        alg_bytes = n ** 3 * esz * occ_frac + W * H * 4 + 1024   # touched bricks once + RGBA8 out + LUT
        step_s = dev_ms * 1e-3 / K                                # one launch per step
        paired = op.last_frame_paired()
        tex_peak = vr.bench_tex_peak(local, dtype, 2 if paired else 1)
        tex3d_peak = vr.bench_tex_peak(local, dtype, 1) if paired else tex_peak   # one 3-D trilinear fetch per sample
        exec_rate = (total_exec / K) / step_s
        line = {
            "metric": "ray samples/s @1920x1080, 512^3 u16" if args.workload == "c2" else f"ray samples/s ({args.workload})",
            "value": all_ref / (dev_ms * 1e-3), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wu,
            "ms_per_step": dev_ms / K, "fps": world * K / (dev_ms * 1e-3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "volume_dtype": dname,
                       "sampler": "exact(8 point fetches + fp32 lerp)" if args.exact else "hw trilinear",
                       "empty_space_skipping": bool(prod["empty_skip"]), "early_ray_termination_eps": prod["early_term_eps"] if prod["early_term"] else 0,
                       "l2": f"inputs larger than L2 ({n ** 3 * esz / 2 ** 20:.0f} MiB volume, no flush)",
                       "frames_in_flight": F,
                       "layout": "paired z-slices, 2-D layered array (+ the 3-D array kept for the exact sampler)" if paired else "3-D array",
                       "parallelism": f"frames of the orbit dealt round-robin to {world} GPU(s), no collective"},
            "executed_samples_per_s": all_exec / (dev_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": alg_bytes / step_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": alg_bytes / step_s / 1e9 / hbm_peak, "traffic": measured_traffic(args.workload),
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback",
                         "algorithmic_bytes_per_launch": alg_bytes},
            "roofline_tex": {"bound": "tex", "achieved": exec_rate / 1e9, "peak": tex_peak / 1e9, "unit": "Gfetch/s",
                             "frac": exec_rate / tex_peak,
                             "peak_source": "vr_bench_tex_peak: L1-resident " + ("paired-layout bilinear fetches (2-D layered, two z-slices per texel)" if paired else "3-D trilinear fetches") + ", measured in this run",
                             "frac_of_3d_trilinear_peak": exec_rate / tex3d_peak,
                             "fetch_only_ms_per_step": fetch_only_ms,
                             "fetch_only_share_of_step": (fetch_only_ms / (dev_ms / K)) if fetch_only_ms else None},
            "e2e": {"value": all_ref / (e2e_ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": 192,
                    "d2h_bytes_per_step": d2h_bytes / K, "frame_bytes": W * H * 4,
                    "d2h": "full frame every step" if args.full_download else "dirty rectangle (vr_download_frame_dirty_async); the host frame is complete and checked against a full download",
                    "ms_per_step": e2e_ms / K, "ms_per_step_full_frame_every_step": e2e_full_ms / K,
                    "value_full_frame_every_step": all_ref / (e2e_full_ms * 1e-3) if world == 1 else None, "fps": world * K / (e2e_ms * 1e-3),
                    "frame_checksum": checksum},
            "gpu_launches": K,
            "clocks": clocks.result(),
            "occupancy": {"cells": list(cells), "cell_voxels": cell_vox, "empty_cells": int(empty)},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(op, vr, uniforms, lut, W, H, my_frames[0], sp)
    op.close()
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


def oracle_orbit_camera(o, k):
    cam = o.camera()
    for _ in range(k):
        cam.rotate(4.0, 0.0)
    return cam


def timed_rows(o, frame, vol, lut, stride, start=0):
    t0 = time.perf_counter()
    r = o.render(frame, vol, lut, row_start=start, row_stride=stride)
    return time.perf_counter() - t0, r


def cpu_baseline(op, vr, uniforms, lut, W, H, k, sp, budget_s=12.0):
    """The same workload on the host cores, a bounded sample of ~budget_s: frames of the orbit (every
    s-th row of each when one frame alone would exceed the budget), the first one also checked against
    the GPU's frame."""
    import numpy as np
    o, kind = load_cpu_oracle()
    vol = op.download_volume()
    _, _, mn, mx = op.volume_info()
    frame = o.frame(W, H, oracle_orbit_camera(o, k), mn, mx)
    t_probe, _ = timed_rows(o, frame, vol, lut, 64)
    stride = max(1, int(math.ceil(t_probe * 64 / budget_s)))
    t, r = timed_rows(o, frame, vol, lut, stride)
    tot_t, tot_n, nframes, kk = t, r["stats"].n_ref, 1, k
    while tot_t < budget_s and nframes < ORBIT_FRAMES:
        kk = (kk + 7) % ORBIT_FRAMES
        t2, r2 = timed_rows(o, o.frame(W, H, oracle_orbit_camera(o, kk), mn, mx), vol, lut, stride)
        tot_t, tot_n, nframes = tot_t + t2, tot_n + r2["stats"].n_ref, nframes + 1
    # parity of the very frame the GPU renders, on the sampled rows (tolerance: tests/common.py)
    op.set_uniforms(*uniforms[k])
    op.record(sp)
    img = op.download()
    a, b = img[0::stride].astype(np.int16), r["rgba8"][0::stride].astype(np.int16)
    d = np.abs(a - b).max(axis=-1).ravel()
    mse = float(np.mean((a - b).astype(np.float64) ** 2))
    return {"value": tot_n / tot_t, "unit": "samples/s", "cores": os.cpu_count(), "kind": kind,
            "sample": f"{nframes} frames of the orbit (frame {k}, then every 7th), every {stride}-th row of {W}x{H} "
                      f"({tot_n} samples, {tot_t:.1f} s, OpenMP on all {os.cpu_count()} host cores)",
            "parity_vs_gpu_frame": {"p99.9_abs_diff_of_255": float(np.percentile(d, 99.9)),
                                    "psnr_db": None if mse == 0 else 10.0 * math.log10(255.0 ** 2 / mse)}}


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
    vol = port.synth(kind, {"c1": 0x5EED0001, "c2": 0x5EED0002, "c3": 0x5EED0003}[args.workload], np_dt, n)
    mn, mx = float(vol.min()), float(vol.max())
    lut = o.discretize(*tf_markers(tfn))
    K, Wu = args.steps, args.warmup
    cams = {}

    def frame_of(k):
        if k not in cams:
            cams[k] = o.frame(W, H, oracle_orbit_camera(o, k), mn, mx)
        return cams[k]

    t_probe, _ = timed_rows(o, frame_of(0), vol, lut, 32)
    budget = 150.0
    stride = max(1, int(math.ceil(t_probe * 32 * (K + Wu) / budget)))
    for j in range(Wu):
        timed_rows(o, frame_of((j * world) % ORBIT_FRAMES), vol, lut, stride, start=j % stride)
    total, t0 = 0, time.perf_counter()
    for j in range(K):
        _, r = timed_rows(o, frame_of((j * world) % ORBIT_FRAMES), vol, lut, stride, start=j % stride)
        total += r["stats"].n_ref
    dt = time.perf_counter() - t0
    v = total / dt
    sample = f"each step = every {stride}-th row of one {W}x{H} orbit frame ({total // K} samples/step), OpenMP on all cores"
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

def run_partitioned(args):
    """c4: 2048^3 u16 fBm replicated, 7680x4320, sort-first tiles (strong scaling).
    c5: 4096^3 u16 fBm split into sub-boxes, 3840x2160, sort-last + direct-send merge over NVLink."""
    import torch
    import torch.distributed as dist

    import volumetric_renderer_b200 as vr
    from volumetric_renderer_b200 import multigpu as mg

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
    lut = product_lut(vr, "B")
    op = vr.OffscreenPass(W, H, device=local)
    op.transfer_function_changed(lut)
    op.set_options(sampler=vr.SAMPLER_HW, empty_skip=int(not args.no_ess), early_term=int(not args.no_ert),
                   early_term_eps=1.0 / 1024.0, collect_stats=1)
    if c4:
        op.synth_volume(vr.SYNTH_FBM, 0x5EED0004, vr.U16, n)
        part = mg.SortFirst(op, rank, world, tile_px=args.tile_px, slots=args.slots)
    else:
        part = mg.SortLast(op, rank, world, (n, n, n), mode=args.merge)
        op.synth_subvolume(vr.SYNTH_FBM, 0x5EED0005, vr.U16)
        part.set_global_density_range()
    # 36-frame orbit, 10 degrees per frame
    cam, proj, frames = vr.Camera(), vr.projection(W, H), []
    for _ in range(36):
        frames.append((cam.get_view(), proj, cam.get_position()))
        cam.rotate((40.0, 0.0))
    stream = torch.cuda.Stream()
    sp = stream.cuda_stream

    def frame(k):
        u = frames[k % 36]
        op.set_uniforms(*u)
        if c4:
            part.render(sp)
        else:
            part.render(u[2], sp)

    # untimed: samples per frame (each rank counts its own tiles / its own samples)
    counts = torch.zeros(2, dtype=torch.float64, device="cuda")
    for k in range(K):
        frame(k)
        part.finish()
        st = op.stats()
        counts += torch.tensor([float(st.n_ref if c4 else st.n_inslice), float(st.n_fetch)], dtype=torch.float64, device="cuda")
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
        frame(k)           # enqueue only: frames are handed off between the ranks by stream-ordered counters
    part.finish()          # device sync + barrier: every frame is complete in rank 0's image (slots)
    dt = time.perf_counter() - t0
    clocks.stop_flag = True
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = tt.item()
    checksum = None
    if rank == 0:
        checksum = int(np_sum_image(op))
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # algorithmic bytes per GPU and frame: its share of the volume once (c4: the 1/N of the replicated volume
        # its tiles' rays cross; c5: its sub-box), but never more than the 8 texels (16 B) each of its executed
        # samples needs (steps are 10 voxels apart at 2048^3 and 20 at 4096^3: most voxels are never touched),
        # + its share of the RGBA8 image (c4) or its float4 partial and its share of the merge (c5).
        step_s = dt / K
        exec_per_gpu = counts[1].item() / K / world
        alg = min(n ** 3 * 2 / world, exec_per_gpu * 16) + W * H * (4 if c4 else 20) / world
        print(json.dumps({
            "metric": f"ray samples/s ({args.workload})", "value": counts[0].item() / dt, "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": Wu, "ms_per_step": step_s * 1e3, "fps": K / dt,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": (f"c4: {n}^3 u16 fBm replicated, {W}x{H}, sort-first {args.tile_px}-px tiles dealt along diagonals, each rank launches its own tiles, peer stores into rank 0's image ({args.slots} slots)"
                                    if c4 else f"c5: {n}^3 u16 fBm in {world} sub-boxes {part.grid}, {W}x{H}, sort-last, direct-send merge ({args.merge})"),
                       "timing": "wall clock over K frames enqueued back to back (frames handed off between ranks by stream-ordered counters in peer memory; nccl merge: collectives), one device sync + barrier at the end, max over ranks",
                       "l2": "inputs larger than L2"},
            "executed_samples_per_s": counts[1].item() / dt,
            "roofline": {"bound": "hbm", "achieved": alg / step_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": alg / step_s / 1e9 / hbm_peak, "traffic": None, "per": "GPU"},
            "gpu_launches": K * (1 if c4 else 2), "clocks": clocks.result(), "frame_checksum": checksum}), flush=True)
    part.close()
    op.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def np_sum_image(op):
    import numpy as np
    return op.download().astype(np.int64).sum()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=ORBIT_FRAMES)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS) + ["c4", "c5"], default="c2")
    ap.add_argument("--size", type=int, default=0, help="c4/c5: volume edge override")
    ap.add_argument("--frame", default="", help="c4/c5: WxH override")
    ap.add_argument("--tile-px", type=int, default=64, help="c4: sort-first partition tile edge in pixels")
    ap.add_argument("--slots", type=int, default=2, help="c4: image slots on the display rank (frames a rank may run ahead)")
    ap.add_argument("--full-download", action="store_true", help="e2e: copy the whole frame to the host every step")
    ap.add_argument("--merge", choices=["p2p", "nccl"], default="p2p", help="c5: peer-memory merge or NCCL exchange")
    ap.add_argument("--exact", action="store_true", help="verification sampler instead of hardware trilinear")
    ap.add_argument("--no-ess", action="store_true")
    ap.add_argument("--no-ert", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
