"""volumetric-renderer_b200 -- B200-native ray-march backend, host-side mirror of the
reference's interface for this path.

The compute lives in libvr_b200.so (CUDA kernels for sm_100a behind the C ABI of
include/vr.h).  This module mirrors the reference's public surface with the same names and
argument meaning so parity tests read like the reference's call sites:

    Vol::Rendering::OffscreenPass   src/rendering/offscreen_pass.h:27-148   -> OffscreenPass
    Vol::Scene::Camera              src/scene/camera.h:9-23                 -> Camera
    Vol::Data::Dataset              src/data/dataset.h:9-13                 -> Dataset
    Vol::Data::NrrdFileParser       src/data/nrrd_file_parser.cpp:21-47     -> OffscreenPass.import_nrrd
    Vol::UI::Components::Gradient   src/ui/components/gradient.h:11-36      -> Gradient

Nothing here computes pixels: every method is a thin call into the C ABI.  Import as
`volumetric_renderer_b200` (shim at the repo root; the directory name has a hyphen).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import (F32, U8, U16, LAYOUT_ARRAY3D, LAYOUT_AUTO, LAYOUT_PAIRED, PARTITION_REPLICAS, PARTITION_SORT_FIRST, PARTITION_SORT_LAST, SAMPLER_EXACT, SAMPLER_HW, SYNTH_BLOBS, SYNTH_BLOBS_FBM,
                   SYNTH_FBM, SYNTH_MARSCHNER_LOBB, FrameInfo, Options, Stats)

lib = _lib.load()
KERNEL_VERSION = lib.vr_kernel_version().decode()

_NP = {U8: np.uint8, U16: np.uint16, F32: np.float32}
_CODE = {np.dtype(np.uint8): U8, np.dtype(np.uint16): U16, np.dtype(np.float32): F32}


class VrError(RuntimeError):
    """The reference throws std::runtime_error on failure; this is its counterpart."""

    def __init__(self, status: int, message: str):
        super().__init__(f"[vr status {status}] {message}")
        self.status = status


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Camera:
    """Vol::Scene::Camera (src/scene/camera.cpp:7-48)."""

    def __init__(self):
        self._h = lib.vr_camera_new()
        if not self._h:
            raise MemoryError("vr_camera_new")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.vr_camera_free(h)

    def rotate(self, delta) -> "Camera":
        lib.vr_camera_rotate(self._h, float(delta[0]), float(delta[1]))
        return self

    def zoom(self, delta: float) -> "Camera":
        lib.vr_camera_zoom(self._h, float(delta))
        return self

    def get_position(self) -> np.ndarray:
        out = np.zeros(3, np.float32)
        lib.vr_camera_get_position(self._h, _ptr(out))
        return out

    def get_view(self) -> np.ndarray:
        out = np.zeros(16, np.float32)
        lib.vr_camera_get_view(self._h, _ptr(out))
        return out


def projection(width: int, height: int) -> np.ndarray:
    """proj of OffscreenPass::update_uniform_buffer (offscreen_pass.cpp:1154-1168)."""
    out = np.zeros(16, np.float32)
    lib.vr_projection(width, height, _ptr(out))
    return out


@dataclass
class Gradient:
    """Vol::UI::Components::Gradient (gradient.h:11-36); default = gradient.cpp:64-70."""
    color_markers: list = field(default_factory=lambda: [(0.0, (0.0, 0.0, 0.0)), (1.0, (1.0, 1.0, 1.0))])
    alpha_markers: list = field(default_factory=lambda: [(0.0, 1.0), (1.0, 1.0)])

    def discretize(self, count: int = 256) -> np.ndarray:
        cm = np.ascontiguousarray([[loc, *rgb] for loc, rgb in self.color_markers], np.float32)
        am = np.ascontiguousarray([[loc, a] for loc, a in self.alpha_markers], np.float32)
        out = np.zeros(count, np.uint32)
        rc = lib.vr_gradient_discretize(_ptr(cm), len(cm), _ptr(am), len(am), count, _ptr(out))
        if rc != _lib.VR_OK:
            raise VrError(rc, "vr_gradient_discretize: bad markers")
        return out


@dataclass
class Dataset:
    """Vol::Data::Dataset (src/data/dataset.h:9-13); `data` is [nz, ny, nx], x fastest.
    uint8 / uint16 / float32 arrays go to the device as they are."""
    data: np.ndarray
    min: float | None = None
    max: float | None = None

    @property
    def dimensions(self):
        return (self.data.shape[2], self.data.shape[1], self.data.shape[0])


class OffscreenPass:
    """Vol::Rendering::OffscreenPass with the Vulkan objects replaced by a CUDA context."""

    def __init__(self, width: int, height: int, device: int = 0):
        h = C.c_void_p()
        rc = lib.vr_create(device, width, height, C.byref(h))
        if rc != _lib.VR_OK:
            raise VrError(rc, lib.vr_last_error(None).decode())
        self._h = h

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.vr_destroy(h)

    __del__ = close

    def _check(self, rc: int):
        if rc != _lib.VR_OK:
            raise VrError(rc, lib.vr_last_error(self._h).decode())

    # ---- the reference's members ------------------------------------------------------
    def record(self, stream: int | None = None, frame_index: int = 0):
        """record(command_buffer, frame_index), offscreen_pass.cpp:163-230 (asynchronous)."""
        self._check(lib.vr_render_frame(self._h, frame_index, stream))

    def framebuffer_size_changed(self, width: int, height: int):
        self._check(lib.vr_resize(self._h, width, height))

    def load_csv(self, paths):
        """Importer::import(FileFormat::CSV) without the dialog (src/data/importer.cpp:30-46)."""
        arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
        self._check(lib.vr_load_csv(self._h, arr, len(paths)))

    def volume_dataset_changed(self, dataset: Dataset):
        a = dataset.data
        if a.dtype not in _CODE:
            a = a.astype(np.float32)  # nrrd_file_parser.cpp:67-77 widens everything else
        a = np.ascontiguousarray(a)
        mn = C.c_float(dataset.min) if dataset.min is not None else None
        mx = C.c_float(dataset.max) if dataset.max is not None else None
        nz, ny, nx = a.shape
        self._check(lib.vr_set_volume(self._h, _ptr(a), _CODE[a.dtype], nx, ny, nz,
                                      C.byref(mn) if mn is not None else None,
                                      C.byref(mx) if mx is not None else None))

    def slicing_changed(self, min_slice, max_slice):
        lo, hi = np.asarray(min_slice, np.float32), np.asarray(max_slice, np.float32)
        self._check(lib.vr_set_slicing(self._h, _ptr(lo), _ptr(hi)))

    def transfer_function_changed(self, data):
        d = np.ascontiguousarray(data, np.uint32)
        self._check(lib.vr_set_transfer_function(self._h, _ptr(d), len(d)))

    def transfer_function_changed_device(self, dev_ptr: int, n: int):
        """vr_set_transfer_function_device: n packed texels already in device memory."""
        self._check(lib.vr_set_transfer_function_device(self._h, C.c_void_p(dev_ptr), n))

    def gradient_changed(self, gradient: "Gradient", count: int = 256):
        """vr_set_gradient: Gradient::discretize on the device + transfer_function_changed (what main_window.cpp:250-257
        does on every edit of the gradient)."""
        cm = np.ascontiguousarray([[loc, *rgb] for loc, rgb in gradient.color_markers], np.float32)
        am = np.ascontiguousarray([[loc, a] for loc, a in gradient.alpha_markers], np.float32)
        self._check(lib.vr_set_gradient(self._h, _ptr(cm), len(cm), _ptr(am), len(am), count))

    def get_transfer_function(self) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(lib.vr_get_transfer_function(self._h, None, 0, C.byref(n)))
        out = np.zeros(n.value, np.uint32)
        self._check(lib.vr_get_transfer_function(self._h, _ptr(out), n.value, None))
        return out

    # ---- the implicit camera pull of update_uniform_buffer, made a push ------------------
    def set_camera(self, camera: Camera):
        self._check(lib.vr_set_camera_orbit(self._h, camera._h))

    def set_uniforms(self, view, proj, camera_position):
        v, p, c = (np.ascontiguousarray(x, np.float32) for x in (view, proj, camera_position))
        self._check(lib.vr_set_camera(self._h, _ptr(v), _ptr(p), _ptr(c)))

    # ---- loader fast path, knobs ---------------------------------------------------------
    def import_nrrd(self, path: str):
        """Importer::import(Nrrd) -> parse -> volume_dataset_changed (importer.cpp:36-46)."""
        self._check(lib.vr_load_nrrd(self._h, str(path).encode()))

    def synth_volume(self, kind: int, seed: int, dtype: int, n):
        nx, ny, nz = (n, n, n) if isinstance(n, int) else n
        self._check(lib.vr_synth_volume(self._h, kind, seed & 0xFFFFFFFF, dtype, nx, ny, nz))

    def set_volume_layout(self, layout: int):
        """Layout of the NEXT volume upload: LAYOUT_AUTO / LAYOUT_ARRAY3D / LAYOUT_PAIRED."""
        self._check(lib.vr_set_volume_layout(self._h, int(layout)))

    def volume_layout(self) -> int:
        """LAYOUT_PAIRED if the paired copy is resident next to the 3-D array, else LAYOUT_ARRAY3D."""
        v = C.c_int(0)
        self._check(lib.vr_get_volume_layout(self._h, C.byref(v), None))
        return v.value

    def last_frame_paired(self) -> bool:
        v, f = C.c_int(0), C.c_int(0)
        self._check(lib.vr_get_volume_layout(self._h, C.byref(v), C.byref(f)))
        return bool(f.value)

    def last_frame_pair_axis(self) -> int | None:
        """Model axis (0 x, 1 y, 2 z) the last frame's pairs ran along; None if it sampled the 3-D array."""
        v, f = C.c_int(0), C.c_int(0)
        self._check(lib.vr_get_volume_layout(self._h, C.byref(v), C.byref(f)))
        return f.value - 1 if f.value else None

    def last_frame_info(self) -> FrameInfo:
        """vr_get_last_frame_info: layout / pair axis / skipping / launch shape of the frame recorded last."""
        fi = FrameInfo()
        self._check(lib.vr_get_last_frame_info(self._h, C.byref(fi)))
        return fi

    def tile_mask(self, frame_index: int = 0):
        """vr_get_tile_mask: (bool array [tiles_y, tiles_x] of the tiles the march set rays up in, ignored flag, tile height);
        None when the frame recorded last into that slot used no mask."""
        import numpy as np
        tx, ty, th = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        if lib.vr_get_tile_mask(self._h, frame_index, None, 0, C.byref(tx), C.byref(ty), C.byref(th)) != 0:
            return None
        n = (tx.value * ty.value + 31) // 32 + 1
        words = np.zeros(n, np.uint32)
        self._check(lib.vr_get_tile_mask(self._h, frame_index, words.ctypes.data_as(C.POINTER(C.c_uint32)), n, None, None, None))
        bits = np.unpackbits(words[:-1].view(np.uint8), bitorder="little")[: tx.value * ty.value].astype(bool)
        return bits.reshape(ty.value, tx.value), bool(words[-1]), th.value

    def set_volume_spacing(self, spacing):
        """vr_set_volume_spacing: honour per-axis sample distances (non-cubic voxels); None = the reference's unit cube."""
        if spacing is None:
            self._check(lib.vr_set_volume_spacing(self._h, None))
        else:
            s = np.asarray(spacing, np.float32)
            self._check(lib.vr_set_volume_spacing(self._h, _ptr(s)))

    def volume_spacing(self):
        """(spacings the NRRD loaded last states, NaN where none; extents of the box in use)"""
        f, e = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self._check(lib.vr_get_volume_spacing(self._h, _ptr(f), _ptr(e)))
        return f, e

    def set_preintegration(self, substeps: int):
        """vr_set_preintegration: classify segments between consecutive samples from a 2-D table (quality mode, changes the
        image, no reference counterpart); 0 = off."""
        self._check(lib.vr_set_preintegration(self._h, substeps))

    def preintegration_table(self) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(lib.vr_get_preintegration_table(self._h, None, 0, C.byref(n)))
        tab = np.zeros((n.value, n.value, 4), np.float32)
        self._check(lib.vr_get_preintegration_table(self._h, _ptr(tab), n.value * n.value, None))
        return tab

    def set_march(self, ray_dist: float, step_size: float):
        self._check(lib.vr_set_march(self._h, ray_dist, step_size))

    def set_density_range(self, mn: float, mx: float):
        self._check(lib.vr_set_density_range(self._h, mn, mx))

    @property
    def options(self) -> Options:
        o = Options()
        self._check(lib.vr_get_options(self._h, C.byref(o)))
        return o

    def set_options(self, **kw):
        o = self.options
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(k)
            setattr(o, k, v)
        self._check(lib.vr_set_options(self._h, C.byref(o)))

    # ---- outputs -------------------------------------------------------------------------
    @property
    def size(self):
        w, h = C.c_uint32(), C.c_uint32()
        lib.vr_get_size(self._h, C.byref(w), C.byref(h))
        return w.value, h.value

    def sync(self):
        self._check(lib.vr_sync(self._h))

    def get_image(self):
        """get_image_view(): (device pointer, pitch in bytes) of the RGBA8 target."""
        p, pitch = C.c_void_p(), C.c_size_t()
        self._check(lib.vr_get_image(self._h, C.byref(p), C.byref(pitch)))
        return p.value, pitch.value

    def get_image_frame(self, frame_index: int) -> int:
        """Device pointer of the RGBA8 target of a frame slot (allocated on first use)."""
        if frame_index:
            self.ipc_export(16 + frame_index)   # allocates the slot
        p, pitch = C.c_void_p(), C.c_size_t()
        self._check(lib.vr_get_image_frame(self._h, frame_index, C.byref(p), C.byref(pitch)))
        return p.value

    def get_raw(self, frame_index: int = 0) -> int:
        p = C.c_void_p()
        self._check(lib.vr_get_raw_frame(self._h, frame_index, C.byref(p)))
        return p.value

    def download(self, out: np.ndarray | None = None) -> np.ndarray:
        w, h = self.size
        if out is None:
            out = np.empty((h, w, 4), np.uint8)
        self._check(lib.vr_download(self._h, _ptr(out)))
        return out

    def present(self, dst_ptr: int, width: int, height: int, pitch_bytes: int = 0, stream: int | None = None, frame_index: int = 0):
        """The viewport's LINEAR / CLAMP_TO_EDGE sampling of the frame (offscreen_pass.cpp:361-375) into a device
        RGBA8 buffer of width x height."""
        self._check(lib.vr_present_frame(self._h, int(frame_index), dst_ptr, int(width), int(height), int(pitch_bytes), stream))

    def write_image(self, path: str):
        """The display image as binary PPM (or 4-channel PAM if the name ends in .pam)."""
        self._check(lib.vr_write_image(self._h, str(path).encode()))

    def bench_fetch_only(self, on: bool):
        """Measurement aid (include/vr.h vr_bench_fetch_only): frames run the fetch-only build of the march kernel."""
        self._check(lib.vr_bench_fetch_only(self._h, int(bool(on))))

    def download_async(self, host_ptr: int, stream: int | None = None, frame_index: int = 0):
        self._check(lib.vr_download_frame_async(self._h, frame_index, host_ptr, stream))

    def download_frame(self, frame_index: int, stream: int | None = None) -> np.ndarray:
        """The whole RGBA8 target of a frame in flight (vr_download_frame_async on `stream`, then a stream sync)."""
        import torch
        w, h = self.size
        out = np.empty((h, w, 4), np.uint8)
        self._check(lib.vr_download_frame_async(self._h, int(frame_index), _ptr(out), stream))
        torch.cuda.synchronize()
        return out

    def download_dirty_async(self, host_ptr: int, stream: int | None = None, frame_index: int = 0) -> int:
        """vr_download_frame_dirty_async: bring a host RGBA8 buffer up to date with the frame, copying only the
        rectangle that can have changed since the last delivery into that buffer.  Returns the bytes copied."""
        n = C.c_size_t()
        self._check(lib.vr_download_frame_dirty_async(self._h, frame_index, host_ptr, stream, C.byref(n)))
        return n.value

    def forget_host_buffer(self, host_ptr: int | None = None):
        """vr_forget_host_buffer: the dirty-rectangle path must not assume anything about that buffer (None: any)."""
        self._check(lib.vr_forget_host_buffer(self._h, host_ptr))

    def download_raw(self) -> np.ndarray:
        w, h = self.size
        out = np.empty((h, w, 4), np.float32)
        self._check(lib.vr_download_raw(self._h, _ptr(out)))
        return out

    def download_volume(self) -> np.ndarray:
        dt, dims, mn, mx = self.volume_info()
        out = np.empty((dims[2], dims[1], dims[0]), _NP[dt])
        self._check(lib.vr_download_volume(self._h, _ptr(out)))
        return out

    def volume_info(self):
        dt, dims, mn, mx = C.c_int(), (C.c_uint32 * 3)(), C.c_float(), C.c_float()
        self._check(lib.vr_get_volume_info(self._h, C.byref(dt), dims, C.byref(mn), C.byref(mx)))
        return dt.value, tuple(dims), mn.value, mx.value

    def stats(self) -> Stats:
        s = Stats()
        self._check(lib.vr_get_stats(self._h, C.byref(s)))
        return s

    def occupancy_info(self):
        cells, cv, empty = (C.c_uint32 * 3)(), C.c_uint32(), C.c_uint64()
        self._check(lib.vr_get_occupancy_info(self._h, cells, C.byref(cv), C.byref(empty)))
        return tuple(cells), cv.value, empty.value

    # ---- multi-GPU partitioning ------------------------------------------------------------
    def set_partition_sort_first(self, rank: int, nranks: int, tile_px: int = 64):
        self._check(lib.vr_set_partition_sort_first(self._h, rank, nranks, tile_px))

    def set_partition_sort_last(self, full_dims, own_org, own_dims):
        f, o, d = (np.asarray(x, np.uint32) for x in (full_dims, own_org, own_dims))
        self._check(lib.vr_set_partition_sort_last(self._h, _ptr(f), _ptr(o), _ptr(d)))

    def set_owned_box(self, org, dims):
        """vr_set_owned_box: the part of the held sub-volume this rank samples in the coming frames (balanced planes)."""
        o, d = (np.asarray(x, np.uint32) for x in (org, dims))
        self._check(lib.vr_set_owned_box(self._h, _ptr(o), _ptr(d)))

    def synth_subvolume(self, kind: int, seed: int, dtype: int):
        self._check(lib.vr_synth_subvolume(self._h, kind, seed & 0xFFFFFFFF, dtype))

    def clear_partition(self):
        self._check(lib.vr_clear_partition(self._h))

    def composite(self, part_ptrs, pixel_begin: int, pixel_end: int, stream: int | None = None, out_ptr: int | None = None,
                  rects=None):
        """K2.  rects: per partial (x0, y0, x1, y1) outside which it is the identity and is not read (vr_composite_rects)."""
        arr = (C.c_void_p * len(part_ptrs))(*part_ptrs)
        r = None
        if rects is not None:
            r = np.ascontiguousarray(rects, np.uint32).reshape(-1)
            assert r.size == 4 * len(part_ptrs)
        self._check(lib.vr_composite_rects(self._h, arr, _ptr(r) if r is not None else None, len(part_ptrs), pixel_begin, pixel_end,
                                           out_ptr, stream))

    def composite_region(self, part_ptrs, rects, region, band_index: int, band_count: int, stream: int | None = None, out_ptr: int | None = None):
        """vr_composite_region: K2 over the pixel region (x0, y0, x1, y1), the bands of 8 rows with index % band_count == band_index."""
        arr = (C.c_void_p * len(part_ptrs))(*part_ptrs)
        r = np.ascontiguousarray(rects, np.uint32).reshape(-1)
        assert r.size == 4 * len(part_ptrs)
        g = np.ascontiguousarray(region, np.uint32)
        self._check(lib.vr_composite_region(self._h, arr, _ptr(r), len(part_ptrs), _ptr(g), band_index, band_count, out_ptr, stream))

    def partition_screen_rect(self, full_dims, org, dims):
        """vr_partition_screen_rect: pixels (x0, y0, x1, y1) the rays through that sub-box can cover under the current camera."""
        f, o, d = (np.asarray(x, np.uint32) for x in (full_dims, org, dims))
        out = np.zeros(4, np.uint32)
        self._check(lib.vr_partition_screen_rect(self._h, _ptr(f), _ptr(o), _ptr(d), _ptr(out)))
        return tuple(int(x) for x in out)

    def ipc_export(self, which: int = 0) -> bytes:
        buf = (C.c_ubyte * 64)()
        self._check(lib.vr_ipc_export(self._h, which, buf))
        return bytes(buf)

    def ipc_open(self, handle: bytes) -> int:
        buf = (C.c_ubyte * 64)(*handle)
        p = C.c_void_p()
        self._check(lib.vr_ipc_open(self._h, buf, C.byref(p)))
        return p.value

    def ipc_close(self, ptr: int):
        self._check(lib.vr_ipc_close(self._h, ptr))

    def set_image_target(self, ptr: int | None, frame_index: int = 0):
        self._check(lib.vr_set_image_target_frame(self._h, frame_index, ptr))

    # ---- stream-ordered hand-off between ranks (include/vr.h: vr_stream_signal / vr_stream_wait) ----
    def flags(self) -> int:
        """Device pointer of this context's block of 256 uint32 counters (ipc_export(2) shares it)."""
        p = C.c_void_p()
        self._check(lib.vr_get_flags(self._h, C.byref(p)))
        return p.value

    def stream_signal(self, flag_ptrs, value: int, stream: int | None = None):
        arr = (C.c_void_p * len(flag_ptrs))(*flag_ptrs)
        self._check(lib.vr_stream_signal(self._h, arr, len(flag_ptrs), value & 0xFFFFFFFF, stream))

    def stream_wait(self, flags_ptr: int, n: int, value: int, stream: int | None = None, timeout_ms: int = 10000):
        self._check(lib.vr_stream_wait(self._h, flags_ptr, n, value & 0xFFFFFFFF, timeout_ms, stream))


class PassGroup:
    """vr_group (include/vr.h, csrc/group.cu): one process driving several GPUs -- the pass the application constructs
    at src/rendering/vulkan_context.cpp:51, spread over `devices` (a device may be named more than once)."""

    def __init__(self, devices, partition: int, width: int, height: int, slots: int = 2):
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = lib.vr_create_group(len(devices), devs, partition, width, height, slots, C.byref(h))
        if rc != _lib.VR_OK:
            raise VrError(rc, lib.vr_group_last_error(None).decode())
        self._h, self.size = h, (width, height)

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.vr_group_destroy(h)

    __del__ = close

    def _check(self, rc: int):
        if rc != _lib.VR_OK:
            raise VrError(rc, lib.vr_group_last_error(self._h).decode())

    def __len__(self):
        return lib.vr_group_size(self._h)

    def set_uniforms(self, view, proj, camera_position):
        v, p, c = (np.ascontiguousarray(x, np.float32) for x in (view, proj, camera_position))
        self._check(lib.vr_group_set_camera(self._h, _ptr(v), _ptr(p), _ptr(c)))

    def set_camera(self, camera: Camera):
        self.set_uniforms(camera.get_view(), projection(*self.size), camera.get_position())

    def set_rank_camera(self, rank: int, camera: Camera):
        """replicas: the camera of one rank's frame."""
        ctx = lib.vr_group_context(self._h, rank)
        if lib.vr_set_camera_orbit(ctx, camera._h) != _lib.VR_OK:
            raise VrError(1, lib.vr_last_error(ctx).decode())

    def transfer_function_changed(self, data):
        d = np.ascontiguousarray(data, np.uint32)
        self._check(lib.vr_group_set_transfer_function(self._h, _ptr(d), len(d)))

    def set_balance(self, fraction: float):
        """vr_group_set_balance: sort-last split planes may follow the camera by this fraction of the edge (next volume)."""
        self._check(lib.vr_group_set_balance(self._h, float(fraction)))

    def slicing_changed(self, min_slice, max_slice):
        lo, hi = np.asarray(min_slice, np.float32), np.asarray(max_slice, np.float32)
        self._check(lib.vr_group_set_slicing(self._h, _ptr(lo), _ptr(hi)))

    def set_options(self, **kw):
        o = Options()
        lib.vr_get_options(lib.vr_group_context(self._h, 0), C.byref(o))
        for k, v in kw.items():
            setattr(o, k, v)
        self._check(lib.vr_group_set_options(self._h, C.byref(o)))

    def volume_dataset_changed(self, dataset: "Dataset"):
        a = np.ascontiguousarray(dataset.data if dataset.data.dtype in _CODE else dataset.data.astype(np.float32))
        mn = C.c_float(dataset.min) if dataset.min is not None else None
        mx = C.c_float(dataset.max) if dataset.max is not None else None
        nz, ny, nx = a.shape
        self._check(lib.vr_group_set_volume(self._h, _ptr(a), _CODE[a.dtype], nx, ny, nz, C.byref(mn) if mn is not None else None,
                                            C.byref(mx) if mx is not None else None))

    def synth_volume(self, kind: int, seed: int, dtype: int, n):
        nx, ny, nz = (n, n, n) if isinstance(n, int) else n
        self._check(lib.vr_group_synth_volume(self._h, kind, seed & 0xFFFFFFFF, dtype, nx, ny, nz))

    def record(self):
        self._check(lib.vr_group_render(self._h))

    def sync(self):
        self._check(lib.vr_group_sync(self._h))

    def download(self, rank: int = 0) -> np.ndarray:
        w, h = self.size
        out = np.empty((h, w, 4), np.uint8)
        self._check(lib.vr_group_download(self._h, rank, _ptr(out)))
        return out


def read_nrrd(path: str) -> np.ndarray:
    """The native NRRD reader alone (host only): array [nz, ny, nx] as it would be uploaded."""
    dt, dims, data = C.c_int(), (C.c_uint32 * 3)(), C.c_void_p()
    err = C.create_string_buffer(512)
    rc = lib.vr_nrrd_read_host(str(path).encode(), C.byref(dt), dims, C.byref(data), err, len(err))
    if rc != _lib.VR_OK:
        raise VrError(rc, err.value.decode())
    try:
        n = dims[0] * dims[1] * dims[2]
        np_dt = np.dtype(_NP[dt.value])
        buf = (C.c_char * (n * np_dt.itemsize)).from_address(data.value)
        return np.frombuffer(buf, np_dt).reshape(dims[2], dims[1], dims[0]).copy()
    finally:
        lib.vr_free_host(data)


def read_csv(paths) -> "Dataset":
    """CsvFileParser(paths).parse() of the reference (src/data/csv_file_parser.cpp:14-50), host only: one CSV file
    per z slice -> Dataset(float32 [nz, ny, nx], min, max) with the reference's min <= 0 <= max."""
    arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
    dims, data, mn, mx = (C.c_uint32 * 3)(), C.c_void_p(), C.c_float(), C.c_float()
    err = C.create_string_buffer(512)
    rc = lib.vr_csv_read_host(arr, len(paths), dims, C.byref(data), C.byref(mn), C.byref(mx), err, len(err))
    if rc != _lib.VR_OK:
        raise VrError(rc, err.value.decode())
    try:
        n = dims[0] * dims[1] * dims[2]
        buf = (C.c_float * n).from_address(data.value) if n else []
        a = np.array(buf, np.float32).reshape(dims[2], dims[1], dims[0])
        return Dataset(a, mn.value, mx.value)
    finally:
        lib.vr_free_host(data)


def sort_first_owner(tile_x, tile_y, nranks: int):
    """vr_sort_first_owner for integer arrays: the rank that renders each partition tile."""
    m = next(m for m in range(3, 3 + 2 * max(nranks, 1) + 3) if np.gcd(m, nranks) == 1)
    assert lib.vr_sort_first_owner(1, 1, nranks) == (1 + m) % nranks
    return (np.asarray(tile_x) + np.asarray(tile_y) * m) % nranks


def partition_box(grid, full_dims, index: int):
    g, f = np.asarray(grid, np.uint32), np.asarray(full_dims, np.uint32)
    org, dims = np.zeros(3, np.uint32), np.zeros(3, np.uint32)
    rc = lib.vr_partition_box(_ptr(g), _ptr(f), index, _ptr(org), _ptr(dims))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_partition_box: bad arguments")
    return tuple(int(x) for x in org), tuple(int(x) for x in dims)


def partition_order(grid, full_dims, camera_position, planes=None) -> np.ndarray:
    g, f = np.asarray(grid, np.uint32), np.asarray(full_dims, np.uint32)
    c = np.asarray(camera_position, np.float32)
    out = np.zeros(int(g[0]) * int(g[1]) * int(g[2]), np.uint32)
    pl = None if planes is None else np.ascontiguousarray(planes, np.uint32)
    rc = lib.vr_partition_order_planes(_ptr(g), _ptr(f), None if pl is None else _ptr(pl), _ptr(c), _ptr(out))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_partition_order: bad arguments")
    return out


def partition_stored_box(grid, full_dims, index: int, margin):
    """The regular sub-box grown by `margin` voxels on its interior faces: what a rank holds so that the split planes
    can follow the camera (vr_partition_balance)."""
    g, f, m = (np.asarray(x, np.uint32) for x in (grid, full_dims, margin))
    org, dims = np.zeros(3, np.uint32), np.zeros(3, np.uint32)
    rc = lib.vr_partition_stored_box(_ptr(g), _ptr(f), index, _ptr(m), _ptr(org), _ptr(dims))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_partition_stored_box: bad arguments (margin too large for the grid?)")
    return tuple(int(x) for x in org), tuple(int(x) for x in dims)


def partition_balance(grid, full_dims, camera_position, margin) -> np.ndarray:
    """Split planes (voxels; x planes, then y, then z) that even out the samples per sub-box for this camera."""
    g, f, m = (np.asarray(x, np.uint32) for x in (grid, full_dims, margin))
    c = np.asarray(camera_position, np.float32)
    out = np.zeros(max(int(g[0]) + int(g[1]) + int(g[2]) - 3, 1), np.uint32)
    rc = lib.vr_partition_balance(_ptr(g), _ptr(f), _ptr(c), _ptr(m), _ptr(out))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_partition_balance: bad arguments")
    return out[: int(g[0]) + int(g[1]) + int(g[2]) - 3]


def partition_box_planes(grid, full_dims, planes, index: int):
    g, f = np.asarray(grid, np.uint32), np.asarray(full_dims, np.uint32)
    pl = None if planes is None else np.ascontiguousarray(planes, np.uint32)
    org, dims = np.zeros(3, np.uint32), np.zeros(3, np.uint32)
    rc = lib.vr_partition_box_planes(_ptr(g), _ptr(f), None if pl is None else _ptr(pl), index, _ptr(org), _ptr(dims))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_partition_box_planes: bad arguments")
    return tuple(int(x) for x in org), tuple(int(x) for x in dims)


def nrrd_spacing(path: str) -> np.ndarray:
    """vr_nrrd_read_spacing: per-axis sample distances a NRRD header states (NaN where none)."""
    out = np.zeros(3, np.float64)
    err = C.create_string_buffer(512)
    rc = lib.vr_nrrd_read_spacing(path.encode(), _ptr(out), err, len(err))
    if rc != _lib.VR_OK:
        raise VrError(rc, err.value.decode())
    return out


def sort_last_region(rects, width: int, height: int, first_into_slot: bool, slot_region: np.ndarray):
    """vr_sort_last_region: the region a sort-last frame has to merge into an image slot; slot_region (uint32[4]) is updated."""
    r = np.ascontiguousarray(rects, np.uint32).reshape(-1)
    assert slot_region.dtype == np.uint32 and slot_region.size == 4 and slot_region.flags.c_contiguous
    out = np.zeros(4, np.uint32)
    lib.vr_sort_last_region(_ptr(r), r.size // 4, width, height, int(first_into_slot), _ptr(slot_region), _ptr(out))
    return tuple(int(x) for x in out)


def minmax_static(device: int, dev_values: int, dtype: int, count: int, dev_min_max: int, stream: int | None = None):
    """vr_minmax_static: {min, max} of `count` values in device memory, written to two floats in device memory at
    dev_min_max; allocates nothing (its scratch is a static device array shared by all calls on the device: do not
    overlap calls).  Asynchronous on `stream`."""
    err = C.create_string_buffer(256)
    rc = lib.vr_minmax_static(device, C.c_void_p(dev_values), dtype, count, C.c_void_p(dev_min_max), stream, err, len(err))
    if rc != _lib.VR_OK:
        raise VrError(rc, err.value.decode())


def bench_tex_peak(device: int, dtype: int, hw_filter=True) -> float:
    """hw_filter: 0 point, 1 trilinear tex3D (LAYOUT_ARRAY3D's fetch), 2 paired bilinear (LAYOUT_PAIRED's fetch)."""
    r = C.c_double()
    rc = lib.vr_bench_tex_peak(device, dtype, int(hw_filter), C.byref(r))
    if rc != _lib.VR_OK:
        raise VrError(rc, "vr_bench_tex_peak failed")
    return r.value
