"""ctypes binding of the CPU oracles (oracle/vr_oracle.h).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.

Two builds of the same C API (see oracle/Makefile):
  Oracle("port")       oracle/libvr_oracle_port.so       plain-C restatement
  Oracle("reference")  oracle/_ref/libvr_oracle_ref.so   reference sources compiled as CPU code
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libvr_oracle_port.so")
REF_SO = os.path.join(HERE, "_ref", "libvr_oracle_ref.so")
REFERENCE_ROOT = "/root/reference"

U8, U16, F32 = 0, 1, 2
_NP_DTYPE = {U8: np.uint8, U16: np.uint16, F32: np.float32}
_DTYPE_OF = {np.dtype(np.uint8): U8, np.dtype(np.uint16): U16, np.dtype(np.float32): F32}

DEFAULT_RAY_DIST = 1.8   # res/shaders/volume.frag:29
DEFAULT_STEP = 0.005     # res/shaders/volume.frag:30


class Frame(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("view", C.c_float * 16), ("proj", C.c_float * 16),
        ("camera_position", C.c_float * 3),
        ("min_density", C.c_float), ("max_density", C.c_float),
        ("min_slice", C.c_float * 3), ("max_slice", C.c_float * 3),
        ("ray_dist", C.c_float), ("step_size", C.c_float),
    ]


class Volume(C.Structure):
    _fields_ = [("data", C.c_void_p), ("dtype", C.c_int),
                ("nx", C.c_uint32), ("ny", C.c_uint32), ("nz", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("covered_pixels", C.c_uint64), ("n_ref", C.c_uint64), ("n_inslice", C.c_uint64)]


def build(kind: str = "port") -> str:
    """Compile the oracle if needed; returns the .so path.  'reference' needs /root/reference."""
    if kind == "port":
        subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)
        return PORT_SO
    if os.path.isdir(REFERENCE_ROOT):
        subprocess.run(["make", "-s", "-j8", "-C", HERE, "ref"], check=True)
    if not os.path.exists(REF_SO):
        raise FileNotFoundError(f"{REF_SO} missing and {REFERENCE_ROOT} absent: cannot build the reference oracle")
    return REF_SO


class Oracle:
    def __init__(self, kind: str = "port"):
        path = PORT_SO if kind == "port" else REF_SO
        if not os.path.exists(path):
            path = build(kind)
        self.lib = lib = C.CDLL(path)
        self.kind = kind
        lib.vro_kind.restype = C.c_char_p
        assert lib.vro_kind().decode() == kind
        lib.vro_render.restype = C.c_int
        lib.vro_render.argtypes = [C.POINTER(Frame), C.POINTER(Volume), C.c_void_p, C.c_uint32,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Stats), C.c_int]
        lib.vro_render_rows.restype = C.c_int
        lib.vro_render_rows.argtypes = lib.vro_render.argtypes + [C.c_uint32, C.c_uint32]
        lib.vro_camera_new.restype = C.c_void_p
        lib.vro_camera_free.argtypes = [C.c_void_p]
        lib.vro_camera_rotate.argtypes = [C.c_void_p, C.c_float, C.c_float]
        lib.vro_camera_zoom.argtypes = [C.c_void_p, C.c_float]
        lib.vro_camera_get_position.argtypes = [C.c_void_p, C.c_void_p]
        lib.vro_camera_get_view.argtypes = [C.c_void_p, C.c_void_p]
        lib.vro_projection.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p]
        lib.vro_frame_from_camera.argtypes = [C.POINTER(Frame), C.c_void_p]
        lib.vro_gradient_discretize.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.vro_load_nrrd.restype = C.c_int
        lib.vro_load_nrrd.argtypes = [C.c_char_p, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                      C.POINTER(C.c_void_p)]
        lib.vro_free.argtypes = [C.c_void_p]
        lib.vro_synth.restype = C.c_int
        lib.vro_synth.argtypes = [C.c_int, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]

    # -- camera ----------------------------------------------------------------
    def camera(self) -> "Camera":
        return Camera(self)

    def projection(self, width: int, height: int) -> np.ndarray:
        out = np.zeros(16, np.float32)
        self.lib.vro_projection(width, height, out.ctypes.data)
        return out

    def frame(self, width, height, cam: "Camera", min_density, max_density,
              min_slice=(0.0, 0.0, 0.0), max_slice=(1.0, 1.0, 1.0),
              ray_dist=DEFAULT_RAY_DIST, step_size=DEFAULT_STEP) -> Frame:
        f = Frame()
        f.width, f.height = width, height
        self.lib.vro_frame_from_camera(C.byref(f), cam.h)
        f.min_density, f.max_density = min_density, max_density
        f.min_slice[:] = min_slice
        f.max_slice[:] = max_slice
        f.ray_dist, f.step_size = ray_dist, step_size
        return f

    # -- transfer function -----------------------------------------------------
    def discretize(self, color_markers, alpha_markers, count=256) -> np.ndarray:
        """color_markers: [(loc, r, g, b)], alpha_markers: [(loc, a)] -> uint32[count]"""
        cm = np.ascontiguousarray(np.array(color_markers, np.float32).reshape(-1, 4))
        am = np.ascontiguousarray(np.array(alpha_markers, np.float32).reshape(-1, 2))
        out = np.zeros(count, np.uint32)
        self.lib.vro_gradient_discretize(cm.ctypes.data, len(cm), am.ctypes.data, len(am), count, out.ctypes.data)
        return out

    # -- render ----------------------------------------------------------------
    def render(self, frame: Frame, volume: np.ndarray, lut: np.ndarray, want_raw=False, want_iters=False,
               nthreads=0, row_start=0, row_stride=1):
        """volume: ndarray [nz, ny, nx] (x fastest) of uint8/uint16/float32.
        Returns dict(rgba8=[H,W,4] u8, raw=[H,W,4] f32 | None, iters=[H,W] u16 | None, stats=Stats)"""
        assert volume.ndim == 3 and volume.flags.c_contiguous
        vol = Volume(volume.ctypes.data, _DTYPE_OF[volume.dtype], volume.shape[2], volume.shape[1], volume.shape[0])
        lut = np.ascontiguousarray(lut, np.uint32)
        H, W = frame.height, frame.width
        rgba8 = np.zeros((H, W, 4), np.uint8)
        raw = np.zeros((H, W, 4), np.float32) if want_raw else None
        iters = np.zeros((H, W), np.uint16) if want_iters else None
        st = Stats()
        rc = self.lib.vro_render_rows(C.byref(frame), C.byref(vol), lut.ctypes.data, len(lut), rgba8.ctypes.data,
                                      raw.ctypes.data if want_raw else None,
                                      iters.ctypes.data if want_iters else None, C.byref(st), nthreads,
                                      row_start, row_stride)
        if rc != 0:
            raise RuntimeError(f"vro_render failed: {rc}")
        return dict(rgba8=rgba8, raw=raw, iters=iters, stats=st)

    # -- pre-integrated classification (port only: no reference counterpart) -----
    def preint_table(self, lut: np.ndarray, substeps: int) -> np.ndarray:
        lut = np.ascontiguousarray(lut, np.uint32)
        n = len(lut)
        tab = np.zeros((n, n, 4), np.float32)
        self.lib.vro_preint_table.restype = C.c_int
        self.lib.vro_preint_table.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        rc = self.lib.vro_preint_table(lut.ctypes.data, n, substeps, tab.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"vro_preint_table failed: {rc}")
        return tab

    def render_preint(self, frame: Frame, volume: np.ndarray, lut: np.ndarray, substeps: int, nthreads=0):
        assert volume.ndim == 3 and volume.flags.c_contiguous
        vol = Volume(volume.ctypes.data, _DTYPE_OF[volume.dtype], volume.shape[2], volume.shape[1], volume.shape[0])
        lut = np.ascontiguousarray(lut, np.uint32)
        H, W = frame.height, frame.width
        rgba8 = np.zeros((H, W, 4), np.uint8)
        raw = np.zeros((H, W, 4), np.float32)
        self.lib.vro_render_preint.restype = C.c_int
        self.lib.vro_render_preint.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
        rc = self.lib.vro_render_preint(C.byref(frame), C.byref(vol), lut.ctypes.data, len(lut), substeps, rgba8.ctypes.data, raw.ctypes.data, nthreads)
        if rc != 0:
            raise RuntimeError(f"vro_render_preint failed: {rc}")
        return dict(rgba8=rgba8, raw=raw)

    # -- non-cubic voxels (port only: the reference ignores a file's spacings) -----
    def render_spacing(self, frame: Frame, extent, volume: np.ndarray, lut: np.ndarray, nthreads=0):
        assert volume.ndim == 3 and volume.flags.c_contiguous
        vol = Volume(volume.ctypes.data, _DTYPE_OF[volume.dtype], volume.shape[2], volume.shape[1], volume.shape[0])
        lut = np.ascontiguousarray(lut, np.uint32)
        ext = (C.c_float * 3)(*[float(x) for x in extent])
        H, W = frame.height, frame.width
        rgba8 = np.zeros((H, W, 4), np.uint8)
        raw = np.zeros((H, W, 4), np.float32)
        st = Stats()
        self.lib.vro_render_spacing.restype = C.c_int
        self.lib.vro_render_spacing.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        rc = self.lib.vro_render_spacing(C.byref(frame), ext, C.byref(vol), lut.ctypes.data, len(lut), rgba8.ctypes.data, raw.ctypes.data,
                                         C.byref(st), nthreads)
        if rc != 0:
            raise RuntimeError(f"vro_render_spacing failed: {rc}")
        return dict(rgba8=rgba8, raw=raw, stats=st)

    # -- data ------------------------------------------------------------------
    def load_nrrd(self, path: str):
        dims = (C.c_uint32 * 3)()
        mn, mx = C.c_float(), C.c_float()
        ptr = C.c_void_p()
        rc = self.lib.vro_load_nrrd(path.encode(), dims, C.byref(mn), C.byref(mx), C.byref(ptr))
        if rc != 0:
            raise RuntimeError(f"vro_load_nrrd failed: {rc}")
        n = dims[0] * dims[1] * dims[2]
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(n,)).copy()
        self.lib.vro_free(ptr)
        return arr.reshape(dims[2], dims[1], dims[0]), mn.value, mx.value

    def present_bilinear(self, src: np.ndarray, width: int, height: int) -> np.ndarray:
        """The viewport sampler (LINEAR, CLAMP_TO_EDGE) applied to an RGBA8 frame [h, w, 4] -> [height, width, 4]."""
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.empty((height, width, 4), np.uint8)
        self.lib.vro_present_bilinear.restype = C.c_int
        self.lib.vro_present_bilinear.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32]
        rc = self.lib.vro_present_bilinear(src.ctypes.data, src.shape[1], src.shape[0], dst.ctypes.data, width, height)
        if rc != 0:
            raise RuntimeError(f"vro_present_bilinear failed: {rc}")
        return dst

    def load_csv(self, paths):
        """CsvFileParser(paths).parse() of the reference build; raises RuntimeError(e.what()) like the importer's popup."""
        arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
        dims = (C.c_uint32 * 3)()
        mn, mx = C.c_float(), C.c_float()
        ptr = C.c_void_p()
        msg = C.create_string_buffer(256)
        self.lib.vro_load_csv.restype = C.c_int
        self.lib.vro_load_csv.argtypes = [C.POINTER(C.c_char_p), C.c_uint32, C.c_void_p, C.POINTER(C.c_float),
                                          C.POINTER(C.c_float), C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
        rc = self.lib.vro_load_csv(arr, len(paths), dims, C.byref(mn), C.byref(mx), C.byref(ptr), msg, len(msg))
        if rc != 0:
            raise RuntimeError(msg.value.decode() or f"vro_load_csv failed: {rc}")
        n = dims[0] * dims[1] * dims[2]
        a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(max(n, 1),))[:n].copy()
        self.lib.vro_free(ptr)
        return a.reshape(dims[2], dims[1], dims[0]), mn.value, mx.value

    def raster_cube(self, frame: Frame):
        """oracle/raster.c: the reference's cube drawn by a software rasteriser under Vulkan's rules.
        Returns (mask [H,W] u8: 0 or 1 + face, tex [H,W,3] f32: interpolated in_tex_coords, depth [H,W] f32)."""
        h, w = frame.height, frame.width
        mask, tex, depth = np.zeros((h, w), np.uint8), np.zeros((h, w, 3), np.float32), np.zeros((h, w), np.float32)
        self.lib.vro_raster_cube.restype = C.c_int
        self.lib.vro_raster_cube.argtypes = [C.POINTER(Frame), C.c_void_p, C.c_void_p, C.c_void_p]
        if self.lib.vro_raster_cube(C.byref(frame), mask.ctypes.data, tex.ctypes.data, depth.ctypes.data) < 0:
            raise RuntimeError("vro_raster_cube failed")
        return mask, tex, depth

    def entry_map(self, frame: Frame):
        """vro_pixel_entry (the analytic coverage + entry the oracle and the kernel use) for every pixel."""
        h, w = frame.height, frame.width
        mask, tex = np.zeros((h, w), np.uint8), np.zeros((h, w, 3), np.float32)
        self.lib.vro_entry_map.restype = C.c_int
        self.lib.vro_entry_map.argtypes = [C.POINTER(Frame), C.c_void_p, C.c_void_p]
        if self.lib.vro_entry_map(C.byref(frame), mask.ctypes.data, tex.ctypes.data) < 0:
            raise RuntimeError("vro_entry_map failed")
        return mask, tex

    def synth(self, kind: int, seed: int, dtype, n, nthreads=0) -> np.ndarray:
        nx, ny, nz = (n, n, n) if isinstance(n, int) else n
        code = dtype if isinstance(dtype, int) else _DTYPE_OF[np.dtype(dtype)]
        out = np.empty((nz, ny, nx), _NP_DTYPE[code])
        rc = self.lib.vro_synth(kind, seed & 0xFFFFFFFF, code, nx, ny, nz, out.ctypes.data, nthreads)
        if rc != 0:
            raise RuntimeError(f"vro_synth failed: {rc}")
        return out


class Camera:
    def __init__(self, oracle: Oracle):
        self.o = oracle
        self.h = C.c_void_p(oracle.lib.vro_camera_new())

    def __del__(self):
        try:
            self.o.lib.vro_camera_free(self.h)
        except Exception:
            pass

    def rotate(self, dx, dy):
        self.o.lib.vro_camera_rotate(self.h, dx, dy)
        return self

    def zoom(self, delta):
        self.o.lib.vro_camera_zoom(self.h, delta)
        return self

    @property
    def position(self) -> np.ndarray:
        out = np.zeros(3, np.float32)
        self.o.lib.vro_camera_get_position(self.h, out.ctypes.data)
        return out

    @property
    def view(self) -> np.ndarray:
        out = np.zeros(16, np.float32)
        self.o.lib.vro_camera_get_view(self.h, out.ctypes.data)
        return out


# Transfer functions of SURVEY.md section 8(d)
TF_0 = np.array([0xFFFFFFFF], np.uint32)  # start-up LUT, offscreen_pass.cpp:119
TF_A_MARKERS = ([(0.0, 0, 0, 0), (1.0, 1, 1, 1)], [(0.0, 1.0), (1.0, 1.0)])  # gradient.cpp:64-70
TF_B_MARKERS = ([(0.0, 0, 0, 0.5), (0.5, 1, 0.5, 0), (1.0, 1, 1, 1)], [(0.0, 0.0), (0.3, 0.0), (1.0, 0.08)])
TF_C_MARKERS = ([(0.0, 0, 0, 0.5), (0.5, 1, 0.5, 0), (1.0, 1, 1, 1)], [(0.0, 0.0), (0.05, 0.0), (1.0, 0.6)])
