"""ctypes declarations for libvr_b200.so (include/vr.h).

The library is the product; this module only declares its entry points.  There is no
fallback of any kind: if the shared library is missing, importing the package raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VR_LIB_PATH") or os.path.join(HERE, "libvr_b200.so")  # VR_LIB_PATH: an alternative build (experiments)

VR_OK = 0
U8, U16, F32 = 0, 1, 2
SAMPLER_HW, SAMPLER_EXACT = 0, 1
SYNTH_MARSCHNER_LOBB, SYNTH_FBM, SYNTH_BLOBS, SYNTH_BLOBS_FBM = 0, 1, 2, 3
LAYOUT_AUTO, LAYOUT_ARRAY3D, LAYOUT_PAIRED = 0, 1, 2
PARTITION_REPLICAS, PARTITION_SORT_FIRST, PARTITION_SORT_LAST = 0, 1, 2


class Options(C.Structure):
    _fields_ = [("sampler", C.c_int), ("empty_skip", C.c_int), ("early_term", C.c_int),
                ("early_term_eps", C.c_float), ("collect_stats", C.c_int), ("write_raw", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("covered_pixels", C.c_uint64), ("n_ref", C.c_uint64), ("n_inslice", C.c_uint64),
                ("n_fetch", C.c_uint64)]


class SortLastWiring(C.Structure):
    """vr_sort_last_wiring (include/vr.h)"""
    _fields_ = [("rank", C.c_uint32), ("world", C.c_uint32), ("display", C.c_uint32), ("slots", C.c_uint32),
                ("grid", C.c_uint32 * 3), ("full", C.c_uint32 * 3), ("margin", C.c_uint32 * 3),
                ("flags", C.POINTER(C.c_void_p)), ("raw", C.POINTER(C.c_void_p)), ("image", C.POINTER(C.c_void_p)),
                ("pixel_begin", C.c_uint64), ("pixel_end", C.c_uint64)]


class FrameInfo(C.Structure):
    _fields_ = [("layout", C.c_int32), ("pair_axis", C.c_int32), ("empty_skip_active", C.c_int32), ("exact_sampler", C.c_int32),
                ("bbox", C.c_uint32 * 4), ("march_blocks", C.c_uint32), ("clear_blocks", C.c_uint32), ("target_bytes", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/vr.h declares
_P = C.c_void_p
_F = C.POINTER(C.c_float)
_U = C.POINTER(C.c_uint32)
SIGNATURES = {
    "vr_create": (C.c_int, [C.c_int, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "vr_destroy": (C.c_int, [_P]),
    "vr_last_error": (C.c_char_p, [_P]),
    "vr_abi_version": (C.c_int, []),
    "vr_kernel_version": (C.c_char_p, []),
    "vr_render": (C.c_int, [_P, _P]),
    "vr_render_frame": (C.c_int, [_P, C.c_uint32, _P]),
    "vr_get_last_frame_info": (C.c_int, [_P, C.POINTER(FrameInfo)]),
    "vr_get_tile_mask": (C.c_int, [_P, C.c_uint32, C.POINTER(C.c_uint32), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                   C.POINTER(C.c_uint32)]),
    "vr_get_image_frame": (C.c_int, [_P, C.c_uint32, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "vr_download_frame_async": (C.c_int, [_P, C.c_uint32, _P, _P]),
    "vr_download_frame_dirty_async": (C.c_int, [_P, C.c_uint32, _P, _P, C.POINTER(C.c_size_t)]),
    "vr_forget_host_buffer": (C.c_int, [_P, _P]),
    "vr_sync": (C.c_int, [_P]),
    "vr_set_camera": (C.c_int, [_P, _P, _P, _P]),
    "vr_resize": (C.c_int, [_P, C.c_uint32, C.c_uint32]),
    "vr_set_volume": (C.c_int, [_P, _P, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _F, _F]),
    "vr_set_volume_device": (C.c_int, [_P, _P, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _F, _F]),
    "vr_load_nrrd": (C.c_int, [_P, C.c_char_p]),
    "vr_nrrd_read_host": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), _U, C.POINTER(_P), C.c_char_p, C.c_size_t]),
    "vr_nrrd_read_spacing": (C.c_int, [C.c_char_p, _P, C.c_char_p, C.c_size_t]),
    "vr_free_host": (None, [_P]),
    "vr_sort_first_owner": (C.c_uint32, [C.c_uint32, C.c_uint32, C.c_uint32]),
    "vr_present_frame": (C.c_int, [_P, C.c_uint32, _P, C.c_uint32, C.c_uint32, C.c_size_t, _P]),
    "vr_load_csv": (C.c_int, [_P, C.POINTER(C.c_char_p), C.c_uint32]),
    "vr_csv_read_host": (C.c_int, [C.POINTER(C.c_char_p), C.c_uint32, _U, C.POINTER(_P), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                   C.c_char_p, C.c_size_t]),
    "vr_get_volume_info": (C.c_int, [_P, C.POINTER(C.c_int), _U, _F, _F]),
    "vr_set_density_range": (C.c_int, [_P, C.c_float, C.c_float]),
    "vr_set_slicing": (C.c_int, [_P, _P, _P]),
    "vr_set_transfer_function": (C.c_int, [_P, _P, C.c_uint32]),
    "vr_set_transfer_function_device": (C.c_int, [_P, _P, C.c_uint32]),
    "vr_set_gradient": (C.c_int, [_P, _P, C.c_uint32, _P, C.c_uint32, C.c_uint32]),
    "vr_get_transfer_function": (C.c_int, [_P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "vr_set_march": (C.c_int, [_P, C.c_float, C.c_float]),
    "vr_set_volume_spacing": (C.c_int, [_P, _P]),
    "vr_get_volume_spacing": (C.c_int, [_P, _P, _P]),
    "vr_set_preintegration": (C.c_int, [_P, C.c_uint32]),
    "vr_get_preintegration_table": (C.c_int, [_P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "vr_set_volume_layout": (C.c_int, [_P, C.c_int]),
    "vr_get_volume_layout": (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "vr_get_options": (C.c_int, [_P, C.POINTER(Options)]),
    "vr_set_options": (C.c_int, [_P, C.POINTER(Options)]),
    "vr_get_image": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "vr_download": (C.c_int, [_P, _P]),
    "vr_download_async": (C.c_int, [_P, _P, _P]),
    "vr_get_raw": (C.c_int, [_P, C.POINTER(_P)]),
    "vr_get_raw_frame": (C.c_int, [_P, C.c_uint32, C.POINTER(_P)]),
    "vr_download_raw": (C.c_int, [_P, _P]),
    "vr_write_image": (C.c_int, [_P, C.c_char_p]),
    "vr_get_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "vr_get_size": (C.c_int, [_P, _U, _U]),
    "vr_get_occupancy_info": (C.c_int, [_P, _U, _U, C.POINTER(C.c_uint64)]),
    "vr_camera_new": (_P, []),
    "vr_camera_free": (None, [_P]),
    "vr_camera_rotate": (None, [_P, C.c_float, C.c_float]),
    "vr_camera_zoom": (None, [_P, C.c_float]),
    "vr_camera_get_position": (None, [_P, _P]),
    "vr_camera_get_view": (None, [_P, _P]),
    "vr_projection": (None, [C.c_uint32, C.c_uint32, _P]),
    "vr_set_camera_orbit": (C.c_int, [_P, _P]),
    "vr_gradient_discretize": (C.c_int, [_P, C.c_uint32, _P, C.c_uint32, C.c_uint32, _P]),
    "vr_synth_volume": (C.c_int, [_P, C.c_int, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32]),
    "vr_download_volume": (C.c_int, [_P, _P]),
    "vr_set_partition_sort_first": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.c_uint32]),
    "vr_set_partition_sort_last": (C.c_int, [_P, _P, _P, _P]),
    "vr_synth_subvolume": (C.c_int, [_P, C.c_int, C.c_uint32, C.c_int]),
    "vr_clear_partition": (C.c_int, [_P]),
    "vr_composite": (C.c_int, [_P, _P, C.c_uint32, C.c_uint64, C.c_uint64, _P]),
    "vr_composite_to": (C.c_int, [_P, _P, C.c_uint32, C.c_uint64, C.c_uint64, _P, _P]),
    "vr_partition_screen_rect": (C.c_int, [_P, _P, _P, _P, _P]),
    "vr_composite_rects": (C.c_int, [_P, _P, _P, C.c_uint32, C.c_uint64, C.c_uint64, _P, _P]),
    "vr_get_flags": (C.c_int, [_P, C.POINTER(_P)]),
    "vr_stream_signal": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32, _P]),
    "vr_stream_wait": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32, _P]),
    "vr_sync_uses_memops": (C.c_int, []),
    "vr_set_image_target_frame": (C.c_int, [_P, C.c_uint32, _P]),
    "vr_ipc_export": (C.c_int, [_P, C.c_int, _P]),
    "vr_ipc_open": (C.c_int, [_P, _P, C.POINTER(_P)]),
    "vr_ipc_close": (C.c_int, [_P, _P]),
    "vr_set_image_target": (C.c_int, [_P, _P]),
    "vr_partition_box": (C.c_int, [_P, _P, C.c_uint32, _P, _P]),
    "vr_partition_order": (C.c_int, [_P, _P, _P, _P]),
    "vr_set_owned_box": (C.c_int, [_P, _P, _P]),
    "vr_composite_region": (C.c_int, [_P, _P, _P, C.c_uint32, _P, C.c_uint32, C.c_uint32, _P, _P]),
    "vr_sort_last_region": (None, [_P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _P, _P]),
    "vr_sort_last_frame": (C.c_int, [_P, C.POINTER(SortLastWiring), C.c_uint64, _P, _P, _P, _P]),
    "vr_minmax_static_max_count": (C.c_uint64, []),
    "vr_minmax_static": (C.c_int, [C.c_int, _P, C.c_int, C.c_uint64, _P, _P, C.c_char_p, C.c_size_t]),
    "vr_partition_stored_box": (C.c_int, [_P, _P, C.c_uint32, _P, _P, _P]),
    "vr_partition_balance": (C.c_int, [_P, _P, _P, _P, _P]),
    "vr_partition_box_planes": (C.c_int, [_P, _P, _P, C.c_uint32, _P, _P]),
    "vr_partition_order_planes": (C.c_int, [_P, _P, _P, _P, _P]),
    "vr_create_group": (C.c_int, [C.c_uint32, C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "vr_group_destroy": (C.c_int, [_P]),
    "vr_group_last_error": (C.c_char_p, [_P]),
    "vr_group_size": (C.c_uint32, [_P]),
    "vr_group_context": (_P, [_P, C.c_uint32]),
    "vr_group_set_camera": (C.c_int, [_P, _P, _P, _P]),
    "vr_group_set_transfer_function": (C.c_int, [_P, _P, C.c_uint32]),
    "vr_group_set_slicing": (C.c_int, [_P, _P, _P]),
    "vr_group_set_options": (C.c_int, [_P, C.POINTER(Options)]),
    "vr_group_set_volume": (C.c_int, [_P, _P, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _F, _F]),
    "vr_group_synth_volume": (C.c_int, [_P, C.c_int, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32]),
    "vr_group_set_balance": (C.c_int, [_P, C.c_float]),
    "vr_group_render": (C.c_int, [_P]),
    "vr_group_sync": (C.c_int, [_P]),
    "vr_group_download": (C.c_int, [_P, C.c_uint32, _P]),
    "vr_bench_tex_peak": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "vr_bench_fetch_only": (C.c_int, [_P, C.c_int]),
    "vr_debug_ray_context": (C.c_int, [_P, _P]),
}


def load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C {HERE}` (or __graft_entry__.build()). "
            "There is no CPU or PyTorch fallback for the ray-march path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    return lib
