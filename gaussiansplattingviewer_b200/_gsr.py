"""ctypes binding of csrc/libgsr.so (C ABI: include/gsr.h).  Loads the CUDA library or raises —
there is no fallback implementation."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libgsr.so")

GSR_FLAG_CLAMP_UNORM = 1
GSR_FLAG_KEYS64 = 2
GSR_FLAG_DEBUG = 4
GSR_FLAG_OFFICIAL = 8

GSR_SORT_STATIC_MAX_PAIRS = 1 << 22

GSR_DBG_VSOUT = 0
GSR_DBG_SETUP = 1
GSR_DBG_INST_KEYS = 2
GSR_DBG_INST_VALS = 3
GSR_DBG_RANGES = 4

GSR_TUNE_FRAME_FORK = 6
GSR_TUNE_TILE = 7
GSR_TUNE_FUSED_DEPTH_HIST = 9
GSR_TUNE_PDL = 10
GSR_TUNE_FUSED_RGB8 = 11
GSR_TUNE_GRAPH = 12


class Stats(C.Structure):
    _fields_ = [
        ("n_gaussians", C.c_longlong), ("n_visible", C.c_longlong), ("n_instances", C.c_longlong),
        ("capacity", C.c_longlong), ("overflow", C.c_int), ("tiles_x", C.c_int), ("tiles_y", C.c_int),
        ("ms_preprocess", C.c_float), ("ms_sort_depth", C.c_float), ("ms_duplicate", C.c_float),
        ("ms_sort_tiles", C.c_float), ("ms_ranges", C.c_float), ("ms_blend", C.c_float), ("ms_total", C.c_float),
    ]


class PlyLayout(C.Structure):
    # gsr_ply_layout: x y z | f_dc_0..2 | f_rest_0..44 | opacity | scale_0..2 | rot_0..3
    _fields_ = [("row_bytes", C.c_int32), ("big_endian", C.c_int32), ("offset", C.c_int32 * 59)]


_fp = C.POINTER(C.c_float)
_vp = C.c_void_p
_ll = C.c_longlong

# name -> (restype, argtypes); every symbol include/gsr.h declares
SIGNATURES = {
    "gsr_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "gsr_destroy": (C.c_int, [_vp]),
    "gsr_last_error": (C.c_char_p, []),
    "gsr_version": (C.c_int, []),
    "gsr_set_gaussians": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _ll, C.c_int, C.c_int]),
    "gsr_load_ply_rows": (C.c_int, [_vp, _vp, _ll, C.POINTER(PlyLayout), C.c_int, _fp, _fp, _vp]),
    "gsr_get_gaussians": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "gsr_set_resolution": (C.c_int, [_vp, C.c_int, C.c_int]),
    "gsr_set_camera": (C.c_int, [_vp, _fp, _fp, _fp, C.c_float, C.c_float, C.c_float]),
    "gsr_set_params": (C.c_int, [_vp, C.c_float, C.c_int, C.c_uint]),
    "gsr_set_capacity": (C.c_int, [_vp, _ll]),
    "gsr_render": (C.c_int, [_vp, _vp, _vp]),
    "gsr_render_timed": (C.c_int, [_vp, _vp, _vp]),
    "gsr_render_host": (C.c_int, [_vp, _vp, _vp]),
    "gsr_render_rgb8_host": (C.c_int, [_vp, _vp, _vp]),
    "gsr_render_r32f_host": (C.c_int, [_vp, _vp, _vp]),
    "gsr_render_rgb8_host_async": (C.c_int, [_vp, _vp, _vp]),
    "gsr_host_sync": (C.c_int, [_vp]),
    "gsr_capture_stereo_async": (C.c_int, [_vp, _fp, _fp, _fp, _fp, _vp, _vp, _vp, C.c_int, _vp]),
    "gsr_render_batch": (C.c_int, [_vp, _fp, _fp, C.c_int, _vp, _vp]),
    "gsr_sort_depth": (C.c_int, [_vp, _fp, _vp, C.c_int, _vp]),
    "gsr_freeze_order": (C.c_int, [_vp, _fp, _vp]),
    "gsr_thaw_order": (C.c_int, [_vp]),
    "gsr_sort_pairs": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _ll, C.c_int, C.c_int, _vp]),
    "gsr_sort_pairs_static": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _ll, C.c_int, C.c_int, _vp]),
    "gsr_get_stats": (C.c_int, [_vp, C.POINTER(Stats)]),
    "gsr_debug_fetch": (C.c_int, [_vp, C.c_int, _vp, _ll, C.POINTER(_ll)]),
    "gsr_scene_buffer": (C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_ll)]),
    "gsr_alloc_gaussians": (C.c_int, [_vp, _ll, C.c_int]),
    "gsr_launch_count": (_ll, [_vp]),
    "gsr_measure_peaks": (C.c_int, [_vp, _fp, _fp, _fp]),
    "gsr_tune": (C.c_int, [_vp, C.c_int, C.c_int]),
    "gsr_count_pairs": (C.c_int, [_vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
}

_lib = None


def load():
    """Returns the loaded library; raises RuntimeError if libgsr.so is missing or does not export the ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m gaussiansplattingviewer_b200.build` "
            "(the renderer has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RuntimeError(f"libgsr.so does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


GSR_EOVERFLOW = -4


class GsrOverflow(RuntimeError):
    """GSR_EOVERFLOW: a frame did not fit the tile-instance buffers (they have been grown): submit it again."""


def check(rc: int):
    if rc != 0:
        msg = load().gsr_last_error()
        text = f"gsr error {rc}: {msg.decode() if msg else ''}"
        raise (GsrOverflow(text) if rc == GSR_EOVERFLOW else RuntimeError(text))
