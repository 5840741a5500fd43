"""ctypes binding of libforma_rt.so — every symbol include/forma_rt.h, forma_fx.h and forma_frag.h declare,
nothing else.

The library is built in-tree by fo-rma_b200/csrc/Makefile (`__graft_entry__.build()`); importing
this module fails loudly when it is missing.  There is no CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# FORMA_RT_LIB: another build of the same library (occupancy / flag experiments); never a different implementation
LIB_PATH = os.environ.get("FORMA_RT_LIB") or os.path.join(_HERE, "libforma_rt.so")

OK, EINVAL, EJSON, ECUDA, EIO, ENOMEM = 0, -1, -2, -3, -4, -5
NO_JITTER, PINHOLE, HOST_ONLY, SRGB_OUT = 1, 2, 4, 8
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_LBVH = 0, 1, 2
SPHERE, PLANE, TRIANGLE = 0, 1, 2
NOALLOC_MAX_PIXELS = 262144  # FORMA_RT_NOALLOC_MAX_PIXELS


class Options(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_uint32), ("device", ctypes.c_int32), ("seed", ctypes.c_uint64),
                ("flags", ctypes.c_uint32), ("accel", ctypes.c_uint32)]


class Object(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("material", ctypes.c_int32), ("id", ctypes.c_int32),
                ("fuzz", ctypes.c_float), ("color", ctypes.c_float * 3), ("a", ctypes.c_float * 3),
                ("b", ctypes.c_float * 3), ("c", ctypes.c_float * 3)]


class Camera(ctypes.Structure):
    _fields_ = [("position", ctypes.c_float * 3), ("lower_left_corner", ctypes.c_float * 3),
                ("aspect", ctypes.c_float), ("horizontal", ctypes.c_float * 3), ("vertical", ctypes.c_float * 3),
                ("u", ctypes.c_float * 3), ("v", ctypes.c_float * 3), ("w", ctypes.c_float * 3),
                ("lens_radius", ctypes.c_float), ("focus_dist", ctypes.c_float), ("radius", ctypes.c_float),
                ("rotation", ctypes.c_float)]


class Stats(ctypes.Structure):
    _fields_ = [("rays", ctypes.c_uint64), ("paths", ctypes.c_uint64), ("kernel_ms", ctypes.c_float),
                ("total_ms", ctypes.c_float), ("gpu_launches", ctypes.c_uint32), ("n_objects", ctypes.c_uint32),
                ("n_spheres", ctypes.c_uint32), ("n_planes", ctypes.c_uint32), ("n_triangles", ctypes.c_uint32),
                ("bvh_nodes", ctypes.c_uint32), ("accel_in_use", ctypes.c_uint32), ("sm_count", ctypes.c_uint32),
                ("sm_clock_khz", ctypes.c_uint32), ("reserved", ctypes.c_uint32), ("culled_paths", ctypes.c_uint64)]


class WalkStats(ctypes.Structure):
    _fields_ = [("rays", ctypes.c_uint64), ("node_visits", ctypes.c_uint64), ("leaf_tests", ctypes.c_uint64),
                ("node_bytes", ctypes.c_uint32), ("leaf_bytes", ctypes.c_uint32), ("bvh_nodes", ctypes.c_uint32),
                ("l2_read_gbs", ctypes.c_float)]


assert ctypes.sizeof(Object) == 64 and ctypes.sizeof(Camera) == 104

_vp, _u8p, _fp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_float)
_u32, _u64, _int, _sz = ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t

# name -> (restype, argtypes): the full surface of include/forma_rt.h
SYMBOLS = {
    "forma_rt_create": (_int, [ctypes.POINTER(Options), _u32, _u32, ctypes.POINTER(_vp)]),
    "forma_rt_destroy": (None, [_vp]),
    "forma_rt_last_error": (ctypes.c_char_p, []),
    "forma_rt_abi_version": (_u32, []),
    "forma_rt_load_builtin": (_int, [_vp, _int]),
    "forma_rt_load_json": (_int, [_vp, ctypes.c_char_p, _sz]),
    "forma_rt_load_objects": (_int, [_vp, _vp, _sz]),
    "forma_rt_get_objects": (_int, [_vp, _vp, _sz, ctypes.POINTER(_sz)]),
    "forma_rt_set_object": (_int, [_vp, _u32, _fp, _fp]),
    "forma_rt_camera_new": (_int, [_vp]),
    "forma_rt_camera_look": (_int, [_vp, _fp, _fp, _fp, ctypes.c_float]),
    "forma_rt_camera_json": (_int, [_vp]),
    "forma_rt_camera_orbit": (_int, [_vp, _fp]),
    "forma_rt_camera_translate": (_int, [_vp, _fp]),
    "forma_rt_camera_get": (_int, [_vp, ctypes.POINTER(Camera)]),
    "forma_rt_camera_set": (_int, [_vp, ctypes.POINTER(Camera)]),
    "forma_rt_update": (_int, [_vp, ctypes.c_uint8, ctypes.c_float, _vp]),
    "forma_rt_flag_error": (_int, [_vp, ctypes.POINTER(_u32)]),
    "forma_rt_host_alloc": (_int, [_vp, _sz, ctypes.POINTER(_vp)]),
    "forma_rt_host_free": (_int, [_vp, _vp]),
    "forma_rt_render_device_noalloc": (_int, [_vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "forma_rt_set_partition": (_int, [_vp, _u32, _u32]),
    "forma_rt_accumulate": (_int, [_vp, _u32, _u32, _vp, _vp, ctypes.POINTER(_u32)]),
    "forma_rt_accumulate_reset": (_int, [_vp]),
    "forma_rt_update_progressive": (_int, [_vp, ctypes.c_uint8, ctypes.c_float, _vp, ctypes.POINTER(_u32)]),
    "forma_rt_srgb_encode_device": (_int, [_vp, _vp, _sz, _vp]),
    "forma_rt_srgb_table": (_int, [_vp]),
    "forma_rt_render": (_int, [_vp, _u32, _u32, _u32, _vp, _vp]),
    "forma_rt_save_image": (_int, [_vp, _u32, ctypes.c_char_p]),
    "forma_rt_primary_hits": (_int, [_vp, _vp, _vp]),
    "forma_rt_render_mt": (_int, [_vp, _u32, _vp]),
    "forma_rt_render_mt_average": (_int, [_vp, _u32, _vp]),
    "forma_rt_save_image_mt": (_int, [_vp, _u32, ctypes.c_char_p]),
    "forma_rt_render_device": (_int, [_vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "forma_rt_tonemap_device": (_int, [_vp, _vp, _u32, _u64, _u64, _vp, _vp]),
    "forma_rt_reduce_tonemap_device": (_int, [_vp, ctypes.POINTER(_vp), _u32, _u32, _u64, _u64, _vp, _vp, _vp]),
    "forma_rt_flag_set_device": (_int, [_vp, _vp, _u32, _vp]),
    "forma_rt_flag_wait_device": (_int, [_vp, ctypes.POINTER(_vp), _u32, _u32, _vp]),
    "forma_rt_ipc_export": (_int, [_vp, _vp, _vp]),
    "forma_rt_ipc_open": (_int, [_vp, _vp, ctypes.POINTER(_vp)]),
    "forma_rt_ipc_close": (_int, [_vp, _vp]),
    "forma_rt_device_alloc": (_int, [_vp, _sz, ctypes.POINTER(_vp)]),
    "forma_rt_device_free": (_int, [_vp, _vp]),
    "forma_rt_synchronize": (_int, [_vp]),
    "forma_rt_stats_get": (_int, [_vp, ctypes.POINTER(Stats)]),
    "forma_rt_write_png": (_int, [ctypes.c_char_p, _vp, _u32, _u32]),
    "forma_rt_walk_stats_get": (_int, [_vp, _u32, _u32, ctypes.POINTER(WalkStats)]),
}


class FxStats(ctypes.Structure):
    _fields_ = [("n_effects", ctypes.c_uint32), ("n_passes", ctypes.c_uint32), ("gpu_launches", ctypes.c_uint32),
                ("kernel_ms", ctypes.c_float), ("total_ms", ctypes.c_float), ("bytes_moved", ctypes.c_uint64)]


FX_UNFUSED = 1
# name -> (restype, argtypes): the full surface of include/forma_fx.h (same library)
SYMBOLS_FX = {
    "forma_fx_create": (_int, [_int, _u32, _u32, _u32, ctypes.POINTER(_vp)]),
    "forma_fx_destroy": (None, [_vp]),
    "forma_fx_resize": (_int, [_vp, _u32, _u32]),
    "forma_fx_update_effects": (_int, [_vp, _vp, _u32]),
    "forma_fx_run": (_int, [_vp, _vp, _vp, ctypes.c_float]),
    "forma_fx_run_device": (_int, [_vp, _vp, _vp, ctypes.c_float, _vp]),
    "forma_fx_apply_device": (_int, [_vp, _int, _vp, _vp, ctypes.c_float, _vp]),
    "forma_fx_run_traced": (_int, [_vp, _vp, _u32, _u32, _u32, ctypes.c_float, _vp, _vp]),
    "forma_fx_effect_name": (ctypes.c_char_p, [_int]),
    "forma_fx_synchronize": (_int, [_vp]),
    "forma_fx_stats_get": (_int, [_vp, ctypes.POINTER(FxStats)]),
}


class FragParams(ctypes.Structure):
    _fields_ = [("time", ctypes.c_float), ("res_x", ctypes.c_float), ("res_y", ctypes.c_float),
                ("aspect", ctypes.c_float), ("flags", ctypes.c_uint32)]


class FragStats(ctypes.Structure):
    _fields_ = [("rays", ctypes.c_uint64), ("gpu_launches", ctypes.c_uint32), ("kernel_ms", ctypes.c_float),
                ("total_ms", ctypes.c_float)]


FRAG_RT, FRAG_RT2 = 1, 2
FRAG_SRGB_SURFACE = 1
# name -> (restype, argtypes): the full surface of include/forma_frag.h (same library)
SYMBOLS_FRAG = {
    "forma_frag_create": (_int, [_int, _u32, _u32, ctypes.POINTER(_vp)]),
    "forma_frag_destroy": (None, [_vp]),
    "forma_frag_resize": (_int, [_vp, _u32, _u32]),
    "forma_frag_render": (_int, [_vp, _int, ctypes.POINTER(FragParams), _vp, _vp]),
    "forma_frag_render_device": (_int, [_vp, _int, ctypes.POINTER(FragParams), _vp, _vp, _vp]),
    "forma_frag_light_direction": (_int, [ctypes.c_float, _vp]),
    "forma_frag_synchronize": (_int, [_vp]),
    "forma_frag_stats_get": (_int, [_vp, ctypes.POINTER(FragStats)]),
}


def build_library(verbose: bool = False) -> str:
    """Compile csrc/ for sm_100a into libforma_rt.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc")]
    subprocess.check_call(cmd, stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C fo-rma_b200/csrc` or __graft_entry__.build(). "
                "forma_b200 has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in {**SYMBOLS, **SYMBOLS_FX, **SYMBOLS_FRAG}.items():
            fn = getattr(L, name)  # AttributeError here = header/library skew
            fn.restype, fn.argtypes = res, args
        if L.forma_rt_abi_version() != 3:
            raise ImportError("libforma_rt.so ABI version mismatch")
        _lib = L
    return _lib


class FormaError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"forma_rt error {code}: {message}")
        self.code = code


def check(rc: int) -> None:
    if rc != 0:
        raise FormaError(rc, lib().forma_rt_last_error().decode("utf-8", "replace"))
