"""ctypes binding of libwgsl_rt_oracle.so (TEST INFRASTRUCTURE ONLY — see wgsl_rt_oracle.cpp header).

Importers allowed: tests/, __graft_entry__.smoke(), the cpu legs of tools/perf_frag.py.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libwgsl_rt_oracle.so")
_lib = None

RT = 1   # src/shaders/rt.wgsl
RT2 = 2  # src/shaders/rt2.wgsl


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "wgsl_rt_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        f, vp, u32 = ctypes.c_float, ctypes.c_void_p, ctypes.c_uint32
        for name in ("wgo_log2", "wgo_exp2"):
            getattr(L, name).argtypes = [f]
            getattr(L, name).restype = f
        L.wgo_pow.argtypes = [f, f]
        L.wgo_pow.restype = f
        L.wgo_light2.argtypes = [f, vp]
        L.wgo_light2.restype = None
        L.wgo_render.argtypes = [ctypes.c_int, u32, u32, f, f, f, f, ctypes.c_int, vp, vp, vp, vp, u32]
        _lib = L
    return _lib


def log2(x: float) -> float:
    return lib().wgo_log2(x)


def exp2(x: float) -> float:
    return lib().wgo_exp2(x)


def pow(x: float, y: float) -> float:  # noqa: A001 - the WGSL builtin's name
    return lib().wgo_pow(x, y)


def light2(time: float) -> np.ndarray:
    out = np.zeros(3, np.float32)
    lib().wgo_light2(time, out.ctypes.data)
    return out


def render(shader: int, w: int, h: int, time: float = 0.0, srgb: bool = False, resolution=None, aspect: float = 0.0,
           threads: int = 1, want_bounces: bool = False):
    """One frame of fs_main: returns (rgba8 (h, w, 4) u8, linear (h, w, 4) f32, rays) [+ bounces (h, w) u32]."""
    rgba = np.zeros((h, w, 4), np.uint8)
    lin = np.zeros((h, w, 4), np.float32)
    nb = np.zeros((h, w), np.uint32) if want_bounces else None
    rays = ctypes.c_uint64(0)
    rx, ry = (0.0, 0.0) if resolution is None else resolution
    rc = lib().wgo_render(int(shader), w, h, rx, ry, aspect, time, int(srgb), lin.ctypes.data, rgba.ctypes.data,
                          nb.ctypes.data if want_bounces else None, ctypes.addressof(rays), threads)
    if rc:
        raise ValueError("wgo_render: bad argument")
    return (rgba, lin, rays.value, nb) if want_bounces else (rgba, lin, rays.value)
