"""ctypes binding of libpostfx_oracle.so (TEST INFRASTRUCTURE ONLY — see postfx_oracle.cpp header).

Importers allowed: tests/, __graft_entry__.smoke(), the cpu legs of tools/perf_postfx.py.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpostfx_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "postfx_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        vp, u32 = ctypes.c_void_p, ctypes.c_uint32
        L.fxo_apply.argtypes = [ctypes.c_int, vp, vp, u32, u32, ctypes.c_float, u32]
        L.fxo_chain.argtypes = [vp, u32, vp, vp, u32, u32, ctypes.c_float, u32]
        _lib = L
    return _lib


def apply(effect: int, frame: np.ndarray, time: float = 0.0, out: np.ndarray | None = None, threads: int = 1):
    """One dispatch of one filter over an (h, w, 4) uint8 frame; texels it does not write keep `out`'s value."""
    src = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w, _ = src.shape
    dst = np.zeros_like(src) if out is None else out
    rc = lib().fxo_apply(int(effect), src.ctypes.data, dst.ctypes.data, w, h, ctypes.c_float(time), threads)
    if rc:
        raise ValueError("fxo_apply: bad argument")
    return dst


def chain(effects, frame: np.ndarray, time: float = 0.0, threads: int = 1):
    """The active effects in order, on the reference's ping-pong textures, first-frame state (all zero)."""
    src = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w, _ = src.shape
    dst = np.zeros_like(src)
    e = np.asarray([int(x) for x in effects], dtype=np.uint8)
    rc = lib().fxo_chain(e.ctypes.data if e.size else None, e.size, src.ctypes.data, dst.ctypes.data, w, h,
                         ctypes.c_float(time), threads)
    if rc:
        raise ValueError("fxo_chain: bad argument")
    return dst
