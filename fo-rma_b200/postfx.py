"""Host-side mirror of the reference's post-processing stage over include/forma_fx.h.

Reference surface (src/rendering/post_processor.rs, src/shader_utils.rs):
    Effect                                   shader_utils.rs:58-71   (same members, same order)
    effect_to_name(effect)                   shader_utils.rs:73-88
    PostProcessor::new(device, size, ..)     post_processor.rs:42-99
    PostProcessor::update_effects(.., map)   post_processor.rs:133-197  (active effects, map order)
    PostProcessor::run(.., width, height)    post_processor.rs:101-129
    PostProcessor::resize(.., size, ..)      post_processor.rs:199-262
The wgpu device / queue / texture-view arguments have no counterpart: frames are RGBA8 numpy arrays (h, w, 4) on the
host, or device pointers.  Every call goes to libforma_rt.so; there is no CPU path here.
"""
from __future__ import annotations

import ctypes
import enum
from typing import Iterable, Mapping

import numpy as np

from . import _lib
from ._lib import check


class Effect(enum.IntEnum):
    NONE = 0
    NOISE = 1
    PIXELATE = 2
    INVERT_COLOR = 3
    WAVE = 4
    INTERLACE = 5
    FLIP_AXIS = 6
    GRAYSCALE = 7
    STEP = 8
    WATERCOLOR = 9
    CHROMOSTEREOPSIS = 10
    ANAGLYPH = 11


def effect_to_name(effect: int) -> str:
    name = _lib.lib().forma_fx_effect_name(int(effect))
    if name is None:
        raise ValueError(f"unknown effect {effect}")
    return name.decode()


class PostProcessor:
    """PostProcessor (post_processor.rs:31-40): the intermediate frames plus the list of active effects."""

    def __init__(self, width: int, height: int, *, device: int = 0, unfused: bool = False):
        h = ctypes.c_void_p()
        check(_lib.lib().forma_fx_create(device, width, height, _lib.FX_UNFUSED if unfused else 0, ctypes.byref(h)))
        self._h = h
        self.width, self.height, self.device = width, height, device
        self.effects: list[Effect] = []

    def close(self) -> None:
        if getattr(self, "_h", None):
            _lib.lib().forma_fx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def update_effects(self, effect_to_active: Mapping[int, bool] | Iterable[int]) -> None:
        """`effect_to_active` as in the reference (an insertion-ordered map effect -> bool; the active ones run in map
        order), or simply the active effects in order."""
        if isinstance(effect_to_active, Mapping):
            active = [e for e, on in effect_to_active.items() if on]
        else:
            active = list(effect_to_active)
        arr = np.asarray([int(e) for e in active], dtype=np.uint8)
        check(_lib.lib().forma_fx_update_effects(self._h, arr.ctypes.data if arr.size else None, arr.size))
        self.effects = [Effect(int(e)) for e in arr]

    def resize(self, width: int, height: int) -> None:
        check(_lib.lib().forma_fx_resize(self._h, width, height))
        self.width, self.height = width, height

    def run(self, frame: np.ndarray, time: float = 0.0, out: np.ndarray | None = None) -> np.ndarray:
        """One PostProcessor::run over a host frame (h, w, 4) uint8; returns the filtered frame.  With no active
        effect the reference dispatches nothing: `out` is returned as it was (zeros when not given)."""
        src = np.ascontiguousarray(frame, dtype=np.uint8)
        if src.shape != (self.height, self.width, 4):
            raise ValueError(f"frame must be ({self.height}, {self.width}, 4) uint8, got {src.shape}")
        dst = np.zeros_like(src) if out is None else out
        if dst.shape != src.shape or dst.dtype != np.uint8 or not dst.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous uint8 array of the frame's shape")
        check(_lib.lib().forma_fx_run(self._h, src.ctypes.data, dst.ctypes.data, ctypes.c_float(time)))
        return dst

    def run_device(self, d_src: int, d_dst: int, time: float = 0.0, stream: int | None = None) -> None:
        check(_lib.lib().forma_fx_run_device(self._h, d_src, d_dst, ctypes.c_float(time), stream))

    def apply_device(self, effect: int, d_src: int, d_dst: int, time: float = 0.0, stream: int | None = None) -> None:
        check(_lib.lib().forma_fx_apply_device(self._h, int(effect), d_src, d_dst, ctypes.c_float(time), stream))

    def run_traced(self, model, spp: int, *, sample_offset: int = 0, max_depth: int = 50, time: float = 0.0,
                   want_rgb: bool = False):
        """Trace a frame with `model` (a tracer.TraceModel of the same size and device) and filter it on the device.
        Returns the RGBA8 frame, or (rgba, rgb) with the unfiltered RGB8 frame when want_rgb."""
        rgba = np.empty((self.height, self.width, 4), dtype=np.uint8)
        rgb = np.empty((self.height, self.width, 3), dtype=np.uint8) if want_rgb else None
        check(_lib.lib().forma_fx_run_traced(self._h, model._h, spp, sample_offset, max_depth, ctypes.c_float(time),
                                             rgb.ctypes.data if want_rgb else None, rgba.ctypes.data))
        return (rgba, rgb) if want_rgb else rgba

    def synchronize(self) -> None:
        check(_lib.lib().forma_fx_synchronize(self._h))

    def stats(self) -> _lib.FxStats:
        st = _lib.FxStats()
        check(_lib.lib().forma_fx_stats_get(self._h, ctypes.byref(st)))
        return st
