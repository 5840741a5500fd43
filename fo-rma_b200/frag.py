"""Host-side mirror of the reference's fragment-shader ray tracers over include/forma_frag.h.

Reference surface: `State` of src/frontend/wnt.rs — new (wnt.rs:81-237, builds the pipeline around
src/shaders/rt.wgsl), resize (wnt.rs:243-250), render (wnt.rs:259-302: one full-screen draw of `fs_main`) — and the
two shaders themselves, src/shaders/rt.wgsl and src/shaders/rt2.wgsl.  The window, surface and wgpu device have no
counterpart: a frame is an RGBA8 (h, w, 4) and / or f32 (h, w, 4) numpy array, or a device pointer.  Every call goes
to libforma_rt.so; there is no CPU path here.
"""
from __future__ import annotations

import ctypes
import enum

import numpy as np

from . import _lib
from ._lib import check


class Shader(enum.IntEnum):
    RT = 1   # src/shaders/rt.wgsl
    RT2 = 2  # src/shaders/rt2.wgsl


def light_direction(time: float) -> np.ndarray:
    """rt2.wgsl:144 — the light direction `radiance` derives from uniforms.time."""
    out = np.zeros(3, np.float32)
    check(_lib.lib().forma_frag_light_direction(ctypes.c_float(time), out.ctypes.data))
    return out


class State:
    """wnt.rs:64-78 `State`, headless: the render target's size and the device frames behind it."""

    def __init__(self, width: int, height: int, *, device: int = 0, shader: int = Shader.RT, srgb_surface: bool = False):
        h = ctypes.c_void_p()
        check(_lib.lib().forma_frag_create(device, width, height, ctypes.byref(h)))
        self._h = h
        self.width, self.height, self.device = width, height, device
        self.shader, self.srgb_surface = Shader(int(shader)), srgb_surface

    def close(self) -> None:
        if getattr(self, "_h", None):
            _lib.lib().forma_frag_destroy(self._h)
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

    def resize(self, width: int, height: int) -> None:
        """State::resize: a zero-sized request is ignored, as wnt.rs:244 does."""
        check(_lib.lib().forma_frag_resize(self._h, width, height))
        if width > 0 and height > 0:
            self.width, self.height = width, height

    def _params(self, time, resolution, aspect):
        rx, ry = (0.0, 0.0) if resolution is None else resolution
        return _lib.FragParams(time, rx, ry, aspect, _lib.FRAG_SRGB_SURFACE if self.srgb_surface else 0)

    def render(self, time: float = 0.0, *, shader: int | None = None, resolution=None, aspect: float = 0.0,
               want_linear: bool = False):
        """State::render: one full-screen pass of fs_main.  Returns the RGBA8 frame, or (rgba8, linear) with
        fs_main's f32 output when want_linear.  `resolution` / `aspect` default to the shaders' constants."""
        p = self._params(time, resolution, aspect)
        rgba = np.empty((self.height, self.width, 4), np.uint8)
        lin = np.empty((self.height, self.width, 4), np.float32) if want_linear else None
        check(_lib.lib().forma_frag_render(self._h, int(self.shader if shader is None else shader), ctypes.byref(p),
                                           rgba.ctypes.data, lin.ctypes.data if want_linear else None))
        return (rgba, lin) if want_linear else rgba

    def render_device(self, d_rgba8: int | None, d_linear: int | None, time: float = 0.0, *, shader: int | None = None,
                      resolution=None, aspect: float = 0.0, stream: int | None = None) -> None:
        p = self._params(time, resolution, aspect)
        check(_lib.lib().forma_frag_render_device(self._h, int(self.shader if shader is None else shader),
                                                  ctypes.byref(p), d_rgba8, d_linear, stream))

    def synchronize(self) -> None:
        check(_lib.lib().forma_frag_synchronize(self._h))

    def stats(self) -> _lib.FragStats:
        st = _lib.FragStats()
        check(_lib.lib().forma_frag_stats_get(self._h, ctypes.byref(st)))
        return st
