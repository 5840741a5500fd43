"""Host-side mirror of the reference's tracer module (src/cpu_ray_tracer/tracer.rs) over the C ABI.

Same names, argument meaning and error behaviour as the reference's public surface:

    create_model(width, height) -> TraceModel          tracer.rs:19-28
    update(model, keys, delta_time)                     tracer.rs:30-55   (fills model.pixels)
    save_image(model, sample)                           tracer.rs:160-187 (writes out/basic.png)

`TraceModel` keeps the reference's public fields (scene, width, height, pixels); `Scene` exposes
`camera` and `objects` with `translate` / `rotate` on an object (shapes/hitable.rs:12-13).  Where the
reference panics (`unwrap`), these raise FormaError.  Everything renders on the GPU through
libforma_rt.so; there is no CPU path here.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import _lib
from ._lib import FormaError, check  # noqa: F401

MAX_DEPTH = 50  # tracer.rs:10
CHANNEL_COUNT = 3  # tracer.rs:9

OBJECT_DTYPE = np.dtype([("kind", "<i4"), ("material", "<i4"), ("id", "<i4"), ("fuzz", "<f4"),
                         ("color", "<f4", (3,)), ("a", "<f4", (3,)), ("b", "<f4", (3,)), ("c", "<f4", (3,))])


def _f3(v):
    return (ctypes.c_float * 3)(*[float(x) for x in v])


class ObjectRef:
    """objects[i] of Scene.objects: Hitable::translate / rotate (plane.rs:62-68 assign; others no-op)."""

    def __init__(self, model: "TraceModel", index: int):
        self._m, self._i = model, index

    def translate(self, v):
        check(_lib.lib().forma_rt_set_object(self._m._h, self._i, _f3(v), None))

    def rotate(self, v):
        check(_lib.lib().forma_rt_set_object(self._m._h, self._i, None, _f3(v)))

    def record(self) -> np.ndarray:
        return self._m.scene.object_records()[self._i]


class _Objects:
    def __init__(self, model):
        self._m = model

    def __len__(self):
        n = ctypes.c_size_t(0)
        check(_lib.lib().forma_rt_get_objects(self._m._h, None, 0, ctypes.byref(n)))
        return n.value

    def __getitem__(self, i):
        if not 0 <= i < len(self):
            raise IndexError(i)
        return ObjectRef(self._m, i)


class CameraRef:
    """Scene.camera (camera.rs:7-21)."""

    def __init__(self, model):
        self._m = model

    def get(self) -> _lib.Camera:
        cam = _lib.Camera()
        check(_lib.lib().forma_rt_camera_get(self._m._h, ctypes.byref(cam)))
        return cam

    def as_array(self) -> np.ndarray:
        return np.frombuffer(bytes(self.get()), dtype=np.float32).copy()

    def set_array(self, a: np.ndarray):
        a = np.ascontiguousarray(a, dtype=np.float32)
        assert a.size == 26
        cam = _lib.Camera.from_buffer_copy(a.tobytes())
        check(_lib.lib().forma_rt_camera_set(self._m._h, ctypes.byref(cam)))

    def orbit(self, delta):  # camera.rs:97-122
        check(_lib.lib().forma_rt_camera_orbit(self._m._h, _f3(delta)))

    def translate(self, delta):  # camera.rs:74-95
        check(_lib.lib().forma_rt_camera_translate(self._m._h, _f3(delta)))

    def reset(self):  # Camera::new
        check(_lib.lib().forma_rt_camera_new(self._m._h))

    def look(self, look_from, look_at, up, v_fov):
        check(_lib.lib().forma_rt_camera_look(self._m._h, _f3(look_from), _f3(look_at), _f3(up), float(v_fov)))

    def use_json(self):
        check(_lib.lib().forma_rt_camera_json(self._m._h))


class Scene:
    """cpu_ray_tracer/scene.rs:4-7."""

    def __init__(self, model):
        self._m = model
        self.camera = CameraRef(model)
        self.objects = _Objects(model)

    def load_builtin(self, which: int):
        check(_lib.lib().forma_rt_load_builtin(self._m._h, which))

    def load_json(self, text):
        """scene_loader::construct_scene_from_json (basics/scene_loader.rs:3-7) + the RT mapping."""
        if isinstance(text, str):
            text = text.encode("utf-8")
        check(_lib.lib().forma_rt_load_json(self._m._h, text, len(text)))

    def load_objects(self, records: np.ndarray):
        r = np.ascontiguousarray(records, dtype=OBJECT_DTYPE)
        check(_lib.lib().forma_rt_load_objects(self._m._h, r.ctypes.data, len(r)))

    def object_records(self) -> np.ndarray:
        n = len(self.objects)
        out = np.zeros(n, dtype=OBJECT_DTYPE)
        cnt = ctypes.c_size_t(0)
        check(_lib.lib().forma_rt_get_objects(self._m._h, out.ctypes.data, n, ctypes.byref(cnt)))
        return out


class TraceModel:
    """tracer.rs:12-17: { scene, width, height, pixels } — pixels is RGB8, row-major, row 0 = top."""

    def __init__(self, width: int, height: int, *, device: int = 0, seed: int = 0, flags: int = 0,
                 accel: int = _lib.ACCEL_AUTO):
        L = _lib.lib()
        opts = _lib.Options(ctypes.sizeof(_lib.Options), device, seed, flags, accel)
        h = ctypes.c_void_p()
        check(L.forma_rt_create(ctypes.byref(opts), width, height, ctypes.byref(h)))
        self._h = h
        self.width, self.height = width, height
        self.pixels = np.zeros(width * height * CHANNEL_COUNT, dtype=np.uint8)  # tracer.rs:26
        self.scene = Scene(self)

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().forma_rt_destroy(self._h)
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

    # ---- added entry points (documented in include/forma_rt.h) ----
    def render(self, spp: int, *, sample_offset: int = 0, max_depth: int = MAX_DEPTH, want_rgb=True,
               want_linear=False, rgb_out: np.ndarray | None = None):
        """Returns (rgb [h,w,3] u8 | None, linear_rgba [h,w,4] f32 | None)."""
        rgb = None
        if want_rgb:
            rgb = rgb_out if rgb_out is not None else np.empty((self.height, self.width, 3), dtype=np.uint8)
        lin = np.empty((self.height, self.width, 4), dtype=np.float32) if want_linear else None
        check(_lib.lib().forma_rt_render(self._h, spp, sample_offset, max_depth,
                                         rgb.ctypes.data if rgb is not None else None,
                                         lin.ctypes.data if lin is not None else None))
        return rgb, lin

    def accumulate(self, spp_more: int, *, max_depth: int = MAX_DEPTH, want_linear=False):
        """Progressive refinement: spp_more samples on top of the handle's running sum.
        Returns (rgb of the average so far, linear_rgba | None, samples accumulated)."""
        rgb = np.empty((self.height, self.width, 3), dtype=np.uint8)
        lin = np.empty((self.height, self.width, 4), dtype=np.float32) if want_linear else None
        n = ctypes.c_uint32(0)
        check(_lib.lib().forma_rt_accumulate(self._h, spp_more, max_depth, rgb.ctypes.data,
                                             lin.ctypes.data if lin is not None else None, ctypes.byref(n)))
        return rgb, lin, n.value

    def set_partition(self, rank: int, world: int):
        """Tile-interleaved multi-GPU split: render only the 32x4 tiles t with t % world == rank."""
        check(_lib.lib().forma_rt_set_partition(self._h, rank, world))

    def accumulate_reset(self):
        check(_lib.lib().forma_rt_accumulate_reset(self._h))

    def srgb_encode_device(self, d_rgb: int, nbytes: int, stream: int = 0):
        """color_utils::srgb_encode (color_utils.rs:49-66) in place on a device RGB8 buffer."""
        check(_lib.lib().forma_rt_srgb_encode_device(self._h, d_rgb, nbytes, stream or None))

    def render_mt(self, frame: int = 0) -> np.ndarray:
        """tracer::render_mt (tracer.rs:83-134): the (4 * (h // 4), w, 3) u8 rows of one banded 1-spp frame of a fresh
        get_simple_scene() through this model's camera."""
        out = np.zeros((4 * (self.height // 4), self.width, 3), dtype=np.uint8)
        buf = out if out.size else np.zeros(1, np.uint8)
        check(_lib.lib().forma_rt_render_mt(self._h, frame, buf.ctypes.data))
        return out

    def render_mt_average(self, sample: int) -> np.ndarray:
        """save_image_mt's loops (tracer.rs:139-154) without the PNG: (h, w, 3) u8."""
        out = np.empty((self.height, self.width, 3), dtype=np.uint8)
        check(_lib.lib().forma_rt_render_mt_average(self._h, sample, out.ctypes.data))
        return out

    def primary_hits(self):
        ids = np.empty((self.height, self.width), dtype=np.int32)
        ts = np.empty((self.height, self.width), dtype=np.float32)
        check(_lib.lib().forma_rt_primary_hits(self._h, ids.ctypes.data, ts.ctypes.data))
        return ids, ts

    def stats(self) -> _lib.Stats:
        s = _lib.Stats()
        check(_lib.lib().forma_rt_stats_get(self._h, ctypes.byref(s)))
        return s

    def walk_stats(self, spp: int, *, max_depth: int = MAX_DEPTH) -> _lib.WalkStats:
        """Instrumented (untimed) frame: node visits / leaf tests of the LBVH walk, plus this device's measured
        streaming-from-L2 read bandwidth (config D's L2 roofline)."""
        ws = _lib.WalkStats()
        check(_lib.lib().forma_rt_walk_stats_get(self._h, spp, max_depth, ctypes.byref(ws)))
        return ws

    # device-resident variants: pointers are raw CUDA addresses (e.g. torch.Tensor.data_ptr())
    def render_device(self, spp, sample_offset, max_depth, d_rgb: int | None, d_sum: int | None, stream: int = 0):
        check(_lib.lib().forma_rt_render_device(self._h, spp, sample_offset, max_depth, d_rgb, d_sum, stream or None))

    def render_device_noalloc(self, spp, sample_offset, max_depth, d_rgb: int | None, d_sum: int | None, stream: int = 0):
        """render_device without any allocation (static device scratch; width * height <= _lib.NOALLOC_MAX_PIXELS, the
        scene resident from an earlier render call).  Calls that overlap on one GPU race on the scratch."""
        check(_lib.lib().forma_rt_render_device_noalloc(self._h, spp, sample_offset, max_depth, d_rgb, d_sum, stream or None))

    def tonemap_device(self, d_sum: int, spp_total: int, d_rgb: int, pixel_begin=0, pixel_end=None, stream: int = 0):
        pixel_end = self.width * self.height if pixel_end is None else pixel_end
        check(_lib.lib().forma_rt_tonemap_device(self._h, d_sum, spp_total, pixel_begin, pixel_end, d_rgb, stream or None))

    def reduce_tonemap_device(self, peer_ptrs, spp_total, pixel_begin, pixel_end, d_rgb, d_sum_out=None, stream: int = 0):
        arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
        check(_lib.lib().forma_rt_reduce_tonemap_device(self._h, arr, len(peer_ptrs), spp_total, pixel_begin, pixel_end,
                                                        d_rgb, d_sum_out, stream or None))

    def flag_set_device(self, d_flag: int, value: int, stream: int = 0):
        """*d_flag = value once everything queued on `stream` so far is complete and visible system-wide."""
        check(_lib.lib().forma_rt_flag_set_device(self._h, d_flag, value & 0xFFFFFFFF, stream or None))

    def flag_wait_device(self, d_flags, value: int, stream: int = 0):
        """Later work on `stream` starts once every flag in d_flags has reached `value`."""
        arr = (ctypes.c_void_p * len(d_flags))(*d_flags)
        check(_lib.lib().forma_rt_flag_wait_device(self._h, arr, len(d_flags), value & 0xFFFFFFFF, stream or None))

    def ipc_export(self, d_ptr: int) -> bytes:
        buf = (ctypes.c_uint8 * 64)()
        check(_lib.lib().forma_rt_ipc_export(self._h, d_ptr, buf))
        return bytes(buf)

    def ipc_open(self, handle: bytes) -> int:
        buf = (ctypes.c_uint8 * 64).from_buffer_copy(handle)
        p = ctypes.c_void_p()
        check(_lib.lib().forma_rt_ipc_open(self._h, buf, ctypes.byref(p)))
        return p.value

    def ipc_close(self, d_ptr: int):
        check(_lib.lib().forma_rt_ipc_close(self._h, d_ptr))

    def synchronize(self):
        check(_lib.lib().forma_rt_synchronize(self._h))

    def flag_error(self) -> int:
        """0, or 1 + the index of the flag a flag_wait_device of this handle gave up on (raises FormaError then)."""
        n = ctypes.c_uint32(0)
        check(_lib.lib().forma_rt_flag_error(self._h, ctypes.byref(n)))
        return n.value

    def host_alloc(self, nbytes: int) -> np.ndarray:
        """Page-locked host buffer owned by the handle, as a uint8 array (a `pixels` buffer at the pinned D2H rate)."""
        p = ctypes.c_void_p()
        check(_lib.lib().forma_rt_host_alloc(self._h, nbytes, ctypes.byref(p)))
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), shape=(nbytes,))

    def host_free(self, buf: np.ndarray) -> None:
        check(_lib.lib().forma_rt_host_free(self._h, buf.ctypes.data))


def create_model(width: int, height: int, **kw) -> TraceModel:
    """tracer::create_model (tracer.rs:19-28)."""
    return TraceModel(width, height, **kw)


def update(model: TraceModel, keys: int, delta_time: float) -> None:
    """tracer::update (tracer.rs:30-55): keys = 00EQADWS; replaces model.pixels with a 1-spp frame."""
    pixels = np.empty(model.width * model.height * CHANNEL_COUNT, dtype=np.uint8)
    check(_lib.lib().forma_rt_update(model._h, keys & 0xFF, float(delta_time), pixels.ctypes.data))
    model.pixels = pixels


def update_progressive(model: TraceModel, keys: int, delta_time: float) -> int:
    """tracer::update with progressive refinement (added mode): one more sample per call while the camera and
    the scene stand still, a fresh 1-spp frame otherwise.  Returns the samples accumulated in model.pixels."""
    pixels = np.empty(model.width * model.height * CHANNEL_COUNT, dtype=np.uint8)
    n = ctypes.c_uint32(0)
    check(_lib.lib().forma_rt_update_progressive(model._h, keys & 0xFF, float(delta_time), pixels.ctypes.data, ctypes.byref(n)))
    model.pixels = pixels
    return n.value


def srgb_table() -> np.ndarray:
    """color_utils::srgb_encode (color_utils.rs:49-66) as its u8 -> u8 table."""
    t = np.zeros(256, dtype=np.uint8)
    check(_lib.lib().forma_rt_srgb_table(t.ctypes.data))
    return t


def save_image_mt(model: TraceModel, sample: int, path: str = "out/basic_mt.png") -> None:
    """tracer::save_image_mt (tracer.rs:136-158).  Raises FormaError(EIO) where the reference panics."""
    check(_lib.lib().forma_rt_save_image_mt(model._h, sample, os.fsencode(path)))


def save_image(model: TraceModel, sample: int, path: str = "out/basic.png") -> None:
    """tracer::save_image (tracer.rs:160-187).  Raises FormaError(EIO) where the reference panics
    (e.g. ./out missing, tracer.rs:186 unwrap)."""
    check(_lib.lib().forma_rt_save_image(model._h, sample, os.fsencode(path)))
