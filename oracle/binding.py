"""ctypes binding of liboracle.so (TEST INFRASTRUCTURE ONLY — see forma_oracle.cpp header).

Importers allowed: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from .scene_oracle import OBJECT_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

FLAG_NO_JITTER, FLAG_PINHOLE, FLAG_SEQ_RNG, FLAG_MT_ROWS = 1, 2, 4, 8
FLAG_BVH = 16  # checker-side accelerator (forma_oracle.cpp OBvh): the scan's results, plane-free finite lists only
CAMERA_FLOATS = 26
# field offsets (in floats) inside the 26-float camera block
CAM_FIELDS = {
    "position": (0, 3), "lower_left_corner": (3, 6), "aspect": (6, 7), "horizontal": (7, 10),
    "vertical": (10, 13), "u": (13, 16), "v": (16, 19), "w": (19, 22), "lens_radius": (22, 23),
    "focus_dist": (23, 24), "radius": (24, 25), "rotation": (25, 26),
}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "forma_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        fp = ctypes.POINTER(ctypes.c_float)
        L.oracle_vec3_length_squared.restype = ctypes.c_float
        L.oracle_vec3_length.restype = ctypes.c_float
        L.oracle_vec3_dot.restype = ctypes.c_float
        L.oracle_u32_to_f32.restype = ctypes.c_float
        L.oracle_u32_to_f32.argtypes = [ctypes.c_uint32]
        L.oracle_schlick.restype = ctypes.c_float
        L.oracle_schlick.argtypes = [ctypes.c_float, ctypes.c_float]
        L.oracle_as_u8.argtypes = [ctypes.c_float]
        L.oracle_refract.argtypes = [fp, fp, ctypes.c_float, fp]
        L.oracle_vec3_scale.argtypes = [fp, ctypes.c_float, fp]
        L.oracle_vec3_lscale.argtypes = [ctypes.c_float, fp, fp]
        L.oracle_vec3_div.argtypes = [fp, ctypes.c_float, fp]
        L.oracle_object_hit.argtypes = [ctypes.c_void_p, fp, fp, ctypes.c_float, ctypes.c_float, fp]
        L.oracle_camera_new.argtypes = [ctypes.c_uint32, ctypes.c_uint32, fp]
        L.oracle_camera_look.argtypes = [fp, fp, fp, ctypes.c_float, ctypes.c_float, fp]
        L.oracle_camera_orbit.argtypes = [fp, fp]
        L.oracle_camera_translate.argtypes = [fp, fp]
        L.oracle_update_camera.argtypes = [fp, ctypes.c_uint8, ctypes.c_float]
        L.oracle_builtin_scene.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
        L.oracle_primary_hits.argtypes = [ctypes.c_void_p, ctypes.c_uint64, fp, ctypes.c_uint32, ctypes.c_uint32,
                                          ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_render.restype = ctypes.c_uint64
        L.oracle_render.argtypes = [ctypes.c_void_p, ctypes.c_uint64, fp, ctypes.c_uint32, ctypes.c_uint32,
                                    ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint64,
                                    ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                                    ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_tonemap.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p]
        L.oracle_philox4x32_10.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_philox4x32_r.argtypes = [ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_path_rounds.restype = ctypes.c_int
        assert L.oracle_object_size() == OBJECT_DTYPE.itemsize
        assert L.oracle_camera_floats() == CAMERA_FLOATS
        _lib = L
    return _lib


def _f3(v):
    return (ctypes.c_float * 3)(*[float(x) for x in v])


def _fptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def vec3_op(name, a, b):
    out = (ctypes.c_float * 3)()
    getattr(lib(), "oracle_vec3_" + name)(_f3(a), _f3(b), out)
    return np.array(out[:], dtype=np.float32)


def philox(k0, k1, ctr, rounds=10):
    c = np.array(ctr, dtype=np.uint32)
    o = np.zeros(4, dtype=np.uint32)
    lib().oracle_philox4x32_r(rounds, k0, k1, c.ctypes.data, o.ctypes.data)
    return o


def path_rounds() -> int:
    """Philox rounds of the path's counter RNG (oracle and kernels: 7)."""
    return lib().oracle_path_rounds()


def camera_new(w, h) -> np.ndarray:
    cam = np.zeros(CAMERA_FLOATS, dtype=np.float32)
    lib().oracle_camera_new(w, h, _fptr(cam))
    return cam


def camera_look(look_from, look_at, up, v_fov, aspect) -> np.ndarray:
    cam = np.zeros(CAMERA_FLOATS, dtype=np.float32)
    lib().oracle_camera_look(_f3(look_from), _f3(look_at), _f3(up), float(v_fov), float(aspect), _fptr(cam))
    return cam


def camera_orbit(cam: np.ndarray, delta):
    lib().oracle_camera_orbit(_fptr(cam), _f3(delta))
    return cam


def camera_translate(cam: np.ndarray, delta):
    lib().oracle_camera_translate(_fptr(cam), _f3(delta))
    return cam


def update_camera(cam: np.ndarray, keys: int, dt: float):
    lib().oracle_update_camera(_fptr(cam), keys, dt)
    return cam


def cam_field(cam, name):
    a, b = CAM_FIELDS[name]
    return cam[a:b]


def builtin_scene(which: int) -> np.ndarray:
    buf = np.zeros(16, dtype=OBJECT_DTYPE)
    n = lib().oracle_builtin_scene(which, buf.ctypes.data, len(buf))
    if n < 0:
        raise ValueError("unknown builtin scene %d" % which)
    return buf[:n].copy()


def object_hit(obj: np.ndarray, origin, direction, t_min=0.001, t_max=3.4028234663852886e38, rec=None):
    r = np.zeros(7, dtype=np.float32) if rec is None else np.asarray(rec, dtype=np.float32).copy()
    o = np.ascontiguousarray(obj.reshape(1))
    hit = lib().oracle_object_hit(o.ctypes.data, _f3(origin), _f3(direction), t_min, t_max, _fptr(r))
    return bool(hit), r


def primary_hits(objs: np.ndarray, cam: np.ndarray, w: int, h: int):
    objs = np.ascontiguousarray(objs)
    ids = np.empty(w * h, dtype=np.int32)
    ts = np.empty(w * h, dtype=np.float32)
    lib().oracle_primary_hits(objs.ctypes.data, len(objs), _fptr(cam), w, h, ids.ctypes.data, ts.ctypes.data)
    return ids.reshape(h, w), ts.reshape(h, w)


def render(objs: np.ndarray, cam: np.ndarray, w: int, h: int, spp: int, *, sample_offset=0, max_depth=50,
           seed=0, flags=0, y0=0, y1=None, nthreads=1, want_rgb=True, want_linear=True):
    """Returns (rgb u8 [h,w,3] | None, linear_sum f32 [h,w,3] | None, rays)."""
    objs = np.ascontiguousarray(objs)
    y1 = h if y1 is None else y1
    rgb = np.zeros((h, w, 3), dtype=np.uint8) if want_rgb else None
    lin = np.zeros((h, w, 3), dtype=np.float32) if want_linear else None
    rays = lib().oracle_render(objs.ctypes.data, len(objs), _fptr(cam), w, h, spp, sample_offset, max_depth, seed,
                               flags, y0, y1, nthreads,
                               rgb.ctypes.data if want_rgb else None, lin.ctypes.data if want_linear else None)
    return rgb, lin, int(rays)


def tonemap(linear_sum: np.ndarray, spp: int) -> np.ndarray:
    lin = np.ascontiguousarray(linear_sum, dtype=np.float32)
    npix = lin.size // 3
    rgb = np.zeros(lin.shape, dtype=np.uint8)
    lib().oracle_tonemap(lin.ctypes.data, npix, spp, rgb.ctypes.data)
    return rgb


def render_mt(cam: np.ndarray, w: int, h: int, frame: int = 0, seed: int = 0) -> np.ndarray:
    """tracer::render_mt (tracer.rs:83-134): one 1-spp frame of a FRESH get_simple_scene() (:105 - not the model's
    scene) through the model's camera, in four bands; returns the (4 * (h // 4), w, 3) u8 rows it produces."""
    rgb, _, _ = render(builtin_scene(0), cam, w, h, 1, sample_offset=frame, seed=seed, flags=FLAG_MT_ROWS,
                       want_linear=False)
    return rgb[:4 * (h // 4)]


def save_image_mt(cam: np.ndarray, w: int, h: int, sample: int, seed: int = 0) -> np.ndarray:
    """tracer::save_image_mt's loops (tracer.rs:139-154): the u8 frames of `sample` render_mt calls averaged AFTER
    gamma in f32 (acc += p as f32 / sample as f32), `as u8`; rows render_mt does not produce stay 0."""
    acc = np.zeros((h, w, 3), dtype=np.float32)
    for k in range(sample):
        f = render_mt(cam, w, h, k, seed)
        acc[:f.shape[0]] = acc[:f.shape[0]] + f.astype(np.float32) / np.float32(sample)
    return np.minimum(acc, np.float32(255.0)).astype(np.uint8)
