"""CPU ORACLE (test infrastructure only) for the scene front-end: scenes/*.json -> the flat
`Object` list the tracer scans, and the JSON camera -> look-at parameters.

The reference tracer NEVER reads scenes/*.json (cpu_ray_tracer/scene.rs:10-15 hard-codes
scenes::get_simple_scene) and has no triangle intersector, so this JSON -> ray-traceable mapping
is an EXTENSION with no reference behaviour to match ("parity unpinned").  What IS taken from the
reference: the JSON schema (src/basics/scene_loader.rs:9-66), the mesh tables the mesh names
refer to (src/primitives/{cube,quad,tetrahedron,triangle,circle,cylinder}.rs), the mesh-name
dispatch with its quad fallback (src/basics/scene.rs:82-98), cylinder sector_count 30
(scene.rs:90), the TRS order of glam's Mat4::from_scale_rotation_translation
(src/primitives/primitive.rs:65-69), mesh sphere RADIUS 0.5 (src/primitives/sphere.rs:7) and the
default diffuse colour CP0.palette[0] (src/color_utils.rs:101-109, basics/scene.rs:292-294).

Written in numpy float32 scalar arithmetic on purpose: the product's front-end is C++, so the two
implementations share no code and a transcription slip in either shows up as a mismatch.
cosf/sinf come from the platform libm through ctypes (what Rust's f32::cos/sin call on Linux).
"""
from __future__ import annotations

import ctypes
import ctypes.util
import json

import numpy as np

f32 = np.float32

OBJECT_DTYPE = np.dtype(
    [
        ("kind", "<i4"),
        ("material", "<i4"),
        ("id", "<i4"),
        ("fuzz", "<f4"),
        ("color", "<f4", (3,)),
        ("a", "<f4", (3,)),
        ("b", "<f4", (3,)),
        ("c", "<f4", (3,)),
    ],
    align=False,
)
assert OBJECT_DTYPE.itemsize == 64

KIND_SPHERE, KIND_PLANE, KIND_TRIANGLE = 0, 1, 2
MATERIALS = {"lambertian": 0, "metal": 1, "dielectric": 2, "light": 3}
DEFAULT_COLOR = (0.263, 0.208, 0.655)  # color_utils.rs:103

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.cosf.restype = ctypes.c_float
_libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]


def cosf(x) -> np.float32:
    return f32(_libm.cosf(float(x)))


def sinf(x) -> np.float32:
    return f32(_libm.sinf(float(x)))


# ---- mesh tables: positions + triangle index triples -------------------------------------------
def _cube():
    h = 0.5
    P = [
        (-h, -h, h), (h, -h, h), (h, h, h), (-h, h, h),        # front  z=+0.5   cube.rs:9-12
        (-h, -h, -h), (h, -h, -h), (h, h, -h), (-h, h, -h),    # back   z=-0.5   cube.rs:14-17
        (h, -h, -h), (h, h, -h), (h, h, h), (h, -h, h),        # right  x=+0.5   cube.rs:19-22
        (-h, -h, -h), (-h, h, -h), (-h, h, h), (-h, -h, h),    # left   x=-0.5   cube.rs:24-27
        (-h, h, -h), (h, h, -h), (h, h, h), (-h, h, h),        # top    y=+0.5   cube.rs:29-32
        (-h, -h, -h), (h, -h, -h), (h, -h, h), (-h, -h, h),    # bottom y=-0.5   cube.rs:34-37
    ]
    I = [0, 3, 1, 1, 3, 2, 4, 5, 6, 6, 7, 4, 8, 11, 10, 10, 9, 8,     # cube.rs:40-47
         12, 13, 15, 13, 14, 15, 16, 17, 18, 16, 18, 19, 20, 23, 21, 21, 23, 22]
    return P, I


def _quad():  # quad.rs:7-17
    return [(-0.5, -0.5, 0.0), (0.5, -0.5, 0.0), (0.5, 0.5, 0.0), (-0.5, 0.5, 0.0)], [0, 1, 2, 2, 3, 0]


def _tetrahedron():  # tetrahedron.rs:10-32 (X = 1.0)
    X = 1.0
    P = [(X, X, -X), (X, -X, X), (-X, X, X),
         (-X, X, X), (-X, -X, -X), (X, X, -X),
         (-X, X, X), (X, -X, X), (-X, -X, -X),
         (X, X, -X), (-X, -X, -X), (X, -X, X)]
    return P, list(range(12))


def _triangle():  # triangle.rs:6-13
    return [(0.0, 0.5, 0.0), (-0.5, -0.5, 0.0), (0.5, -0.5, 0.0)], [0, 1, 2]


_PI32 = f32(np.pi)  # std::f32::consts::PI


def _sector_xy(i, n):
    # `i as f32 * 2.0 * PI / n as f32`, left to right in f32 (circle.rs:68, cylinder.rs:72)
    ang = f32(i) * f32(2.0) * _PI32 / f32(n)
    return f32(0.5) * cosf(ang), f32(0.5) * sinf(ang)


def _circle():  # circle.rs:10-12,65-107 (36 sectors, centre vertex last)
    n = 36
    P = []
    for i in range(n):
        x, y = _sector_xy(i, n)
        P.append((x, y, 0.0))
    P.append((0.0, 0.0, 0.0))
    I = []
    for i in range(n):
        I += [i, (i + 1) % n, n]
    return P, I


def _cylinder(n=30):  # cylinder.rs:68-190 with sector_count = 30 (basics/scene.rs:90)
    P = []
    for yy in (0.5, -0.5):  # top ring + centre, bottom ring + centre
        for i in range(n):
            x, z = _sector_xy(i, n)
            P.append((x, yy, z))
        P.append((0.0, yy, 0.0))
    for yy in (0.5, -0.5):  # side rings
        for i in range(n):
            x, z = _sector_xy(i, n)
            P.append((x, yy, z))
    I = []
    for i in range(n):
        I += [i, (i + 1) % n, n]
    off = n + 1
    for i in range(n):
        I += [(i + 1) % n + off, i + off, n + off]
    off = 2 * (n + 1)
    for i in range(n):
        I += [(i + 1) % n + off, i + off, (i + 1) % n + off + n]
        I += [i + off + n, (i + 1) % n + off + n, i + off]
    return P, I


_MESHES = {}


def mesh(name):
    """basics/scene.rs:82-98: unknown mesh names fall back to the quad."""
    if name not in _MESHES:
        gen = {"cube": _cube, "triangle": _triangle, "circle": _circle, "cylinder": _cylinder,
               "tetrahedron": _tetrahedron}.get(name, _quad)
        P, I = gen()
        _MESHES[name] = (np.array(P, dtype=f32), np.array(I, dtype=np.int32).reshape(-1, 3))
    return _MESHES[name]


# ---- glam Mat4::from_scale_rotation_translation, f32 (primitive.rs:65-69) -----------------------
def quat_axes(q):
    x, y, z, w = (f32(q[0]), f32(q[1]), f32(q[2]), f32(q[3]))
    x2, y2, z2 = x + x, y + y, z + z
    xx, xy, xz = x * x2, x * y2, x * z2
    yy, yz, zz = y * y2, y * z2, z * z2
    wx, wy, wz = w * x2, w * y2, w * z2
    one = f32(1.0)
    ax = (one - (yy + zz), xy + wz, xz - wy)
    ay = (xy - wz, one - (xx + zz), yz + wx)
    az = (xz + wy, yz - wx, one - (xx + yy))
    return ax, ay, az


def transform_point(p, scale, axes, t):
    ax, ay, az = axes
    sx, sy, sz = f32(scale[0]), f32(scale[1]), f32(scale[2])
    out = []
    for k in range(3):
        cx, cy, cz = ax[k] * sx, ay[k] * sy, az[k] * sz  # scaled columns
        out.append(((cx * f32(p[0]) + cy * f32(p[1])) + cz * f32(p[2])) + f32(t[k]))
    return out


def _vec(d):
    return (f32(d["x"]), f32(d["y"]), f32(d["z"]))


def _quat(d):
    return (f32(d["x"]), f32(d["y"]), f32(d["z"]), f32(d["w"]))


def _id(oi, ti):
    """(object index << 16) | triangle index, wrapped to int32 (object index taken mod 2^16)."""
    v = ((oi << 16) | ti) & 0xFFFFFFFF
    return v - (1 << 32) if v >= (1 << 31) else v


def flatten(scene: dict) -> np.ndarray:
    """objects[] of the scene_loader.rs schema -> Object records, in list order."""
    recs = []
    for oi, o in enumerate(scene["objects"]):
        rt = o.get("rt") or {}
        material = MATERIALS.get(rt.get("material", "lambertian"), 0)
        color = tuple(f32(c) for c in rt.get("color", DEFAULT_COLOR))
        fuzz = f32(rt.get("fuzz", 0.0))
        pos, rot, scale = _vec(o["position"]), _quat(o["rotation"]), _vec(o["scale"])
        if o["mesh"] == "sphere":
            r = np.zeros((), dtype=OBJECT_DTYPE)
            r["kind"], r["material"], r["id"], r["fuzz"] = KIND_SPHERE, material, _id(oi, 0), fuzz
            r["color"] = color
            r["a"] = pos
            r["b"] = (f32(0.5) * scale[0], 0.0, 0.0)
            recs.append(r)
            continue
        P, T = mesh(o["mesh"])
        axes = quat_axes(rot)
        W = [transform_point(p, scale, axes, pos) for p in P]
        for ti, (i0, i1, i2) in enumerate(T):
            r = np.zeros((), dtype=OBJECT_DTYPE)
            r["kind"], r["material"], r["id"], r["fuzz"] = KIND_TRIANGLE, material, _id(oi, ti), fuzz
            r["color"] = color
            r["a"], r["b"], r["c"] = W[i0], W[i1], W[i2]
            recs.append(r)
    if not recs:
        return np.zeros((0,), dtype=OBJECT_DTYPE)
    return np.stack(recs)


def json_camera(scene: dict):
    """EXTENSION: JSON camera -> (look_from, look_at, v_up, v_fov).  forward = R*(0,0,1),
    up = R*(0,1,0), focus distance = |position| (1 if the camera sits at the origin)."""
    cam = scene["camera"]
    pos = _vec(cam["position"])
    _, ay, az = quat_axes(_quat(cam["rotation"]))
    focus = np.sqrt(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2], dtype=f32)
    if not focus > 0:
        focus = f32(1.0)
    at = tuple(pos[k] + focus * az[k] for k in range(3))
    return pos, at, ay, f32(cam["fov"])


def load(path_or_text: str) -> dict:
    if path_or_text.lstrip().startswith("{"):
        return json.loads(path_or_text)
    with open(path_or_text) as f:
        return json.load(f)
