"""The reference's own 7 Vec3 known-answer tests (src/cpu_ray_tracer/primitives.rs:163-254), run
against the oracle's Vec3.  These are the only golden vectors the reference holds for the path."""
import ctypes

import numpy as np

from oracle import binding as ob

f32 = np.float32


def _f3(v):
    return (ctypes.c_float * 3)(*v)


def test_length():  # primitives.rs:163-168
    a = (2.0, -3.0, -1.2)
    assert f32(ob.lib().oracle_vec3_length_squared(_f3(a))) == f32(14.440001)
    assert f32(ob.lib().oracle_vec3_length(_f3(a))) == f32(3.8)


def test_dot():  # primitives.rs:170-179
    a, b = (1.0, 1.0, 1.0), (2.0, -3.0, -0.2)
    dot = lambda x, y: f32(ob.lib().oracle_vec3_dot(_f3(x), _f3(y)))
    assert dot(a, a) == f32(3.0)
    assert dot(b, b) == f32(13.04)
    assert dot(a, b) == f32(-1.2)
    assert dot(b, a) == f32(-1.2)
    assert dot(a, b) == dot(b, a)


def test_cross():  # primitives.rs:181-189
    a, b = (1.0, 2.0, 3.0), (-3.0, -2.0, 1.0)
    assert ob.vec3_op("cross", a, a)[0] == 0.0
    assert ob.vec3_op("cross", b, b)[1] == 0.0
    assert ob.vec3_op("cross", a, b)[0] == 8.0
    assert ob.vec3_op("cross", b, a)[2] == -4.0


def test_addition():  # primitives.rs:191-204
    a, b = (1.0, 2.0, 3.0), (-3.0, -2.0, 1.0)
    assert ob.vec3_op("add", a, b).tolist() == [-2.0, 0.0, 4.0]
    assert ob.vec3_op("add", a, b).tolist() == ob.vec3_op("add", b, a).tolist()


def test_subtraction():  # primitives.rs:206-219
    a, b = (1.0, 2.0, 3.0), (-3.0, -2.0, 1.0)
    assert ob.vec3_op("sub", a, b).tolist() == [4.0, 4.0, 2.0]
    out = (ctypes.c_float * 3)()
    ob.lib().oracle_vec3_lscale(-1.0, _f3(ob.vec3_op("sub", b, a)), out)
    assert ob.vec3_op("sub", a, b).tolist() == list(out)


def test_multiplication():  # primitives.rs:221-241
    a, b = (1.0, 2.0, 3.0), (-3.0, -2.0, 1.0)
    out = (ctypes.c_float * 3)()
    ob.lib().oracle_vec3_scale(_f3(a), 3.0, out)
    assert list(out) == [3.0, 6.0, 9.0]
    assert ob.vec3_op("mul", a, b).tolist() == [-3.0, -4.0, 3.0]


def test_division():  # primitives.rs:243-254
    out = (ctypes.c_float * 3)()
    ob.lib().oracle_vec3_div(_f3((1.0, 2.0, 3.0)), 2.0, out)
    assert list(out) == [0.5, 1.0, 1.5]
