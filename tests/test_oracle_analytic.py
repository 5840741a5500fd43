"""Analytic / known-answer tests that pin the oracle beyond the reference's Vec3 tests (the reference
has none for hit / scatter / camera / write-out: SURVEY §4).  Each cites the reference lines."""
import ctypes

import numpy as np
import pytest

from oracle import binding as ob
from oracle.scene_oracle import OBJECT_DTYPE

f32 = np.float32
FMAX = float(np.finfo(np.float32).max)


def sphere(c, r, mat=0, col=(0.5, 0.5, 0.5), fuzz=0.0, id=0):
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["kind"], o["material"], o["id"], o["fuzz"], o["color"], o["a"], o["b"] = 0, mat, id, fuzz, col, c, (r, 0, 0)
    return o


def plane(pos, ori, size, mat=0, col=(0.5, 0.5, 0.5), id=0):
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["kind"], o["material"], o["id"], o["color"], o["a"], o["b"], o["c"] = 1, mat, id, col, pos, ori, size
    return o


def tri(v0, v1, v2, id=0):
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["kind"], o["id"], o["color"], o["a"], o["b"], o["c"] = 2, id, (0.5, 0.5, 0.5), v0, v1, v2
    return o


# ---- Philox4x32-10: Random123's published known-answer vectors --------------------------------
@pytest.mark.parametrize("key,ctr,out", [
    ((0, 0), (0, 0, 0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF,) * 4, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0xA4093822, 0x299F31D0), (0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
])
def test_philox_kat(key, ctr, out):
    assert tuple(int(x) for x in ob.philox(key[0], key[1], ctr)) == out


# ---- Philox4x32-7, the round count the path draws from: Random123's published known-answer vectors (kat_vectors,
# "philox4x32 7 ...") for the same three inputs ------------------------------------------------------------------
@pytest.mark.parametrize("key,ctr,out", [
    ((0, 0), (0, 0, 0, 0), (0x5F6FB709, 0x0D893F64, 0x4F121F81, 0x4F730A48)),
    ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF,) * 4, (0x5207DDC2, 0x45165E59, 0x4D8EE751, 0x8C52F662)),
    ((0xA4093822, 0x299F31D0), (0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344),
     (0x4DFCCABA, 0x190A87F0, 0xC47362BA, 0xB6B5242A)),
])
def test_philox7_kat(key, ctr, out):
    assert ob.path_rounds() == 7
    assert tuple(int(x) for x in ob.philox(key[0], key[1], ctr, rounds=7)) == out


def test_unit_float_is_rand_standard_f32():
    # rand's Standard f32: (u32 >> 8) * 2^-24, in [0, 1)
    L = ob.lib()
    assert L.oracle_u32_to_f32(0) == 0.0
    assert L.oracle_u32_to_f32(0xFFFFFFFF) == float(f32(16777215) / f32(16777216))
    assert L.oracle_u32_to_f32(0x80000000) == 0.5
    assert L.oracle_u32_to_f32(0xFF) == 0.0


# ---- Sphere::hit (shapes/sphere.rs:23-51) ---------------------------------------------------------
def test_sphere_axis_hit_exact():
    hit, rec = ob.object_hit(sphere((0, 0, -3), 1.0), (0, 0, 0), (0, 0, -1))
    assert hit and rec[0] == 2.0 and rec[1:4].tolist() == [0, 0, -2] and rec[4:7].tolist() == [0, 0, 1]


def test_sphere_unnormalised_direction_scales_t():  # Q2: a = d.d stays in the formula
    hit, rec = ob.object_hit(sphere((0, 0, -3), 1.0), (0, 0, 0), (0, 0, -2))
    assert hit and rec[0] == 1.0


def test_sphere_inside_takes_second_root():  # sphere.rs:40-48
    hit, rec = ob.object_hit(sphere((0, 0, 0), 1.0), (0, 0, 0), (0, 0, -1))
    assert hit and rec[0] == 1.0 and rec[4:7].tolist() == [0, 0, -1]


def test_sphere_tangent_is_a_miss():  # discriminant > 0 is strict (Q3)
    hit, _ = ob.object_hit(sphere((0, 1, -3), 1.0), (0, 0, 0), (0, 0, -1))
    assert not hit


def test_sphere_respects_open_interval():  # root < t_max strict; root > t_min strict
    s = sphere((0, 0, -3), 1.0)
    assert not ob.object_hit(s, (0, 0, 0), (0, 0, -1), t_max=2.0)[0]   # root_one == t_max, root_two beyond
    assert ob.object_hit(s, (0, 0, 0), (0, 0, -1), t_max=2.0000002)[0]
    assert not ob.object_hit(s, (0, 0, 0), (0, 0, -1), t_min=4.0)[0]  # root_two == t_min


# ---- Plane::hit (shapes/plane.rs:24-44) -----------------------------------------------------------
def test_plane_zero_orientation_never_hits():  # scenes.rs:8-15: get_simple_scene's plane until rotate()
    assert not ob.object_hit(plane((-1, 0, 0), (0, 0, 0), (5, 5, 5)), (0, 0, 1), (0, 0, -1))[0]


def test_plane_is_one_sided_and_tests_denom_not_t():  # Q4
    p = plane((0, 0, -2), (0, 0, -1), (5, 5, 5))
    hit, rec = ob.object_hit(p, (0, 0, 0), (0, 0, -1))
    assert hit and rec[0] == 2.0 and rec[4:7].tolist() == [0, 0, 1]  # normal = -orientation
    assert not ob.object_hit(p, (0, 0, 0), (0, 0, 1))[0]             # denom < 0
    # t_max bounds denom (= 1 here), not t: a hit at t = 2 passes t_max = 1.5 but fails t_max = 1.0
    assert ob.object_hit(p, (0, 0, 0), (0, 0, -1), t_max=1.5)[0]
    assert not ob.object_hit(p, (0, 0, 0), (0, 0, -1), t_max=1.0)[0]
    # negative t is accepted: the plane is behind the origin, inside its extent
    hit, rec = ob.object_hit(plane((0, 0, 2), (0, 0, -1), (5, 5, 5)), (0, 0, 0), (0, 0, -1))
    assert hit and rec[0] == -2.0


def test_plane_clobbers_record_on_extent_reject():  # Q5: t, p written before the extent test
    p = plane((0, 0, -2), (0, 0, -1), (1, 1, 0.5))
    before = np.array([7, 1, 2, 3, 4, 5, 6], dtype=np.float32)
    hit, rec = ob.object_hit(p, (3, 0, 0), (0, 0, -1), rec=before)
    assert not hit and rec[0] == 2.0 and rec[1:4].tolist() == [3, 0, -2] and rec[4:7].tolist() == [4, 5, 6]


# ---- triangle extension ---------------------------------------------------------------------------
def test_triangle_double_sided_inclusive_edges():
    t = tri((-1, -1, -2), (1, -1, -2), (0, 1, -2))
    hit, rec = ob.object_hit(t, (0, 0, 0), (0, 0, -1))
    assert hit and rec[0] == 2.0 and rec[4:7].tolist() == [0, 0, 1]
    hit, rec = ob.object_hit(t, (0, 0, -4), (0, 0, 1))       # from behind: normal flips toward the ray
    assert hit and rec[0] == 2.0 and rec[4:7].tolist() == [0, 0, -1]
    assert ob.object_hit(t, (0, -1, 0), (0, 0, -1))[0]         # on the bottom edge
    assert ob.object_hit(t, (-1, -1, 0), (0, 0, -1))[0]        # on a vertex
    assert not ob.object_hit(t, (0, -1.0001, 0), (0, 0, -1))[0]
    assert not ob.object_hit(t, (0, 0, 0), (1, 0, 0))[0]       # parallel: det == 0
    assert not ob.object_hit(tri((0, 0, -2), (0, 0, -2), (0, 0, -2)), (0, 0, 0), (0, 0, -1))[0]  # degenerate


# ---- optics (utility.rs:27-54) ---------------------------------------------------------------------
def test_schlick_normal_incidence():
    r0 = f32((f32(1.0) - f32(1.3)) / (f32(1.0) + f32(1.3)))
    assert f32(ob.lib().oracle_schlick(1.0, 1.3)) == f32(r0 * r0)
    assert abs(ob.lib().oracle_schlick(1.0, 1.3) - 0.017013) < 1e-6
    assert f32(ob.lib().oracle_schlick(0.0, 1.3)) == f32(1.0)  # r0 + (1 - r0) * 1


def test_refract_normal_incidence_and_tir():
    out = (ctypes.c_float * 3)()
    ok = ob.lib().oracle_refract(ob._f3((0, 0, -2)), ob._f3((0, 0, 1)), 1.0 / 1.3, out)
    assert ok == 1 and out[0] == 0 and out[1] == 0 and out[2] == -1.0   # straight through, unit length
    # grazing exit from the dense side: total internal reflection
    v = (1.0, 0.0, -0.1)
    ok = ob.lib().oracle_refract(ob._f3(v), ob._f3((0, 0, 1)), 1.3, out)
    assert ok == 0


def test_reflect():
    out = (ctypes.c_float * 3)()
    ob.lib().oracle_reflect(ob._f3((1, -1, 0)), ob._f3((0, 1, 0)), out)
    assert list(out) == [1.0, 1.0, 0.0]


def test_as_u8_is_rust_saturating_cast():  # tracer.rs:74-76,181-183
    L = ob.lib()
    cases = {1.0 * 255.0: 255, 254.999: 254, 0.9999: 0, -0.1: 0, 300.0: 255, float("nan"): 0, float("inf"): 255,
             float("-inf"): 0, 128.5: 128}
    for f, want in cases.items():
        assert L.oracle_as_u8(f) == want, f


# ---- camera (camera.rs) ------------------------------------------------------------------------------
def test_camera_new_fields():  # camera.rs:24-60 at 600x600 (the reference's de-facto size)
    cam = ob.camera_new(600, 600)
    hh = f32(np.tan(f32(f32(60.0) * f32(3.14159265359) / f32(180.0)) / f32(2.0), dtype=f32))
    assert ob.cam_field(cam, "position").tolist() == [0, 0, 1]
    assert ob.cam_field(cam, "w").tolist() == [0, 0, 1] and ob.cam_field(cam, "u").tolist()[0] == 1.0
    np.testing.assert_allclose(ob.cam_field(cam, "lower_left_corner"), [-hh, -hh, 0.0], rtol=1e-6)
    np.testing.assert_allclose(ob.cam_field(cam, "horizontal"), [2 * hh, 0, 0], rtol=1e-6, atol=0)
    assert ob.cam_field(cam, "lens_radius")[0] == f32(0.05)
    # Q12: stored constants unrelated to the basis
    assert (ob.cam_field(cam, "focus_dist")[0], ob.cam_field(cam, "radius")[0], ob.cam_field(cam, "rotation")[0]) == (2.0, 5.0, 0.0)


def test_first_update_teleports_eye_to_radius_5():  # camera.rs:97-102 via tracer.rs:52 (Q12)
    cam = ob.camera_new(600, 600)
    ob.update_camera(cam, 0, 0.016)
    assert ob.cam_field(cam, "position").tolist() == [5.0, 0.0, 0.0]
    ob.update_camera(cam, 0b010000, 0.5)  # Q: +y
    assert ob.cam_field(cam, "position").tolist() == [5.0, 0.5, 0.0]
    ob.update_camera(cam, 0b000001, 0.25)  # S: radius += dt
    assert ob.cam_field(cam, "radius")[0] == 5.25


# ---- get_color / render (tracer.rs:160-219) -----------------------------------------------------------
def test_sky_straight_up_and_horizon():  # tracer.rs:213-217
    cam = ob.camera_look((0, 0, 0), (0, 1, 0), (0, 0, 1), 1.0, 1.0)
    _, lin, rays = ob.render(np.zeros(0, dtype=OBJECT_DTYPE), cam, 1, 1, 1, flags=ob.FLAG_NO_JITTER | ob.FLAG_PINHOLE)
    assert rays == 1
    np.testing.assert_allclose(lin[0, 0], [0.5, 0.7, 1.0], atol=2e-4)  # ~1 deg of tilt from the half-pixel offset
    assert lin[0, 0, 2] == 1.0


def test_metal_absorption_returns_exact_zero_and_depth_limit():  # Q7, Q9
    # a fully fuzzy metal sphere filling the view: some paths are absorbed -> exact 0 contributions
    objs = np.stack([sphere((0, 0, -1), 0.9, mat=1, col=(1, 1, 1), fuzz=1.0)])
    cam = ob.camera_new(8, 8)
    _, lin, _ = ob.render(objs, cam, 8, 8, 1, seed=5)
    assert (lin == 0.0).any()
    # max_depth = 0: any hit returns black (tracer.rs:205), misses return sky
    _, lin0, rays0 = ob.render(objs, cam, 8, 8, 1, max_depth=0, seed=5)
    assert rays0 == 64 and set(np.unique(lin0[3:5, 3:5])) == {0.0}


def test_render_is_thread_count_and_band_invariant():
    objs, cam = ob.builtin_scene(2), ob.camera_new(40, 24)
    rgb1, lin1, r1 = ob.render(objs, cam, 40, 24, 3, seed=11, nthreads=1)
    rgb8, lin8, r8 = ob.render(objs, cam, 40, 24, 3, seed=11, nthreads=5)
    assert r1 == r8 and np.array_equal(lin1.view(np.uint32), lin8.view(np.uint32)) and np.array_equal(rgb1, rgb8)
    _, top, _ = ob.render(objs, cam, 40, 24, 3, seed=11, y0=0, y1=10)
    assert np.array_equal(top[:10].view(np.uint32), lin1[:10].view(np.uint32)) and not top[10:].any()


def test_sample_offset_composes():
    objs, cam = ob.builtin_scene(0), ob.camera_new(16, 16)
    _, whole, _ = ob.render(objs, cam, 16, 16, 4, seed=2)
    _, a, _ = ob.render(objs, cam, 16, 16, 2, seed=2)
    _, b, _ = ob.render(objs, cam, 16, 16, 2, sample_offset=2, seed=2)
    np.testing.assert_allclose(a + b, whole, rtol=1e-6, atol=1e-6)


def test_sequential_rng_mode_agrees_statistically():
    # mode 1 (thread_rng stand-in, Q14 order) vs the counter RNG: same distribution, different streams
    objs, cam = ob.builtin_scene(2), ob.camera_new(24, 24)
    _, a, _ = ob.render(objs, cam, 24, 24, 64, seed=1)
    _, b, _ = ob.render(objs, cam, 24, 24, 64, seed=1, flags=ob.FLAG_SEQ_RNG)
    assert not np.array_equal(a, b)
    assert abs(float(a.mean() - b.mean())) / 64 < 0.02


def test_oracle_output_pins():
    """tests/golden/oracle_pins.json (tools/make_oracle_pins.py): the oracle's images, hit ids, hit distances and ray
    counts on six fixed cases.  Not reference outputs - a guard against the oracle drifting under the GPU suite."""
    import importlib.util, json, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("pins", os.path.join(root, "tools", "make_oracle_pins.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    pins = json.load(open(os.path.join(root, "tests", "golden", "oracle_pins.json")))
    assert len(pins) == len(mod.CASES)
    for n, w, h, spp, seed in mod.CASES:
        assert mod.run(n, w, h, spp, seed) == pins[f"{n} {w}x{h} {spp}spp seed{seed}"], n
