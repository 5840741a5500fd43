"""-m gpu: the LBVH walk must return exactly what the linear scan returns (same ids, same t bits,
same images), on the oracle at sizes it can finish and against the GPU's own brute-force scan at
BASELINE config D's size (100 000 spheres + ground)."""
import importlib.util
import json
import os

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from oracle import binding as ob
from oracle import scene_oracle as so

from conftest import ROOT, load_scene_pair

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def generator():
    spec = importlib.util.spec_from_file_location("sphere_scene_generator",
                                                  os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def model(txt, w, h, accel, seed=0):
    m = forma_b200.create_model(w, h, seed=seed, accel=accel)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    return m


@pytest.mark.parametrize("name", ["scene_01", "scene_05", "scene_06", "scene_09"])
def test_lbvh_equals_scan_on_reference_scenes(name):
    w, h = 480, 270
    txt, objs, cam = load_scene_pair(name, w, h)
    mb, ms = model(txt, w, h, _lib.ACCEL_LBVH, 8), model(txt, w, h, _lib.ACCEL_BRUTE, 8)
    ids_b, t_b = mb.primary_hits()
    ids_s, t_s = ms.primary_hits()
    assert mb.stats().accel_in_use == _lib.ACCEL_LBVH and ms.stats().accel_in_use == _lib.ACCEL_BRUTE
    assert mb.stats().bvh_nodes > 0
    assert np.array_equal(ids_b, ids_s) and np.array_equal(bits(t_b), bits(t_s))
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    assert np.array_equal(ids_b, ids_o) and np.array_equal(bits(t_b), bits(t_o))
    rb, lb = mb.render(2, want_linear=True)
    rs, ls = ms.render(2, want_linear=True)
    assert mb.stats().rays == ms.stats().rays
    assert np.array_equal(bits(lb), bits(ls)) and np.array_equal(rb, rs)


def test_lbvh_sphere_field_vs_oracle():
    """2 000 spheres + the radius-1000 ground (a "big" primitive kept out of the tree): oracle parity."""
    d = generator().generate(2000, seed=99)
    txt = json.dumps(d)
    w, h = 256, 144
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    cam = ob.camera_look(fr, at, up, fov, w / h)
    m = model(txt, w, h, _lib.ACCEL_AUTO, 21)
    st0 = m.stats()
    ids, t = m.primary_hits()
    assert m.stats().accel_in_use == _lib.ACCEL_LBVH and m.stats().bvh_nodes == 2000 - 1
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    assert np.array_equal(ids, ids_o) and np.array_equal(bits(t), bits(t_o))
    rgb, lin = m.render(2, want_linear=True)
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, 2, seed=21, nthreads=os.cpu_count())
    assert m.stats().rays == rays_o
    assert np.array_equal(bits(lin[..., :3]), bits(lin_o)) and np.array_equal(rgb, rgb_o)


def test_config_d_100k_lbvh_equals_gpu_scan():
    """BASELINE config D at full scene size: 100 000 spheres + ground, 1920x1080 primary rays, and a
    1-spp depth-50 frame at 960x540 — LBVH vs the GPU's own linear scan, bit for bit."""
    txt = json.dumps(generator().generate(100_000, seed=1234))
    w, h = 1920, 1080
    mb, ms = model(txt, w, h, _lib.ACCEL_LBVH, 5), model(txt, w, h, _lib.ACCEL_BRUTE, 5)
    ids_b, t_b = mb.primary_hits()
    ids_s, t_s = ms.primary_hits()
    assert mb.stats().bvh_nodes == 100_000 - 1
    assert np.array_equal(ids_b, ids_s) and np.array_equal(bits(t_b), bits(t_s))
    assert (ids_b >= 0).mean() > 0.5 and len(np.unique(ids_b)) > 5000
    w, h = 480, 270
    mb, ms = model(txt, w, h, _lib.ACCEL_LBVH, 5), model(txt, w, h, _lib.ACCEL_BRUTE, 5)
    rb, lb = mb.render(1, want_linear=True)
    rs, ls = ms.render(1, want_linear=True)
    assert mb.stats().rays == ms.stats().rays > w * h * 2
    assert np.array_equal(bits(lb), bits(ls)) and np.array_equal(rb, rs)


def test_lbvh_rejects_interleaved_planes():
    from test_oracle_analytic import plane, sphere
    objs = np.stack([sphere((0, 0, -3), 1.0), plane((0, 0, -9), (0, 0, -1), (5, 5, 5)), sphere((1, 0, -3), 1.0),
                     sphere((2, 0, -3), 1.0), sphere((3, 0, -3), 1.0), sphere((4, 0, -3), 1.0)])
    m = forma_b200.create_model(32, 32, accel=_lib.ACCEL_LBVH)
    m.scene.load_objects(objs)
    with pytest.raises(forma_b200.tracer.FormaError):
        m.primary_hits()
    m2 = forma_b200.create_model(32, 32, accel=_lib.ACCEL_AUTO)  # AUTO falls back to the scan
    m2.scene.load_objects(objs)
    ids, _ = m2.primary_hits()
    assert m2.stats().accel_in_use == _lib.ACCEL_BRUTE
    assert np.array_equal(ids, ob.primary_hits(objs, ob.camera_new(32, 32), 32, 32)[0])


def test_walk_stats_instrumented_frame():
    """forma_rt_walk_stats_get: the instrumented instantiation walks the same tree for the same rays."""
    rng = np.random.default_rng(3)
    n = 5000
    recs = np.zeros(n, dtype=forma_b200.tracer.OBJECT_DTYPE)
    recs["kind"] = _lib.SPHERE
    recs["material"] = rng.integers(0, 3, n)
    recs["id"] = np.arange(n)
    recs["fuzz"] = rng.uniform(0, 0.5, n)
    recs["color"] = rng.uniform(0.1, 0.9, (n, 3))
    recs["a"] = np.stack([rng.uniform(-6, 6, n), rng.uniform(-3, 3, n), rng.uniform(-9, -2, n)], 1)
    recs["b"][:, 0] = rng.uniform(0.05, 0.2, n)
    w, h, spp = 160, 90, 3
    m = forma_b200.create_model(w, h, seed=2, accel=_lib.ACCEL_LBVH)
    m.scene.load_objects(recs)
    rgb, lin = m.render(spp, want_linear=True)
    rays = m.stats().rays
    ws = m.walk_stats(spp)
    assert ws.rays == rays and ws.bvh_nodes == n - 1 and ws.node_bytes == 32 and ws.leaf_bytes == 16
    assert ws.node_visits > ws.leaf_tests > 0
    # every ray that is not culled by the scene's bounding sphere visits at least the root
    assert ws.node_visits <= rays * 2 * 64 and ws.node_visits >= 0.2 * rays
    assert 500.0 < ws.l2_read_gbs < 40000.0
    # the instrumented frame leaves the product path untouched: the next render is the same frame again
    rgb2, lin2 = m.render(spp, want_linear=True)
    assert np.array_equal(rgb, rgb2) and np.array_equal(lin.view(np.uint32), lin2.view(np.uint32))
    assert m.stats().rays == rays
    # a scanned scene has no walk
    b = forma_b200.create_model(64, 48, seed=1, accel=_lib.ACCEL_BRUTE)
    b.scene.load_builtin(2)
    ws = b.walk_stats(2)
    assert ws.bvh_nodes == 0 and ws.node_visits == 0 and ws.leaf_tests == 0 and ws.rays > 0


def test_config_d_fullsize_band_vs_accelerated_oracle():
    """BASELINE config D as benched - 100 000 spheres + ground, 1920x1080, 64 spp, depth 50 - against the ORACLE (its
    checker-side accelerator, pinned to its own scan by tests/test_oracle_bvh.py) on a two-row band: f32 sums and
    RGB8 bytes bit for bit."""
    d = generator().generate(100_000, seed=1234)
    w, h, spp = 1920, 1080, 64
    m = model(json.dumps(d), w, h, _lib.ACCEL_AUTO, 0)
    rgb, lin = m.render(spp, want_linear=True)
    assert m.stats().accel_in_use == _lib.ACCEL_LBVH and m.stats().bvh_nodes == 100_000 - 1
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    cam = ob.camera_look(fr, at, up, fov, w / h)
    y0, y1 = 640, 642
    rgb_o, lin_o, _ = ob.render(objs, cam, w, h, spp, seed=0, nthreads=os.cpu_count(), flags=ob.FLAG_BVH, y0=y0, y1=y1)
    assert np.array_equal(bits(lin[y0:y1, :, :3]), bits(lin_o[y0:y1]))
    assert np.array_equal(rgb[y0:y1], rgb_o[y0:y1])
    assert len(np.unique(rgb[y0:y1].reshape(-1, 3), axis=0)) > 200  # the band looks at the sphere field


def test_lbvh_negative_and_zero_radius_spheres():
    """Sphere::hit only uses radius * radius (sphere.rs:27): r < 0 hits like |r| (the hollow-sphere idiom), r = 0 never
    has disc > 0 but must not poison the walk's pad.  60 spheres: ACCEL_AUTO builds the tree (spheres + 2 x triangles >= 32)."""
    from test_oracle_analytic import sphere
    rng = np.random.default_rng(11)
    objs = []
    for i in range(60):
        c = rng.uniform(-6, 6, 3) + np.array([0, 0, -12.0])
        r = float(rng.choice([-0.9, -0.4, 0.0, 0.5, 1.1]))
        objs.append(sphere(tuple(c), r, mat=int(rng.integers(0, 4)), col=tuple(rng.uniform(0.2, 0.9, 3)), fuzz=0.1, id=i))
    objs = np.stack(objs)
    w, h = 256, 144
    cam = ob.camera_look((0, 0, 3), (0, 0, -12), (0, 1, 0), 60.0, w / h)
    out = []
    for accel in (_lib.ACCEL_AUTO, _lib.ACCEL_BRUTE):
        m = forma_b200.create_model(w, h, seed=9, accel=accel)
        m.scene.load_objects(objs)
        m.scene.camera.set_array(cam)
        ids, t = m.primary_hits()
        rgb, lin = m.render(3, want_linear=True)
        out.append((ids, bits(t), bits(lin), rgb, m.stats().rays, m.stats().accel_in_use))
    assert out[0][5] == _lib.ACCEL_LBVH and out[1][5] == _lib.ACCEL_BRUTE
    for a, b in zip(out[0][:5], out[1][:5]):
        assert np.array_equal(a, b)
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, 3, seed=9, nthreads=os.cpu_count())
    assert np.array_equal(out[0][0], ids_o) and np.array_equal(out[0][1], bits(t_o))
    assert out[0][4] == rays_o and np.array_equal(out[0][2][..., :3], bits(lin_o)) and np.array_equal(out[0][3], rgb_o)
    assert (ids_o >= 0).mean() > 0.1


@pytest.mark.parametrize("seed", range(12))
def test_culls_and_pads_never_drop_a_hit(seed):
    """Randomized differential test of scene_culled, beam_culled and the walk's pads (hand-derived rounding margins):
    tiny spheres and thin triangles sitting at the corners of the scene's bounding box, cameras 1 to 100 scene radii
    away, with and without the tree - against the oracle's plain scan, which has no cull at all."""
    from test_oracle_analytic import sphere, tri
    rng = np.random.default_rng(1000 + seed)
    R = float(rng.choice([0.5, 5.0, 50.0]))
    objs, k = [], 0
    for sx in (-1, 1):
        for sy in (-1, 1):
            for sz in (-1, 1):
                c = np.array([sx, sy, sz]) * R
                if rng.random() < 0.5:
                    objs.append(sphere(tuple(c), float(rng.choice([1e-3, 1e-2, 0.05])) * R, mat=int(rng.integers(0, 3)), id=k))
                else:
                    e = rng.uniform(-0.02, 0.02, (2, 3)) * R
                    objs.append(tri(tuple(c), tuple(c + e[0]), tuple(c + e[1]), id=k))
                k += 1
    for _ in range(int(rng.choice([0, 30]))):  # with 30 more spheres ACCEL_AUTO builds the tree (spheres + 2 x triangles >= 32)
        objs.append(sphere(tuple(rng.uniform(-0.8, 0.8, 3) * R), float(rng.uniform(0.01, 0.2)) * R, mat=int(rng.integers(0, 3)), id=k))
        k += 1
    objs = np.stack(objs)
    w, h = 96, 54
    dist = float(rng.choice([1.0, 3.0, 30.0, 100.0])) * R * np.sqrt(3.0)
    u = rng.normal(size=3); u /= np.linalg.norm(u)
    at = rng.choice([-1.0, 1.0], 3) * R  # look at a corner
    cam = ob.camera_look(tuple(at + u * dist), tuple(at), (0, 1, 0), float(rng.choice([10.0, 60.0])), w / h)
    m = forma_b200.create_model(w, h, seed=seed, accel=_lib.ACCEL_AUTO)
    m.scene.load_objects(objs)
    m.scene.camera.set_array(cam)
    ids, t = m.primary_hits()
    rgb, lin = m.render(4, want_linear=True)
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, 4, seed=seed, nthreads=os.cpu_count())
    assert np.array_equal(ids, ids_o) and np.array_equal(bits(t), bits(t_o))
    assert m.stats().rays == rays_o and np.array_equal(bits(lin[..., :3]), bits(lin_o)) and np.array_equal(rgb, rgb_o)
