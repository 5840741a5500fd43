"""The oracle's checker-side accelerator (FLAG_BVH, forma_oracle.cpp OBvh) against the oracle's own brute-force scan
(tracer.rs:195-200 restated): f32 sums, RGB8 bytes and ray counts bit for bit.  Only after this may the accelerated
oracle stand in for the scan at sizes the scan cannot finish (config D at 1920x1080, tests/test_gpu_lbvh.py)."""
import importlib.util
import os

import numpy as np
import pytest

from oracle import binding as ob
from oracle import scene_oracle as so

from conftest import ROOT, load_scene_pair


def generator():
    spec = importlib.util.spec_from_file_location("sphere_scene_generator",
                                                  os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def both(objs, cam, w, h, spp, seed, **kw):
    n = os.cpu_count() or 1
    a = ob.render(objs, cam, w, h, spp, seed=seed, nthreads=n, **kw)
    b = ob.render(objs, cam, w, h, spp, seed=seed, nthreads=n, flags=ob.FLAG_BVH, **kw)
    return a, b


def same(a, b):
    assert a[2] == b[2], "ray counts differ"
    assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), "f32 sums differ"
    assert np.array_equal(a[0], b[0])


def test_accelerated_oracle_equals_scan_on_sphere_field():
    """2 000 spheres of config D's generator + the radius-1000 ground sphere, all three materials, depth 50."""
    d = generator().generate(2000, seed=99)
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    w, h = 192, 108
    a, b = both(objs, ob.camera_look(fr, at, up, fov, w / h), w, h, 3, 21)
    same(a, b)
    assert a[2] > w * h * 3 * 1.5  # the frame bounces


@pytest.mark.parametrize("name", ["scene_01", "scene_05"])
def test_accelerated_oracle_equals_scan_on_meshes(name):
    w, h = 96, 54
    _, objs, cam = load_scene_pair(name, w, h)
    a, b = both(objs, cam, w, h, 2, 5)
    same(a, b)


def test_accelerated_oracle_equals_scan_on_adversarial_spheres():
    """Tiny and negative-radius spheres far from the eye, touching spheres with equal t (ties go to the earlier list
    entry), a sphere the eye sits inside."""
    from test_oracle_analytic import sphere
    rng = np.random.default_rng(7)
    objs = [sphere((0, 0, 0), 30.0, mat=2), sphere((0, 0, -4), 1.0), sphere((0, 0, -4), 1.0, mat=1)]
    for _ in range(300):
        c = rng.uniform(-20, 20, 3)
        objs.append(sphere(tuple(c), float(rng.choice([1e-3, 0.05, -0.3, 0.7])), mat=int(rng.integers(0, 4))))
    objs = np.stack(objs)
    w, h = 128, 72
    cam = ob.camera_look((0.5, 1.0, 9.0), (0, 0, 0), (0, 1, 0), 60.0, w / h)
    a, b = both(objs, cam, w, h, 2, 3)
    same(a, b)


def test_lists_with_planes_fall_back_to_the_scan():
    objs = ob.builtin_scene(0)  # a plane first: order-sensitive (Q4 / Q5), never accelerated
    w, h = 64, 36
    a, b = both(objs, ob.camera_new(w, h), w, h, 2, 1)
    same(a, b)
