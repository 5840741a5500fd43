"""-m gpu: forma_rt_render_device_noalloc (static device scratch, no allocation) against forma_rt_render_device.

The entry point has no counterpart in the reference.  It must
give what the existing entry point gives - the tolerance of this suite is bit-exact f32 sums, identical RGB8 bytes and
the same ray count - on two calls in a row (the scratch array keeps the previous call's contents), for a scanned scene
(render_queue_kernel) and for a scene with an LBVH (render_tree_kernel), alone in a fresh process or after the other
tests.  This file is named so that pytest collects it last."""
import os

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib

from conftest import ROOT

pytestmark = pytest.mark.gpu


def bits(t):
    return t.cpu().numpy().view(np.uint32)


def frames(m, spp, offset, call):
    """One frame through `call` into fresh torch buffers; returns (rgb bytes, f32 sum bits, rays)."""
    import torch
    dev = torch.device("cuda", 0)
    npix = m.width * m.height
    rgb = torch.full((npix * 3,), 7, dtype=torch.uint8, device=dev)
    lin = torch.full((npix, 4), -1.0, dtype=torch.float32, device=dev)
    call(spp, offset, 50, rgb.data_ptr(), lin.data_ptr(), 0)
    st = m.stats()  # synchronises the handle's stream, reads the ray counter back
    return rgb.cpu().numpy(), bits(lin), int(st.rays)


def check_pair(m, spp, offset=0):
    want = frames(m, spp, offset, m.render_device)
    for _ in range(2):  # two calls in a row share the static scratch
        got = frames(m, spp, offset, m.render_device_noalloc)
        assert got[2] == want[2]
        assert np.array_equal(got[1], want[1]), "f32 sums differ from forma_rt_render_device"
        assert np.array_equal(got[0], want[0])
    return want


def test_noalloc_scanned_scene_matches_render_device_twice():
    m = forma_b200.create_model(320, 200, device=0, seed=17)
    m.scene.load_builtin(2)  # get_objects: lambertian + metal + dielectric, scanned
    m.scene.camera.orbit((0.0, 0.0, 0.0))
    a = check_pair(m, 5)
    b = check_pair(m, 3, offset=9)  # another sample range right after: nothing stale may leak in
    assert not np.array_equal(a[1], b[1])
    m.close()


def test_noalloc_lbvh_scene_and_chunked_samples():
    w, h = 512, 512  # = FORMA_RT_NOALLOC_MAX_PIXELS: one sample per chunk, the cross-chunk sum lives in the scratch too
    assert w * h == _lib.NOALLOC_MAX_PIXELS
    m = forma_b200.create_model(w, h, device=0, seed=4)
    m.scene.load_json(open(os.path.join(ROOT, "tests", "golden", "scenes", "scene_01.json")).read())
    m.scene.camera.use_json()
    want = check_pair(m, 3)
    assert m.stats().accel_in_use == _lib.ACCEL_LBVH and want[2] > w * h * 3
    m.close()


def test_noalloc_rejects_what_it_cannot_do():
    m = forma_b200.create_model(1920, 1080, device=0, seed=1)
    m.scene.load_builtin(0)
    with pytest.raises(_lib.FormaError) as e:
        m.render_device_noalloc(1, 0, 50, 1, None)
    assert e.value.code == _lib.EINVAL and "FORMA_RT_NOALLOC_MAX_PIXELS" in str(e.value)
    m.close()
    m = forma_b200.create_model(64, 64, device=0, seed=1)
    m.scene.load_builtin(0)
    with pytest.raises(_lib.FormaError) as e:  # nothing has uploaded the scene yet
        m.render_device_noalloc(1, 0, 50, 1, None)
    assert e.value.code == _lib.EINVAL and "not resident" in str(e.value)
    m.render(1)
    m.scene.load_builtin(2)  # a scene change makes it non-resident again
    with pytest.raises(_lib.FormaError):
        m.render_device_noalloc(1, 0, 50, 1, None)
    m.close()


def test_noalloc_honours_the_srgb_output_flag():
    m = forma_b200.create_model(160, 96, device=0, seed=5, flags=_lib.SRGB_OUT)
    m.scene.load_builtin(0)
    m.scene.camera.orbit((0.0, 0.0, 0.0))
    check_pair(m, 4)
    m.close()
