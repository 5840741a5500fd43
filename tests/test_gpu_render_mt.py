"""GPU parity of tracer::render_mt / save_image_mt (tracer.rs:83-158) against the oracle's
restatement: bytes of every frame and of the post-gamma average, ray counts, and the quirks - a fresh
get_simple_scene() whatever the model holds, bands in descending order, rows past 4 * (h / 4) absent / zero."""
import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from forma_b200.tracer import FormaError
from oracle import binding as ob

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h", [(64, 48), (50, 30), (33, 7), (160, 101), (9, 4)])
def test_render_mt_frames_match_the_oracle(w, h):
    m = forma_b200.create_model(w, h, seed=3)
    cam = ob.camera_new(w, h)
    for frame in (0, 1, 7):
        got = m.render_mt(frame)
        want = ob.render_mt(cam, w, h, frame, seed=3)
        assert got.shape == (4 * (h // 4), w, 3)
        assert np.array_equal(got, want), (w, h, frame, int((got != want).any(-1).sum()))
    st = m.stats()
    assert st.gpu_launches == 1 and st.paths == w * 4 * (h // 4) and st.rays >= st.paths


def test_render_mt_ignores_the_models_scene_and_follows_its_camera():
    w, h = 96, 64
    m = forma_b200.create_model(w, h, seed=9)
    a = m.render_mt(2)
    m.scene.load_builtin(2)                       # the model now holds get_objects(): render_mt does not care (tracer.rs:105)
    m.scene.objects[0].translate((3.0, 0.0, 0.0))
    assert np.array_equal(m.render_mt(2), a)
    forma_b200.update(m, 0b1010, 0.05)            # first update: the eye jumps to radius 5 (Q12), then orbits
    cam = ob.camera_new(w, h)
    ob.update_camera(cam, 0b1010, 0.05)
    b = m.render_mt(2)
    assert not np.array_equal(a, b)
    assert np.array_equal(b, ob.render_mt(cam, w, h, 2, seed=9))
    # the ordinary frame of the same model still shows ITS scene
    rgb, _ = m.render(2)
    want, _, _ = ob.render(ob.builtin_scene(2), cam, w, h, 2, seed=9)
    # objects[0] of get_objects() is a sphere: translate is a no-op on it (hitable.rs:12-13)
    assert np.array_equal(rgb, want)


@pytest.mark.parametrize("w,h,sample", [(64, 48, 5), (50, 30, 3), (40, 3, 2)])
def test_save_image_mt_average(w, h, sample, tmp_path):
    from PIL import Image
    m = forma_b200.create_model(w, h, seed=4)
    cam = ob.camera_new(w, h)
    got = m.render_mt_average(sample)
    want = ob.save_image_mt(cam, w, h, sample, seed=4)
    assert np.array_equal(got, want)
    assert not got[4 * (h // 4):].any()           # rows render_mt never produces stay black
    path = tmp_path / "basic_mt.png"
    forma_b200.save_image_mt(m, sample, str(path))
    assert np.array_equal(np.asarray(Image.open(path).convert("RGB")), want)
    with pytest.raises(FormaError) as e:
        forma_b200.save_image_mt(m, sample, str(tmp_path / "missing_dir" / "x.png"))  # tracer.rs:157 unwrap
    assert e.value.code == _lib.EIO
    with pytest.raises(FormaError):
        m.render_mt_average(0)
