"""-m gpu: the guards added in round 2 around the device-resident entry points.

  * a flag wait never spins for ever: it gives up after FORMA_RT_FLAG_TIMEOUT_MS and the handle reports which flag;
  * frames stored through a caller's device pointer that is not 4-byte aligned take the byte path (no misaligned
    32-bit stores) in every write-out kernel: render_kernel's staged tile, accumulate_kernel, tonemap_kernel;
  * forma_rt_host_alloc gives a pinned `pixels` buffer that behaves like any other host buffer."""
import ctypes
import os

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_flag_wait_gives_up_and_names_the_flag(monkeypatch):
    monkeypatch.setenv("FORMA_RT_FLAG_TIMEOUT_MS", "40")  # read when the handle is created
    m = forma_b200.create_model(64, 64, device=0, seed=1)
    p = ctypes.c_void_p()
    _lib.check(_lib.lib().forma_rt_device_alloc(m._h, 256, ctypes.byref(p)))  # zero-filled: flags at +0 and +128
    m.flag_set_device(p.value, 3)
    m.flag_wait_device([p.value], 3)          # already raised: returns at once
    m.synchronize()
    assert m.flag_error() == 0
    m.flag_wait_device([p.value, p.value + 128], 3)  # the second flag is never raised
    with pytest.raises(_lib.FormaError) as e:
        m.synchronize()
    assert e.value.code == _lib.ECUDA and "flag 1" in str(e.value) and "40 ms" in str(e.value)
    with pytest.raises(_lib.FormaError):      # the handle keeps failing instead of queueing more waits
        m.flag_wait_device([p.value], 3)
    m.close()


@pytest.mark.parametrize("scene", ["sky", "bounce", "tree"])
def test_misaligned_frame_pointer_takes_the_byte_path(scene):
    """render_kernel (sky frame), render_queue_kernel + accumulate_kernel (bouncing built-in scene) and
    render_tree_kernel + accumulate_kernel (LBVH scene), then tonemap_kernel: a frame pointer at byte offset 1, 2, 3
    gives the bytes of the aligned frame."""
    import torch
    w, h, spp = 128, 64, 4
    m = forma_b200.create_model(w, h, device=0, seed=6)
    name = {"sky": "scene_09", "tree": "scene_01"}.get(scene)
    if name:
        m.scene.load_json(open(os.path.join(ROOT, "tests", "golden", "scenes", name + ".json")).read())
        m.scene.camera.use_json()
    else:
        m.scene.load_builtin(2)
    dev = torch.device("cuda", 0)
    n = w * h * 3
    want, lin = m.render(spp, want_linear=True)
    lin_d = torch.from_numpy(np.ascontiguousarray(lin)).to(dev)
    for off in (0, 1, 2, 3):
        buf = torch.zeros(n + 8, dtype=torch.uint8, device=dev)
        m.render_device(spp, 0, 50, buf.data_ptr() + off, None, 0)
        m.synchronize()
        got = buf.cpu().numpy()
        assert np.array_equal(got[off:off + n].reshape(h, w, 3), want), (scene, off)
        assert not got[:off].any() and not got[off + n:].any()
        buf.zero_()
        m.tonemap_device(lin_d.data_ptr(), spp, buf.data_ptr() + off, 0, w * h, 0)
        m.synchronize()
        got = buf.cpu().numpy()
        assert np.array_equal(got[off:off + n].reshape(h, w, 3), want), ("tonemap", scene, off)
        assert not got[:off].any() and not got[off + n:].any()
    m.close()


def test_tonemap_ragged_pixel_ranges():
    """tonemap_kernel owns aligned groups of four pixels: ranges that start and end inside a group write exactly
    their own pixels."""
    import torch
    w, h, spp = 96, 40, 3
    m = forma_b200.create_model(w, h, device=0, seed=2)
    m.scene.load_builtin(0)
    want, lin = m.render(spp, want_linear=True)
    dev = torch.device("cuda", 0)
    lin_d = torch.from_numpy(np.ascontiguousarray(lin)).to(dev)
    for begin, end in ((0, w * h), (1, 2), (5, 6), (3, 4 * 50 + 1), (w * h - 7, w * h), (10, 10)):
        buf = torch.full((w * h * 3,), 9, dtype=torch.uint8, device=dev)
        m.tonemap_device(lin_d.data_ptr(), spp, buf.data_ptr(), begin, end, 0)
        m.synchronize()
        got = buf.cpu().numpy().reshape(-1, 3)
        assert np.array_equal(got[begin:end], want.reshape(-1, 3)[begin:end])
        assert (got[:begin] == 9).all() and (got[end:] == 9).all()
    m.close()


def test_pinned_pixels_buffer():
    w, h = 160, 90
    m = forma_b200.create_model(w, h, device=0, seed=3)
    m.scene.load_builtin(2)
    want, _ = m.render(5)
    buf = m.host_alloc(w * h * 3)
    got, _ = m.render(5, rgb_out=buf.reshape(h, w, 3))
    assert np.array_equal(np.asarray(got), want) and np.array_equal(buf.reshape(h, w, 3), want)
    m.host_free(buf)
    with pytest.raises(_lib.FormaError):
        m.host_free(buf)  # not owned any more
    m.close()
