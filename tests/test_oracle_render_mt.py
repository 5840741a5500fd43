"""CPU tests of the oracle's tracer::render_mt / save_image_mt restatement (tracer.rs:83-158)."""
import numpy as np

from oracle import binding as ob


def test_bands_reproduce_the_single_threaded_mapping_when_height_divides_by_four():
    # v = ((t_height - y) + xi) / height + t_id / 4 is the same f32 as ((height - r) + xi) / height here, so with the
    # counter RNG keyed by the output pixel the banded frame equals render's 1-spp frame of get_simple_scene()
    w, h = 64, 48
    cam = ob.camera_new(w, h)
    for frame in (0, 5):
        mt = ob.render_mt(cam, w, h, frame, seed=3)
        rgb, _, _ = ob.render(ob.builtin_scene(0), cam, w, h, 1, sample_offset=frame, seed=3, want_linear=False)
        assert mt.shape == (h, w, 3) and np.array_equal(mt, rgb)


def test_remainder_rows_are_dropped_and_bands_descend():
    w, h = 40, 30                      # t_height = 7: 28 rows, a squeezed image (each band spans 7/30 of v, offset t_id/4)
    cam = ob.camera_new(w, h)
    mt = ob.render_mt(cam, w, h, 0, seed=1)
    assert mt.shape == (28, w, 3)
    # sky on top (bright, blue-ish), ground sphere at the bottom: the FIRST rows belong to the TOP band (t_id = 3)
    assert mt[:3].mean() > mt[-3:].mean() * 0.5 and mt[0, :, 2].mean() >= mt[0, :, 0].mean()
    assert ob.render_mt(cam, 16, 3, 0).shape == (0, 16, 3)   # height < 4: t_height = 0, an empty Vec


def test_save_image_mt_averages_after_gamma():
    w, h, n = 32, 18, 6
    cam = ob.camera_new(w, h)
    got = ob.save_image_mt(cam, w, h, n, seed=2)
    acc = np.zeros((h, w, 3), np.float32)
    for k in range(n):
        f = ob.render_mt(cam, w, h, k, seed=2)
        acc[:16] = acc[:16] + f.astype(np.float32) / np.float32(n)
    assert np.array_equal(got, acc.astype(np.uint8)) and not got[16:].any()
    # averaging u8 values after sqrt is darker than the gamma of the linear average (Jensen): visible at low spp
    rgb, _, _ = ob.render(ob.builtin_scene(0), cam, w, h, n, seed=2, want_linear=False)
    assert got[:16].astype(int).mean() <= rgb[:16].astype(int).mean() + 0.5
