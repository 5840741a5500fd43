"""GPU parity tests proper (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle
on the same seeded inputs.  Bar: BIT-EXACT — hit ids, hit distance t, per-pixel f32 sums, RGB8
bytes and the ray count (the GPU build uses -fmad=false / IEEE div+sqrt, the oracle
-ffp-contract=off, and both use the same counter-based Philox stream)."""
import os

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from oracle import binding as ob
from oracle.scene_oracle import OBJECT_DTYPE

from conftest import load_scene_pair
from test_oracle_analytic import plane, sphere, tri

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_render_equal(m, objs, cam, w, h, spp, seed, **kw):
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, spp, seed=seed, nthreads=NT, **kw)
    rgb_g, lin_g = m.render(spp, want_linear=True, **{k: v for k, v in kw.items() if k != "flags"})
    st = m.stats()
    assert st.rays == rays_o and st.paths == w * h * spp and st.gpu_launches >= 1
    assert np.array_equal(bits(lin_g[..., :3]), bits(lin_o)), "per-pixel f32 sums differ"
    assert (lin_g[..., 3] == spp).all()
    assert np.array_equal(rgb_g, rgb_o)


@pytest.mark.parametrize("which,w,h,spp", [(0, 200, 120, 6), (2, 200, 120, 6), (2, 67, 45, 3), (1, 64, 32, 2)])
def test_builtin_scenes_bit_exact(which, w, h, spp):
    m = forma_b200.create_model(w, h, seed=1234 + which)
    m.scene.load_builtin(which)
    assert_render_equal(m, ob.builtin_scene(which), ob.camera_new(w, h), w, h, spp, 1234 + which)


def test_reference_frontend_sequence_bit_exact():
    """macroquad.rs:12-13,66-68: translate/rotate objects[0] (the plane becomes live), update() orbits."""
    w, h = 160, 96
    m = forma_b200.create_model(w, h, seed=9)
    objs, cam = ob.builtin_scene(0), ob.camera_new(w, h)
    for frame, keys in enumerate([0, 0b001000, 0b010010]):
        m.scene.objects[0].translate((-1.0, 0.0, 0.0)); m.scene.objects[0].rotate((-1.0, 0.0, 0.0))
        objs["a"][0] = (-1.0, 0.0, 0.0); objs["b"][0] = (-1.0, 0.0, 0.0)
        forma_b200.update(m, keys, 0.125)
        ob.update_camera(cam, keys, 0.125)
        assert np.array_equal(bits(m.scene.camera.as_array()), bits(cam))
        rgb_o, _, _ = ob.render(objs, cam, w, h, 1, sample_offset=frame, seed=9, nthreads=NT)
        assert np.array_equal(m.pixels.reshape(h, w, 3), rgb_o)
        assert (rgb_o[..., 0] > rgb_o[..., 2] + 40).any()  # the red metal plane is in view


@pytest.mark.parametrize("name,w,h,spp", [("scene_09", 320, 180, 8), ("scene_01", 160, 90, 2), ("scene_05", 128, 72, 1),
                                          ("scene_02", 96, 54, 2), ("scene_06", 96, 54, 1), ("scene_07", 64, 36, 2)])
def test_json_scenes_bit_exact(name, w, h, spp):
    txt, objs, cam = load_scene_pair(name, w, h)
    m = forma_b200.create_model(w, h, seed=42)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    assert_render_equal(m, objs, cam, w, h, spp, 42)


def test_config_b_primary_hits_full_size():
    """BASELINE config B: scene_05 at 1920x1080, un-jittered pinhole primary rays: ids bit-exact,
    t bit-exact (the bar was 1e-4 relative)."""
    w, h = 1920, 1080
    txt, objs, cam = load_scene_pair("scene_05", w, h)
    m = forma_b200.create_model(w, h)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    ids_g, t_g = m.primary_hits()
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    assert np.array_equal(ids_g, ids_o) and np.array_equal(bits(t_g), bits(t_o))
    hit = ids_g >= 0
    assert 0.3 < hit.mean() <= 1.0 and len(np.unique(ids_g[hit] >> 16)) > 30
    assert (t_g[~hit] == np.finfo(np.float32).max).all()
    # a frame that also contains misses: the built-in camera looks past the scene
    m.scene.camera.reset()
    ids2, t2 = m.primary_hits()
    ids2_o, t2_o = ob.primary_hits(objs, ob.camera_new(w, h), w, h)
    assert np.array_equal(ids2, ids2_o) and np.array_equal(bits(t2), bits(t2_o))
    assert (t2[ids2 < 0] == np.finfo(np.float32).max).all()


def test_mixed_list_order_semantics():
    """Planes between spheres and triangles: segment boundaries, the Q5 record clobber, ties on equal t
    (earlier list entry wins), a negative-t plane accept that blocks everything after it."""
    rng = np.random.default_rng(5)
    objs = [sphere((0, 0, -3), 1.0, mat=0, col=(.8, .3, .3), id=10),
            plane((0, 0, -2.5), (0, 0, -1), (0.6, 0.6, 5), mat=1, col=(.9, .9, .9), id=11),   # in front of sphere 10
            tri((-3, -1, -4), (3, -1, -4), (0, 3, -4), id=12),
            sphere((0, 0, -3), 1.0, mat=1, col=(.1, .9, .1), fuzz=0.3, id=13),                # coincident with 10: loses ties
            plane((5, 0, -2), (0, 0, -1), (0.5, 0.5, 5), mat=0, col=(.2, .2, .9), id=14),     # extent-rejects: clobbers t,p
            tri((-3, -1, -4), (3, -1, -4), (0, 3, -4), id=15),                                 # coincident with 12: loses
            sphere((1.5, 0.5, -2), 0.4, mat=2, col=(.9, .9, .9), id=16),
            plane((0.45, -1.2, 3), (0, 0.2, -1), (0.15, 9, 9), mat=0, col=(.5, .5, .1), id=17),    # behind the eye: negative t
            sphere((-1.5, 0.5, -2), 0.4, mat=3, col=(.1, .1, .1), id=18)]
    for _ in range(6):
        c = rng.uniform(-2, 2, 3); c[2] -= 4
        objs.append(sphere(c, rng.uniform(.2, .6), mat=int(rng.integers(0, 5)), col=rng.uniform(.1, .9, 3),
                           fuzz=rng.uniform(0, 1), id=100 + _))
    objs = np.stack(objs)
    w, h = 96, 64
    m = forma_b200.create_model(w, h, seed=77)
    m.scene.load_objects(objs)
    cam = ob.camera_new(w, h)
    ids_g, t_g = m.primary_hits()
    ids_o, t_o = ob.primary_hits(objs, cam, w, h)
    assert np.array_equal(ids_g, ids_o) and np.array_equal(bits(t_g), bits(t_o))
    # 13 (coincident with 10) shows exactly where plane 11 was accepted first and RAISED closest_so_far (Q4)
    assert {10, 12, 13, 16, 17, 18, 100}.issubset(set(np.unique(ids_g))) and 11 not in ids_g and 15 not in ids_g
    assert (t_g[ids_g == 17] < 0).all()  # Q4: the plane accepts a negative t and blocks everything after it
    assert_render_equal(m, objs, cam, w, h, 8, 77)


def test_primary_beam_cull_edges():
    """render_kernel skips the scan for pixels whose whole primary beam (lens x pixel footprint) misses
    the bounding sphere of the scene: cameras far away, close up, inside the bounding sphere, with a
    fat lens, and grazing spheres at the silhouette must all stay bit-exact."""
    rng = np.random.default_rng(11)
    objs = []
    for i in range(9):  # < 16 primitives: the scan kernel, no ground sphere -> a small bounding sphere
        objs.append(sphere(rng.uniform(-1.5, 1.5, 3), rng.uniform(.2, .7), mat=int(rng.integers(0, 4)),
                           col=rng.uniform(.2, .9, 3), fuzz=rng.uniform(0, .5), id=i))
    objs.append(tri((-2, -2, 0), (2, -2, 0.5), (0, 2.5, -0.5), id=20))
    objs = np.stack(objs)
    w, h = 112, 72
    for k, (eye, at, fov, lens) in enumerate([((0, 0, 9), (0, 0, 0), 40, None), ((6, 3, 4), (0.5, 0, 0), 25, 0.4),
                                              ((0.2, 0.1, 0.3), (1, 0, 0), 90, None), ((0, 0, 2.6), (0, 3, 0), 60, 1.5),
                                              ((30, 0, 0), (0, 2.2, 0), 8, None)]):
        m = forma_b200.create_model(w, h, seed=50 + k)
        m.scene.load_objects(objs)
        cam = ob.camera_look(eye, at, (0, 1, 0), fov, w / h)
        if lens is not None:
            cam[22] = lens
        m.scene.camera.set_array(cam)
        assert_render_equal(m, objs, cam, w, h, 4, 50 + k)
        ids_g, t_g = m.primary_hits()
        ids_o, t_o = ob.primary_hits(objs, cam, w, h)
        assert np.array_equal(ids_g, ids_o) and np.array_equal(bits(t_g), bits(t_o))
        assert (ids_g < 0).any()  # there is sky in every view, so some beams are culled


def test_flags_depth_and_sample_offset():
    w, h = 120, 80
    objs, cam = ob.builtin_scene(2), ob.camera_new(w, h)
    for flags in (_lib.NO_JITTER, _lib.PINHOLE, _lib.NO_JITTER | _lib.PINHOLE):
        m = forma_b200.create_model(w, h, seed=3, flags=flags)
        m.scene.load_builtin(2)
        assert_render_equal(m, objs, cam, w, h, 2, 3, flags=flags)
    m = forma_b200.create_model(w, h, seed=3)
    m.scene.load_builtin(2)
    for depth in (0, 1, 5, 64):
        assert_render_equal(m, objs, cam, w, h, 2, 3, max_depth=depth)
    assert_render_equal(m, objs, cam, w, h, 3, 3, sample_offset=1000)
    with pytest.raises(forma_b200.tracer.FormaError):
        m.render(1, max_depth=65)
    with pytest.raises(forma_b200.tracer.FormaError):
        m.render(0)


def test_nan_propagation_matches():
    """Q10: NaN radiance (zero-length directions) must poison the same pixels on both sides."""
    objs = np.stack([sphere((0, 0, -1), 0.5, mat=1, col=(float("nan"), 1, 1), fuzz=0.0)])
    w, h = 48, 48
    m = forma_b200.create_model(w, h, seed=1)
    m.scene.load_objects(objs)
    rgb_g, lin_g = m.render(2, want_linear=True)
    rgb_o, lin_o, _ = ob.render(objs, ob.camera_new(w, h), w, h, 2, seed=1)
    assert np.isnan(lin_o).any() and np.array_equal(np.isnan(lin_o), np.isnan(lin_g[..., :3]))
    assert np.array_equal(rgb_g, rgb_o) and (rgb_g[np.isnan(lin_o)] == 0).all()
    # a path cut at max_depth is black, but 0 x inf on the way back up is NaN: the kernels may skip the walk
    # back up for black paths only when every colour is finite
    objs = np.stack([sphere((0, -1000.5, -1), 1000.0, mat=0, col=(.5, .5, .5)), sphere((0, 0, -1), 0.5, mat=0, col=(float("inf"), .5, .5))])
    m.scene.load_objects(objs)
    rgb_g, lin_g = m.render(2, max_depth=1, want_linear=True)
    rgb_o, lin_o, _ = ob.render(objs, ob.camera_new(w, h), w, h, 2, seed=1, max_depth=1)
    nan = np.isnan(lin_o)  # NaN payloads differ between SSE (0xFFC00000) and CUDA (0x7FFFFFFF): compare the pattern
    assert nan.any() and np.array_equal(nan, np.isnan(lin_g[..., :3])) and np.array_equal(rgb_g, rgb_o)
    assert np.array_equal(bits(np.where(nan, 0, lin_g[..., :3])), bits(np.where(nan, 0, lin_o)))


def test_save_image_png(tmp_path):  # tracer.rs:160-187
    from PIL import Image
    w, h = 100, 60
    m = forma_b200.create_model(w, h, seed=4)
    p = str(tmp_path / "basic.png")
    forma_b200.save_image(m, 5, p)
    rgb_o, _, _ = ob.render(ob.builtin_scene(0), ob.camera_new(w, h), w, h, 5, seed=4)
    assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), rgb_o)
    with pytest.raises(forma_b200.tracer.FormaError) as e:
        forma_b200.save_image(m, 1, str(tmp_path / "no_such_dir" / "basic.png"))
    assert e.value.code == _lib.EIO


def test_statistical_mode_against_sequential_rng():
    """SURVEY 8(c) protocol 2: the reference draws from thread_rng (OS-seeded), so beside the same-stream
    bit-exact checks the GPU image (counter RNG) is compared with the oracle driven by an independent
    SEQUENTIAL generator consumed in the reference's order (Q14).  Tolerance, stated: the per-pixel RMSE of
    the RGB8 images must not exceed 1.5 x the noise floor = RMSE of two sequential-RNG oracle runs with
    different seeds at the same spp, and PSNR must be >= 30 dB at 512 spp."""
    w, h, spp = 96, 64, 512
    objs, cam = ob.builtin_scene(2), ob.camera_new(w, h)  # metal fuzz .2 / 1.0, two dielectrics (scenes.rs:62-156)
    m = forma_b200.create_model(w, h, seed=2024)
    m.scene.load_builtin(2)
    rgb_g, _ = m.render(spp, want_linear=True)
    rgb_a, _, _ = ob.render(objs, cam, w, h, spp, seed=1, flags=ob.FLAG_SEQ_RNG, nthreads=NT)
    rgb_b, _, _ = ob.render(objs, cam, w, h, spp, seed=2, flags=ob.FLAG_SEQ_RNG, nthreads=NT)

    def rmse(x, y):
        return float(np.sqrt(np.mean((x.astype(np.float64) - y.astype(np.float64)) ** 2)))

    floor = rmse(rgb_a, rgb_b)
    got = max(rmse(rgb_g, rgb_a), rmse(rgb_g, rgb_b))
    psnr = 20.0 * np.log10(255.0 / got)
    assert 0.0 < floor < 8.0, floor
    assert got <= 1.5 * floor, (got, floor)
    assert psnr >= 30.0, psnr
    # and no bias: channel means agree to a fraction of a grey level
    assert np.abs(rgb_g.mean(axis=(0, 1)) - 0.5 * (rgb_a.mean(axis=(0, 1)) + rgb_b.mean(axis=(0, 1)))).max() < 0.6


def test_progressive_accumulation_is_one_render():
    """k calls of accumulate() continue tracer.rs:170-175's sum, so they are bit-identical to one
    render of the total - on the scan path and on the tree path - and update_progressive() restarts exactly
    when the camera or the scene changed."""
    for which, w, h in ((2, 96, 64), (None, 80, 48)):
        m = forma_b200.create_model(w, h, seed=31)
        if which is None:
            txt, objs, cam = load_scene_pair("scene_01", w, h)  # 363 primitives: LBVH, render_tree_kernel
            m.scene.load_json(txt); m.scene.camera.use_json()
        else:
            m.scene.load_builtin(which)
            objs, cam = ob.builtin_scene(which), ob.camera_new(w, h)
        for k, more in enumerate((1, 3, 4)):
            rgb_p, lin_p, n = m.accumulate(more, want_linear=True)
        assert n == 8 and (lin_p[..., 3] == 8).all()
        rgb_o, lin_o, _ = ob.render(objs, cam, w, h, 8, seed=31, nthreads=NT)
        assert np.array_equal(bits(lin_p[..., :3]), bits(lin_o)) and np.array_equal(rgb_p, rgb_o)
        m.accumulate_reset()
        rgb_1, _, n = m.accumulate(2)
        assert n == 2 and np.array_equal(rgb_1, ob.render(objs, cam, w, h, 2, seed=31, nthreads=NT)[0])
    # the interactive loop: still camera -> refinement, a key -> fresh frame
    w, h = 96, 64
    m = forma_b200.create_model(w, h, seed=8)
    cam = ob.camera_new(w, h)
    ob.update_camera(cam, 0, 0.1)  # the first update() teleports the eye (Q12)
    counts = [forma_b200.update_progressive(m, 0, 0.1) for _ in range(3)]
    assert counts == [1, 2, 3]
    assert np.array_equal(m.pixels.reshape(h, w, 3), ob.render(ob.builtin_scene(0), cam, w, h, 3, seed=8, nthreads=NT)[0])
    assert forma_b200.update_progressive(m, 0b000100, 0.1) == 1  # 'A': the camera moved
    ob.update_camera(cam, 0b000100, 0.1)
    assert np.array_equal(m.pixels.reshape(h, w, 3), ob.render(ob.builtin_scene(0), cam, w, h, 1, seed=8, nthreads=NT)[0])
    m.scene.objects[0].rotate((-1.0, 0.0, 0.0))  # the scene changed
    assert forma_b200.update_progressive(m, 0, 0.1) == 1


def test_srgb_encode_on_device():
    """color_utils::srgb_encode as a GPU pass over RGB8, standalone and as an output flag."""
    import torch
    t = forma_b200.srgb_table()
    w, h = 100, 37  # 11100 bytes: a multiple of 4; the odd-length tail is exercised below
    m = forma_b200.create_model(w, h, seed=4)
    m.scene.load_builtin(2)
    rgb, _ = m.render(4)
    ms = forma_b200.create_model(w, h, seed=4, flags=_lib.SRGB_OUT)
    ms.scene.load_builtin(2)
    rgb_s, _ = ms.render(4)
    assert np.array_equal(rgb_s, t[rgb]) and not np.array_equal(rgb_s, rgb)
    for n in (1, 2, 3, 255, 4096 + 3, 1 << 20):
        x = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
        want = t[x.cpu().numpy()]
        m.srgb_encode_device(x.data_ptr(), n)
        m.synchronize()
        torch.cuda.synchronize()
        assert np.array_equal(x.cpu().numpy(), want)


def test_edge_cases_empty_tiny_and_degenerate():
    """Empty list (every ray is sky), a 1x1 and a 1-row frame, a plane-only scene, and a tree over degenerate
    geometry: coincident spheres (ties), a zero-radius sphere, zero-area triangles, a NaN sphere (stays on the
    exact linear scan) - all bit-exact against the oracle on the scan and the LBVH path."""
    empty = np.zeros(0, dtype=OBJECT_DTYPE)
    for w, h, spp in ((1, 1, 5), (37, 1, 3), (64, 40, 2)):
        m = forma_b200.create_model(w, h, seed=3)
        m.scene.load_objects(empty)
        assert_render_equal(m, empty, ob.camera_new(w, h), w, h, spp, 3)
        ids, t = m.primary_hits()
        assert (ids == -1).all() and (t == np.finfo(np.float32).max).all()
    w, h = 72, 40
    only_plane = np.stack([plane((0, 0, -2), (0, 0, -1), (1.5, 1.0, 5), mat=1, col=(.9, .6, .6), id=5)])
    m = forma_b200.create_model(w, h, seed=4)
    m.scene.load_objects(only_plane)
    assert_render_equal(m, only_plane, ob.camera_new(w, h), w, h, 3, 4)
    rng = np.random.default_rng(21)
    objs = [sphere((0, 0, -3), 0.5, mat=0, col=(.8, .2, .2), id=i) for i in range(3)]             # coincident: lowest index wins
    objs += [sphere((1, 0, -3), 0.0, mat=0, id=10), tri((0, 1, -3), (0, 1, -3), (0, 1, -3), id=11),  # zero radius / zero area
             tri((-1, -1, -3), (1, 1, -3), (0, 0, -3), id=12),                                      # collinear
             sphere((float("nan"), 0, -3), 0.5, mat=0, id=13)]                                      # NaN centre: never hit
    for i in range(24):
        c = rng.uniform(-2, 2, 3); c[2] -= 4
        objs.append(sphere(c, rng.uniform(.1, .5), mat=int(rng.integers(0, 4)), col=rng.uniform(.1, .9, 3), fuzz=rng.uniform(0, 1), id=100 + i))
    objs = np.stack(objs)
    cam = ob.camera_new(w, h)
    for accel in (_lib.ACCEL_BRUTE, _lib.ACCEL_LBVH):
        m = forma_b200.create_model(w, h, seed=6, accel=accel)
        m.scene.load_objects(objs)
        assert_render_equal(m, objs, cam, w, h, 3, 6)
        ids, t = m.primary_hits()
        ids_o, t_o = ob.primary_hits(objs, cam, w, h)
        assert np.array_equal(ids, ids_o) and np.array_equal(bits(t), bits(t_o))
        assert 0 in ids and 1 not in ids and 2 not in ids and 13 not in ids
    assert m.stats().accel_in_use == _lib.ACCEL_LBVH


@pytest.mark.parametrize("w,h", [(67, 45), (33, 5), (130, 7), (5, 3), (1, 1), (3, 64)])
def test_job_queue_paths_on_ragged_frames(w, h):
    """accumulate_kernel owns groups of four pixels of a row and stores whole groups as three 32-bit words: widths
    that are not multiples of 4 (byte path, short last group), frames smaller than a tile, one pixel - through
    render_queue_kernel (get_objects, 5 spp: the eye is inside the ground sphere's bounds) and through
    render_tree_kernel (scene_01 with the LBVH forced, 4 spp), bit-exact against the oracle."""
    m = forma_b200.create_model(w, h, seed=77)
    m.scene.load_builtin(2)
    assert_render_equal(m, ob.builtin_scene(2), ob.camera_new(w, h), w, h, 5, 77)
    assert m.stats().gpu_launches == 2  # trace + accumulate: the job-queue scan, not the pixel kernel
    txt, objs, cam = load_scene_pair("scene_01", w, h)
    m = forma_b200.create_model(w, h, seed=78, accel=_lib.ACCEL_LBVH)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    assert_render_equal(m, objs, cam, w, h, 4, 78)
    assert m.stats().accel_in_use == _lib.ACCEL_LBVH
