"""-m gpu: BASELINE.json's full sizes, checked through size-independent properties (the oracle is
too slow there), plus the device-resident entry points the multi-GPU path uses."""
import numpy as np
import pytest

import forma_b200
from oracle import binding as ob

from conftest import load_scene_pair

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def config_c():
    w, h = 1920, 1080
    txt, objs, cam = load_scene_pair("scene_09", w, h)
    m = forma_b200.create_model(w, h, seed=0)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    return m, objs, cam, w, h


def test_config_c_full_size_properties(config_c):
    """scene_09, 1920x1080, 256 spp, depth 50."""
    m, objs, cam, w, h = config_c
    rgb, lin = m.render(256, want_linear=True)
    st = m.stats()
    assert st.paths == w * h * 256 and st.paths <= st.rays < st.paths * 1.01  # almost every path is one sky ray
    # idempotence: the counter RNG makes a re-render identical
    rgb2, lin2 = m.render(256, want_linear=True)
    assert np.array_equal(rgb, rgb2) and np.array_equal(bits(lin), bits(lin2)) and m.stats().rays == st.rays
    # oracle on a band of rows at full spp (bit-exact), rows chosen through the tetrahedron
    ids, _ = m.primary_hits()
    ys = np.nonzero((ids >= 0).any(axis=1))[0]
    y0 = int(ys[len(ys) // 2])
    _, lin_o, _ = ob.render(objs, cam, w, h, 256, seed=0, y0=y0, y1=y0 + 2, nthreads=16, want_rgb=False)
    assert np.array_equal(bits(lin[y0:y0 + 2, :, :3]), bits(lin_o[y0:y0 + 2]))
    # write-out of the whole frame from the sums (oracle tone-map): tracer.rs:177-184
    assert np.array_equal(ob.tonemap(lin[..., :3], 256), rgb)
    # sky pixels: mean of 256 samples of a smooth gradient stays inside the gradient's range
    sky = lin[0, :, :3] / 256
    assert (sky[:, 2] == 1.0).all() and (sky[:, 0] >= 0.5).all() and (sky[:, 0] <= 1.0).all()


def test_sample_split_composes(config_c):
    """8 disjoint sample ranges (what 8 ranks render) sum to the single-GPU frame within f32
    re-association error; the RGB8 frame differs by at most 1 LSB on rare pixels."""
    m, _, _, w, h = config_c
    rgb, lin = m.render(64, want_linear=True)
    acc = np.zeros_like(lin)
    for r in range(8):
        _, part = m.render(8, sample_offset=8 * r, want_rgb=False, want_linear=True)
        acc += part
    assert (acc[..., 3] == 64).all()
    np.testing.assert_allclose(acc[..., :3], lin[..., :3], rtol=2e-6, atol=0)
    d = np.abs(ob.tonemap(acc[..., :3], 64).astype(int) - rgb.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3


def test_device_entry_points_and_fused_reduce():
    import torch
    w, h = 640, 360
    txt, objs, cam = load_scene_pair("scene_01", w, h)
    m = forma_b200.create_model(w, h, seed=5)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    n = w * h
    parts = [torch.empty((n, 4), dtype=torch.float32, device="cuda") for _ in range(4)]
    ts = torch.cuda.Stream()  # a non-default stream: handle 0 would mean "the handle's own stream"
    stream = ts.cuda_stream
    ts.wait_stream(torch.cuda.current_stream())
    for r, p in enumerate(parts):
        m.render_device(2, 2 * r, 50, None, p.data_ptr(), stream)
    rgb = torch.empty((n, 3), dtype=torch.uint8, device="cuda")
    total = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    m.reduce_tonemap_device([p.data_ptr() for p in parts], 8, 0, n, rgb.data_ptr(), total.data_ptr(), stream)
    ts.synchronize()
    want = parts[0].clone()
    for p in parts[1:]:
        want += p  # same rank order as the kernel
    assert torch.equal(total, want)
    host = total.cpu().numpy()
    assert np.array_equal(ob.tonemap(host[:, :3], 8), rgb.cpu().numpy())
    # per-rank parts equal the oracle's sample ranges bit for bit
    _, lin_o, _ = ob.render(objs, cam, w, h, 2, sample_offset=4, seed=5, nthreads=16, want_rgb=False)
    assert np.array_equal(bits(parts[2].cpu().numpy()[:, :3]), bits(lin_o.reshape(n, 3)))
    # stripe-wise tonemap (what each rank does after a reduce-scatter)
    rgb2 = torch.zeros((n, 3), dtype=torch.uint8, device="cuda")
    for a, b in ((0, n // 3), (n // 3, n)):
        m.tonemap_device(total.data_ptr(), 8, rgb2.data_ptr(), a, b, stream)
    ts.synchronize()
    assert torch.equal(rgb, rgb2)


def test_tile_interleaved_partition_is_bit_exact():
    """SURVEY 8(e) tile-interleave: handles with partitions (r, n) writing ONE device frame reproduce the
    unpartitioned render bit for bit - scan kernel and tree kernel, RGB8 and f32 sums, ragged frame size."""
    import torch
    from forma_b200 import distributed as fd
    for scene, w, h, spp in (("builtin2", 203, 77, 3), ("scene_01", 160, 90, 2)):
        def model():
            m = forma_b200.create_model(w, h, seed=12)
            if scene.startswith("builtin"):
                m.scene.load_builtin(2)
            else:
                txt, _, _ = load_scene_pair(scene, w, h)
                m.scene.load_json(txt); m.scene.camera.use_json()
            return m
        whole = model()
        rgb_w, lin_w = whole.render(spp, want_linear=True)
        rays_w = whole.stats().rays
        for world in (2, 3):
            rgb = torch.full((h, w, 3), 7, dtype=torch.uint8, device="cuda")
            lin = torch.full((h, w, 4), -1.0, dtype=torch.float32, device="cuda")
            ts = torch.cuda.Stream()
            ts.wait_stream(torch.cuda.current_stream())
            rays = paths = 0
            own = fd.owner_map(w, h, world)
            for r in range(world):
                m = model()
                m.set_partition(r, world)
                m.render_device(spp, 0, 50, rgb.data_ptr(), lin.data_ptr(), ts.cuda_stream)
                ts.synchronize()
                st = m.stats()
                rays += st.rays; paths += st.paths
                assert st.paths == int((own == r).sum()) * spp
                # pixels of the ranks still to come are untouched
                later = torch.from_numpy(own > r).cuda()
                assert (rgb[later] == 7).all() and (lin[later] == -1.0).all()
            assert rays == rays_w and paths == w * h * spp
            assert np.array_equal(rgb.cpu().numpy(), rgb_w)
            assert np.array_equal(bits(lin.cpu().numpy()), bits(lin_w))


def test_stream_ordered_flags():
    """forma_rt_flag_set_device / forma_rt_flag_wait_device: a wait queued on one stream holds back the work behind
    it until another stream raises the flag (the ordering the multi-GPU exchange relies on, here on one device)."""
    import ctypes
    import torch
    from forma_b200 import _lib
    m = forma_b200.create_model(64, 48, seed=0)
    p = ctypes.c_void_p()
    _lib.check(_lib.lib().forma_rt_device_alloc(m._h, 256, ctypes.byref(p)))
    flag_a, flag_b = p.value, p.value + 128
    dev = torch.device("cuda:0")
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    marker = torch.zeros(1, dtype=torch.int32, device=dev)
    # already satisfied: 0 >= 0, then set 5 and wait for 5 on the same stream
    m.flag_wait_device([flag_a, flag_b], 0, s1.cuda_stream)
    m.flag_set_device(flag_a, 5, s1.cuda_stream)
    m.flag_wait_device([flag_a], 5, s1.cuda_stream)
    s1.synchronize()
    # s1 waits for flag_b >= 7, then writes the marker; s2 raises flag_b only after we have looked at the marker
    m.flag_wait_device([flag_b], 7, s1.cuda_stream)
    with torch.cuda.stream(s1):
        marker.fill_(1)
    import time
    time.sleep(0.2)
    assert not s1.query()                      # still held back by the wait
    m.flag_set_device(flag_b, 6, s2.cuda_stream)  # not enough
    s2.synchronize()
    time.sleep(0.1)
    assert not s1.query()
    m.flag_set_device(flag_b, 7, s2.cuda_stream)
    s2.synchronize()
    s1.synchronize()
    assert int(marker.item()) == 1
    with pytest.raises(forma_b200.tracer.FormaError):
        m.flag_wait_device([], 1)
    with pytest.raises(forma_b200.tracer.FormaError):
        m.flag_set_device(flag_a + 1, 1)


def test_config_e_4k_sample_ranges_and_reduce_vs_oracle():
    """BASELINE config E's per-GPU unit of work - scene_09 at 3840x2160, 512 spp of a sample range - for two ranks'
    ranges [0, 512) and [512, 1024), then the exchange step (rank-order sum + tone-map, forma_rt_reduce_tonemap_device)
    - against the oracle on a two-row band through the tetrahedron: per-range f32 sums bit for bit, and the RGB8 band
    of the reduced frame equal to the oracle's tone-map (tracer.rs:177-184) of the same rank-order sum."""
    import torch
    w, h, spp = 3840, 2160, 512
    txt, objs, cam = load_scene_pair("scene_09", w, h)
    m = forma_b200.create_model(w, h, seed=0)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    ids, _ = m.primary_hits()
    ys = np.nonzero((ids >= 0).any(axis=1))[0]
    y0 = int(ys[len(ys) // 2])
    dev = torch.device("cuda", 0)
    npix = w * h
    parts = [torch.zeros((npix, 4), dtype=torch.float32, device=dev) for _ in range(2)]
    rgb = torch.zeros(npix * 3, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for r, p in enumerate(parts):
        m.render_device(spp, r * spp, 50, None, p.data_ptr(), stream)
    m.reduce_tonemap_device([p.data_ptr() for p in parts], 2 * spp, 0, npix, rgb.data_ptr(), None, stream)
    torch.cuda.synchronize()
    band = slice(y0 * w, (y0 + 2) * w)
    lin_o = []
    for r in range(2):
        _, lo, _ = ob.render(objs, cam, w, h, spp, sample_offset=r * spp, seed=0, y0=y0, y1=y0 + 2, nthreads=16, want_rgb=False)
        lin_o.append(lo[y0:y0 + 2].reshape(-1, 3))
        assert np.array_equal(bits(parts[r][band, :3].cpu().numpy()), bits(lin_o[r]))
    want = ob.tonemap(lin_o[0] + lin_o[1], 2 * spp)  # f32 add in rank order, as the kernel does
    got = rgb.cpu().numpy().reshape(h * w, 3)[band]
    assert np.array_equal(got, want)
    assert (parts[0][:, 3] == spp).all().item()
