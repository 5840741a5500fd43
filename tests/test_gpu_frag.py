"""GPU parity of the fragment-shader ray tracers (fo-rma_b200/csrc/frag.cu, through include/forma_frag.h) against
the CPU restatement of rt.wgsl / rt2.wgsl (oracle/wgsl_rt_oracle.cpp).  Both sides evaluate the same IEEE f32
operations in the same order (pow through the same fixed exp2 / log2 recipe), so every comparison is bit-exact:
f32 frames by their bit patterns, RGBA8 frames by bytes, ray counts as integers."""
import ctypes

import numpy as np
import pytest

from forma_b200 import _lib
from forma_b200.frag import Shader, State, light_direction
from forma_b200.tracer import FormaError
from oracle import wgsl_rt_binding as wb

pytestmark = pytest.mark.gpu

SIZES = [(1, 1), (3, 2), (16, 8), (17, 9), (333, 7), (7, 333), (320, 240)]  # tile-aligned, ragged, thin


def same(got_u8, got_lin, ref):
    """Bytes equal; f32 frames equal bit for bit, a NaN matching a NaN (IEEE leaves sign and payload of a generated
    NaN to the implementation: x86 makes 0/0 = 0xffc00000, the GPU 0x7fffffff)."""
    ref_u8, ref_lin, _ = ref[:3]
    a, b = got_lin.view(np.uint32), ref_lin.view(np.uint32)
    bad = (a != b) & ~(np.isnan(got_lin) & np.isnan(ref_lin))
    if bad.any() or not np.array_equal(got_u8, ref_u8):
        ys, xs = np.nonzero(bad.any(-1))
        print("f32 mismatches:", int(bad.any(-1).sum()), "u8 mismatches:", int((got_u8 != ref_u8).any(-1).sum()))
        for y, x in list(zip(ys, xs))[:8]:
            print((x, y), got_lin[y, x], ref_lin[y, x], got_u8[y, x], ref_u8[y, x])
        return False
    return True


@pytest.mark.parametrize("shader", [Shader.RT, Shader.RT2])
@pytest.mark.parametrize("w,h", SIZES)
def test_frame_matches_the_oracle(shader, w, h):
    for srgb in (False, True):
        with State(w, h, shader=shader, srgb_surface=srgb) as st:
            for res in ((w, h), None, (640, 480)):
                u8, lin = st.render(2.5, resolution=res, want_linear=True)
                ref = wb.render(int(shader), w, h, time=2.5, srgb=srgb, resolution=res, threads=4)
                assert same(u8, lin, ref), (shader, w, h, srgb, res)
                s = st.stats()
                assert s.rays == ref[2] and s.gpu_launches == 1 and s.kernel_ms > 0


@pytest.mark.parametrize("shader", [Shader.RT, Shader.RT2])
def test_full_size_frame_of_the_shaders_constants(shader):
    # 1600 x 1200 as rt.wgsl:163 / rt2.wgsl:176-177 assume; includes rt.wgsl's pixels that run all 1000 loop trips
    w, h = 1600, 1200
    with State(w, h, shader=shader, srgb_surface=True) as st:
        u8, lin = st.render(0.0, want_linear=True)
        ref = wb.render(int(shader), w, h, time=0.0, srgb=True, threads=8, want_bounces=True)
        assert same(u8, lin, ref)
        assert st.stats().rays == ref[2]
        if shader == Shader.RT:
            assert ref[3].max() == 1000
        # rgba8 alone (no f32 frame requested) gives the same bytes
        assert np.array_equal(st.render(0.0), ref[0])


def test_rt2_time_values():
    w, h = 200, 150
    with State(w, h, shader=Shader.RT2) as st:
        for t in (0.0, 0.016, 1.0, 15.70796, 31.4159, 59.94, 3600.0, -2.5, 1e7):
            u8, lin = st.render(t, resolution=(w, h), want_linear=True)
            assert same(u8, lin, wb.render(wb.RT2, w, h, time=t, resolution=(w, h), threads=4)), t
            assert np.array_equal(light_direction(t).view(np.uint32), wb.light2(t).view(np.uint32))
        # infinite time: cos / sin are NaN, every lit term is NaN, the unorm target stores 0
        u8, lin = st.render(float("inf"), resolution=(w, h), want_linear=True)
        ref = wb.render(wb.RT2, w, h, time=float("inf"), resolution=(w, h))
        assert np.array_equal(u8, ref[0]) and np.array_equal(np.isnan(lin), np.isnan(ref[1]))


def test_aspect_override_and_wide_frames():
    for shader in (Shader.RT, Shader.RT2):
        w, h = 384, 108
        with State(w, h, shader=shader) as st:
            for aspect in (0.0, 16 / 9, 1.0):
                u8, lin = st.render(4.0, resolution=(w, h), aspect=aspect, want_linear=True)
                ref = wb.render(int(shader), w, h, time=4.0, resolution=(w, h), aspect=aspect, threads=4)
                assert same(u8, lin, ref), (shader, aspect)


def test_resize_and_shader_switch_on_one_handle():
    with State(64, 48) as st:
        a = st.render(1.0, resolution=(64, 48))
        assert np.array_equal(a, wb.render(wb.RT, 64, 48, time=1.0, resolution=(64, 48))[0])
        st.resize(0, 10)          # wnt.rs:244: ignored
        assert (st.width, st.height) == (64, 48)
        st.resize(150, 31)
        b = st.render(1.0, shader=Shader.RT2, resolution=(150, 31))
        assert b.shape == (31, 150, 4)
        assert np.array_equal(b, wb.render(wb.RT2, 150, 31, time=1.0, resolution=(150, 31))[0])
        c = st.render(1.0, resolution=(64, 48))  # back to rt.wgsl, now a 150 x 31 window on the 64 x 48 mapping
        assert np.array_equal(c, wb.render(wb.RT, 150, 31, time=1.0, resolution=(64, 48))[0])


def test_device_frames_on_a_caller_stream():
    import torch
    w, h = 256, 160
    dev = torch.device("cuda:0")
    d_u8 = torch.zeros((h, w, 4), dtype=torch.uint8, device=dev)
    d_lin = torch.zeros((h, w, 4), dtype=torch.float32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    with State(w, h, shader=Shader.RT2) as st:
        with torch.cuda.stream(stream):
            st.render_device(d_u8.data_ptr(), d_lin.data_ptr(), 7.0, resolution=(w, h), stream=stream.cuda_stream)
        stream.synchronize()
        ref = wb.render(wb.RT2, w, h, time=7.0, resolution=(w, h), threads=4)
        assert same(d_u8.cpu().numpy(), d_lin.cpu().numpy(), ref)
        assert st.stats().rays == ref[2]
        # one output only
        d_u8.zero_()
        st.render_device(d_u8.data_ptr(), None, 7.0, resolution=(w, h), shader=Shader.RT)
        st.synchronize()
        assert np.array_equal(d_u8.cpu().numpy(), wb.render(wb.RT, w, h, time=7.0, resolution=(w, h), threads=4)[0])


def test_repeatable_and_errors():
    with State(120, 90, shader=Shader.RT) as st:
        a = st.render(0.0)
        assert np.array_equal(a, st.render(0.0))
        with pytest.raises(FormaError) as e:
            st.render(0.0, shader=3)
        assert e.value.code == _lib.EINVAL
        p = _lib.FragParams(float("nan"), 0, 0, 0, 0)
        buf = np.empty((90, 120, 4), np.uint8)
        assert _lib.lib().forma_frag_render(st._h, 1, ctypes.byref(p), buf.ctypes.data, None) == _lib.EINVAL
        p = _lib.FragParams(0, 0, 0, 0, 2)
        assert _lib.lib().forma_frag_render(st._h, 1, ctypes.byref(p), buf.ctypes.data, None) == _lib.EINVAL
        # NULL params = time 0, the shaders' constants, unorm target
        assert _lib.lib().forma_frag_render(st._h, 1, None, buf.ctypes.data, None) == 0
        assert np.array_equal(buf, a)
    with pytest.raises(FormaError):
        State(0, 5)
    assert _lib.lib().forma_frag_render(None, 1, None, None, None) == _lib.EINVAL
