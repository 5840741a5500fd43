"""GPU parity of the post-fx kernels (fo-rma_b200/csrc/postfx.cu, through include/forma_fx.h) against the CPU
restatement of the WGSL filters (oracle/postfx_oracle.cpp).  Byte work: every comparison is bit-exact."""
import ctypes
import itertools

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from forma_b200.postfx import Effect, PostProcessor
from forma_b200.tracer import FormaError
from oracle import postfx_binding as fo

pytestmark = pytest.mark.gpu

SIZES = [(1, 1), (3, 2), (37, 53), (64, 64), (130, 31), (256, 144)]  # w % 4 != 0, tiny, square, w < 10


def frame(w, h, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


def special_frame(w, h):
    """Values around the thresholds the shaders test (0.4 -> 102, step edges at multiples of 51, r == b)."""
    vals = np.array([0, 1, 50, 51, 52, 101, 102, 103, 127, 128, 153, 204, 254, 255], np.uint8)
    rng = np.random.default_rng(7)
    f = vals[rng.integers(0, len(vals), (h, w, 4))]
    f[::3, :, 2] = f[::3, :, 0]  # r == b rows for chromostereopsis
    return f


@pytest.mark.parametrize("w,h", SIZES)
def test_every_filter_matches_the_oracle(w, h):
    with PostProcessor(w, h) as pp:
        for e in Effect:
            pp.update_effects([e])
            for f, t in ((frame(w, h, int(e)), 1.5), (special_frame(w, h), 0.0)):
                got = pp.run(f, t)
                assert np.array_equal(got, fo.chain([e], f, t)), (e.name, w, h)
            assert pp.stats().gpu_launches == 1


def test_noise_time_values():
    w, h = 200, 90
    f = frame(w, h, 3)
    with PostProcessor(w, h) as pp:
        pp.update_effects([Effect.NOISE])
        for t in (0.0, 0.016, 1.0, 59.94, 3600.0, 1e7, -2.5, float("inf"), float("nan")):
            assert np.array_equal(pp.run(f, t), fo.chain([Effect.NOISE], f, t)), t


def test_fused_unfused_and_oracle_agree_on_chains():
    rng = np.random.default_rng(42)
    chains = [[Effect.NOISE, Effect.INVERT_COLOR], [Effect.PIXELATE, Effect.WAVE, Effect.STEP],
              [Effect.GRAYSCALE, Effect.ANAGLYPH, Effect.WATERCOLOR], [Effect.NONE, Effect.NONE],
              [Effect.WATERCOLOR, Effect.NONE, Effect.INTERLACE, Effect.PIXELATE, Effect.CHROMOSTEREOPSIS],
              list(Effect), list(reversed(Effect)), [Effect.INVERT_COLOR] * 20]
    chains += [[Effect(int(x)) for x in rng.integers(0, 12, rng.integers(2, 9))] for _ in range(12)]
    for (w, h) in [(64, 64), (96, 96), (77, 77), (120, 68), (68, 120)]:
        f = frame(w, h, w + h)
        with PostProcessor(w, h) as fused, PostProcessor(w, h, unfused=True) as plain:
            for ch in chains:
                fused.update_effects(ch)
                plain.update_effects(ch)
                exp = fo.chain(ch, f, 2.75)
                a, b = fused.run(f, 2.75), plain.run(f, 2.75)
                assert np.array_equal(b, exp), ("unfused", w, h, [e.name for e in ch])
                assert np.array_equal(a, exp), ("fused", w, h, [e.name for e in ch])
                assert plain.stats().n_passes == len(ch)
                has_ragged_flip = Effect.FLIP_AXIS in ch and w != h
                assert fused.stats().n_passes <= len(ch)
                if has_ragged_flip:
                    assert fused.stats().n_passes == len(ch)  # falls back to the reference's wiring


def test_fusion_cuts_the_passes():
    with PostProcessor(64, 64) as pp:
        pp.update_effects([Effect.PIXELATE, Effect.GRAYSCALE, Effect.NOISE, Effect.INVERT_COLOR])
        assert pp.stats().n_passes == 1
        pp.update_effects([Effect.GRAYSCALE, Effect.PIXELATE, Effect.NOISE, Effect.ANAGLYPH, Effect.STEP])
        assert pp.stats().n_passes == 3
        pp.update_effects([Effect.GRAYSCALE, Effect.FLIP_AXIS, Effect.STEP])
        assert pp.stats().n_passes == 3
        pp.update_effects([Effect.INVERT_COLOR] * 40)
        assert pp.stats().n_passes == 3  # at most 16 filters per pass
        pp.update_effects([])
        assert pp.stats().n_passes == 0


def test_flip_axis_ragged_and_stale_texels():
    for (w, h) in [(50, 20), (20, 50), (33, 32)]:
        f = frame(w, h, 5)
        with PostProcessor(w, h) as pp:
            for ch in ([Effect.FLIP_AXIS], [Effect.WAVE, Effect.NONE, Effect.FLIP_AXIS, Effect.NONE],
                       [Effect.FLIP_AXIS, Effect.FLIP_AXIS], [Effect.NOISE, Effect.FLIP_AXIS, Effect.GRAYSCALE]):
                pp.update_effects(ch)
                assert np.array_equal(pp.run(f, 0.5), fo.chain(ch, f, 0.5)), (w, h, ch)
                # a second run starts from the same first-frame state
                assert np.array_equal(pp.run(f, 0.5), fo.chain(ch, f, 0.5)), (w, h, ch)


def test_empty_chain_and_resize():
    f = frame(40, 30, 1)
    with PostProcessor(40, 30) as pp:
        out = np.full_like(f, 9)
        assert (pp.run(f, 0.0, out=out) == 9).all()  # nothing dispatched: output untouched
        pp.update_effects({Effect.NONE: False, Effect.WATERCOLOR: True, Effect.NOISE: False, Effect.STEP: True})
        assert pp.effects == [Effect.WATERCOLOR, Effect.STEP]
        assert np.array_equal(pp.run(f, 0.0), fo.chain([Effect.WATERCOLOR, Effect.STEP], f, 0.0))
        pp.resize(123, 45)  # watercolor's circles depend on the frame size
        g = frame(123, 45, 2)
        assert np.array_equal(pp.run(g, 0.0), fo.chain([Effect.WATERCOLOR, Effect.STEP], g, 0.0))
        with pytest.raises(ValueError):
            pp.run(f)


def test_error_paths():
    with pytest.raises(FormaError):
        PostProcessor(0, 10)
    with pytest.raises(FormaError):
        PostProcessor(10, 10, device=99)
    with PostProcessor(8, 8) as pp:
        with pytest.raises(FormaError):
            pp.update_effects([12])
        with pytest.raises(FormaError):
            pp.update_effects([0] * 65)
        with pytest.raises(FormaError):
            pp.run_device(0, 0)
        h = ctypes.c_void_p()
        assert _lib.lib().forma_fx_create(0, 8, 8, 2, ctypes.byref(h)) == _lib.EINVAL


def test_device_frames_and_single_dispatch():
    import torch
    w, h = 254, 100  # w % 4 != 0: the scalar path; torch buffers are 256-byte aligned
    f = frame(w, h, 8)
    src = torch.from_numpy(f).cuda()
    dst = torch.full_like(src, 7)
    with PostProcessor(w, h) as pp:
        pp.update_effects([Effect.ANAGLYPH, Effect.NOISE])
        torch.cuda.synchronize()
        pp.run_device(src.data_ptr(), dst.data_ptr(), 4.0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(dst.cpu().numpy(), fo.chain([Effect.ANAGLYPH, Effect.NOISE], f, 4.0))
        # one dispatch: texels flip_axis cannot reach keep the destination's bytes
        dst.fill_(7)
        torch.cuda.synchronize()
        pp.apply_device(Effect.FLIP_AXIS, src.data_ptr(), dst.data_ptr())
        pp.synchronize()
        exp = fo.apply(Effect.FLIP_AXIS, f, out=np.full_like(f, 7))
        assert np.array_equal(dst.cpu().numpy(), exp)
        # a misaligned device frame (offset by one texel) takes the scalar path and gives the same bytes
        w2 = 252
        big = torch.zeros(w2 * h * 4 + 16, dtype=torch.uint8, device="cuda")
        g = frame(w2, h, 9)
        big[4:4 + w2 * h * 4] = torch.from_numpy(g).cuda().reshape(-1)
        out = torch.zeros(w2 * h * 4 + 16, dtype=torch.uint8, device="cuda")
        pp.resize(w2, h)
        torch.cuda.synchronize()
        pp.run_device(big.data_ptr() + 4, out.data_ptr() + 4, 4.0)
        pp.synchronize()
        assert np.array_equal(out[4:4 + w2 * h * 4].cpu().numpy().reshape(h, w2, 4),
                              fo.chain([Effect.ANAGLYPH, Effect.NOISE], g, 4.0))


def test_traced_frame_through_the_chain():
    w, h, spp = 160, 90, 4
    m = forma_b200.create_model(w, h, device=0, seed=5)
    m.scene.load_builtin(2)
    rgb_ref, _ = m.render(spp)
    chain = [Effect.WATERCOLOR, Effect.ANAGLYPH, Effect.NOISE]
    with PostProcessor(w, h) as pp:
        pp.update_effects(chain)
        rgba, rgb = pp.run_traced(m, spp, time=1.0, want_rgb=True)
        assert np.array_equal(rgb, rgb_ref)
        src = np.concatenate([rgb_ref, np.full((h, w, 1), 255, np.uint8)], -1)
        assert np.array_equal(rgba, fo.chain(chain, src, 1.0))
        pp.update_effects([])
        assert np.array_equal(pp.run_traced(m, spp), src)  # no active effect: the frame as traced, alpha 255
    with PostProcessor(w + 1, h) as pp, pytest.raises(FormaError):
        pp.run_traced(m, spp)


def test_full_hd_chain_property():
    """At 1920x1080 (oracle: ~1 s per filter): invert twice is the identity on RGB, flip twice on a square frame is
    the identity, and the fused chain equals the oracle."""
    w, h = 1920, 1080
    f = frame(w, h, 77)
    with PostProcessor(w, h) as pp:
        pp.update_effects([Effect.INVERT_COLOR, Effect.INVERT_COLOR])
        out = pp.run(f)
        assert np.array_equal(out[..., :3], f[..., :3]) and (out[..., 3] == 255).all()
        ch = [Effect.WATERCOLOR, Effect.GRAYSCALE, Effect.ANAGLYPH, Effect.STEP]
        pp.update_effects(ch)
        assert np.array_equal(pp.run(f, 12.5), fo.chain(ch, f, 12.5, threads=8))
        assert pp.stats().n_passes == 2
    with PostProcessor(1024, 1024) as pp:
        g = frame(1024, 1024, 78)
        pp.update_effects([Effect.FLIP_AXIS, Effect.FLIP_AXIS])
        assert np.array_equal(pp.run(g), g)
