"""CPU tests of the post-fx checker (oracle/postfx_oracle.cpp) and of the forma_fx C ABI surface.

The reference holds no golden data for its compute filters and cannot run here (wgpu), so the restatement is checked
against an INDEPENDENT numpy evaluation of the simple shaders, hand-derived values, and structural properties; its
outputs are then pinned (tests/golden/postfx_pins.json) so that it cannot drift silently.
"""
import hashlib
import json
import os
import re

import numpy as np
import pytest

from forma_b200 import _lib
from forma_b200.postfx import Effect, effect_to_name
from oracle import postfx_binding as fo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F = np.float32


def frame(w, h, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


def dec(u8):
    return u8.astype(F) / F(255.0)


def enc(x):
    x = np.where(np.isnan(x), F(0), x).astype(F)
    x = np.clip(x, F(0), F(1))
    return (x * F(255.0) + F(0.5)).astype(np.uint8)


def luma(c):
    return c[..., 0] * F(0.299) + c[..., 1] * F(0.587) + c[..., 2] * F(0.114)


def test_library_exports_every_forma_fx_symbol():
    hdr = open(os.path.join(ROOT, "include", "forma_fx.h")).read()
    declared = set(re.findall(r"FORMA_FX_API[^;]*?\b(forma_fx_\w+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS_FX), declared ^ set(_lib.SYMBOLS_FX)
    L = _lib.lib()
    for name in declared:
        assert getattr(L, name)


def test_effect_names_follow_shader_utils():
    # shader_utils.rs:73-88, in enum order (note "flipaxis", not "flip_axis")
    names = ["none", "noise", "pixelate", "invert_color", "wave", "interlace", "flipaxis", "grayscale", "step",
             "watercolor", "chromostereopsis", "anaglyph"]
    assert [effect_to_name(e) for e in Effect] == names
    assert [int(e) for e in Effect] == list(range(12))
    with pytest.raises(ValueError):
        effect_to_name(12)


def test_fx_create_without_a_device_fails_loudly():
    import ctypes
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = ctypes.c_void_p()
    rc = _lib.lib().forma_fx_create(0, 8, 8, 0, ctypes.byref(h))
    assert rc == _lib.ECUDA and not h.value
    assert b"no CPU fallback" in _lib.lib().forma_rt_last_error()


def test_unorm_round_trip_is_identity():
    u = np.arange(256, dtype=np.uint8)
    assert np.array_equal(enc(dec(u)), u)
    f = np.zeros((1, 256, 4), np.uint8)
    f[0, :, :] = u[:, None]
    assert np.array_equal(fo.apply(Effect.NONE, f), f)


@pytest.mark.parametrize("w,h", [(1, 1), (7, 5), (64, 64), (131, 77)])
def test_simple_filters_against_numpy(w, h):
    f = frame(w, h, seed=w * 1000 + h)
    c = dec(f)
    one = np.ones((h, w), F)
    # invert_color.comp.wgsl
    exp = enc(np.stack([F(1) - c[..., 0], F(1) - c[..., 1], F(1) - c[..., 2], one], -1))
    assert np.array_equal(fo.apply(Effect.INVERT_COLOR, f), exp)
    # grayscale.comp.wgsl
    y = luma(c)
    assert np.array_equal(fo.apply(Effect.GRAYSCALE, f), enc(np.stack([y, y, y, c[..., 3]], -1)))
    # step.comp.wgsl
    b = np.floor(y / F(0.2)) * F(0.2)
    assert np.array_equal(fo.apply(Effect.STEP, f), enc(np.stack([b, b, b, c[..., 3]], -1)))
    # interlace.comp.wgsl: even rows black, alpha kept
    exp = f.copy()
    exp[0::2, :, :3] = 0
    assert np.array_equal(fo.apply(Effect.INTERLACE, f), exp)
    # pixelate.comp.wgsl: block origin
    yy, xx = np.mgrid[0:h, 0:w]
    assert np.array_equal(fo.apply(Effect.PIXELATE, f), f[(yy // 8) * 8, (xx // 8) * 8])
    # chromostereopsis.comp.wgsl
    r = (f[..., 0].astype(int) > f[..., 2].astype(int))
    exp = np.zeros_like(f)
    exp[..., 0] = np.where(r, 255, 0)
    exp[..., 2] = np.where(r, 0, 255)
    exp[..., 3] = 255
    assert np.array_equal(fo.apply(Effect.CHROMOSTEREOPSIS, f), exp)
    # anaglyph.comp.wgsl: red from x - 10, blue from x + 10, zero outside the frame
    exp = np.zeros_like(f)
    exp[..., 3] = 255
    if w > 10:
        exp[:, 10:, 0] = f[:, :-10, 0]
        exp[:, :-10, 2] = f[:, 10:, 2]
    assert np.array_equal(fo.apply(Effect.ANAGLYPH, f), exp)
    # wave.comp.wgsl
    s = np.stack([y, y, y, one], -1)
    s = np.where((c[..., 0] > F(0.4))[..., None], np.stack([one, c[..., 1], F(0.1) * c[..., 2], one], -1), s)
    s = np.where((c[..., 1] > F(0.4))[..., None], np.stack([F(0.1) * one, c[..., 1], F(0.1) * c[..., 2], one], -1), s)
    s = np.where((c[..., 2] > F(0.4))[..., None], np.stack([F(0.1) * c[..., 0], c[..., 1], one, one], -1), s)
    assert np.array_equal(fo.apply(Effect.WAVE, f), enc(s))


def test_noise_against_numpy():
    w, h, t = 97, 41, F(3.25)
    f = frame(w, h, 5)
    c = dec(f)
    yy, xx = np.mgrid[0:h, 0:w]
    fx, fy = xx.astype(F) / F(10), yy.astype(F) / F(10)
    tt = t * F(0.05)
    vx, vy = fx * F(0.3183099) + tt, fy * F(0.3678794) + tt
    inner = vx * vy * (vx + vy)
    f1 = inner - np.floor(inner)
    f2 = F(23) * f1
    n = (f2 - np.floor(f2)) / F(20)
    exp = enc(np.stack([c[..., 0] + n, c[..., 1] + n, c[..., 2] + n, np.ones((h, w), F)], -1))
    assert np.array_equal(fo.apply(Effect.NOISE, f, float(t)), exp)
    # the other filters ignore the time uniform
    for e in Effect:
        if e != Effect.NOISE:
            assert np.array_equal(fo.apply(e, f, 0.0), fo.apply(e, f, 99.0))
    # a NaN uniform makes NaN colours, which textureStore writes as 0
    out = fo.apply(Effect.NOISE, f, float("nan"))
    assert not out[..., :3].any() and (out[..., 3] == 255).all()


def _fract(x):
    return x - np.floor(x)


def _hash_wc(px, py):
    # watercolor.comp.wgsl:7-10
    px, py = F(px), F(py)
    x, y, z = _fract(px * F(0.1031)), _fract(py * F(0.1031)), _fract(px * F(0.1031))
    d = x * (y + F(33.333)) + y * (z + F(33.333)) + z * (x + F(33.333))
    return _fract((x + y) * d)


@pytest.mark.parametrize("w,h", [(160, 120), (33, 200)])
def test_watercolor_against_numpy(w, h):
    f = frame(w, h, 21)
    f[: h // 2] //= 4  # leave headroom so that the additions show
    c = dec(f)
    r, a = c[..., 0].copy(), c[..., 3].copy()
    yy, xx = np.mgrid[0:h, 0:w]
    posx, posy = xx.astype(F), yy.astype(F)
    for i in range(50):
        fi = F(i)
        qw, qh = F(w) / F(4), F(h) / F(4)
        cx = qw + _hash_wc(fi * F(123.45), F(67.89)) * qw * F(2)
        cy = qh + _hash_wc(fi * F(123.45) + F(1), F(67.89) + F(0)) * qh * F(2)
        radius = F(10) + _hash_wc(fi * F(234.56), F(78.9)) * F(200)
        red = _hash_wc(fi * F(345.67), F(89.01))
        dx, dy = posx - cx, posy - cy
        inside = np.sqrt(dx * dx + dy * dy) <= radius
        r = np.where(inside, r + red * F(0.05), r)
        a = np.where(inside, a + F(0.05), a)
    exp = enc(np.stack([r, c[..., 1], c[..., 2], a], -1))
    out = fo.apply(Effect.WATERCOLOR, f)
    assert np.array_equal(out, exp)
    assert np.array_equal(out[..., 1:3], f[..., 1:3])  # only red and alpha ever change


def test_flip_axis_square_and_ragged():
    f = frame(32, 32, 3)
    assert np.array_equal(fo.apply(Effect.FLIP_AXIS, f), f.transpose(1, 0, 2))
    # non-square: (x, y) -> (y, x) only where the target exists; the rest of dst is left alone
    w, h = 50, 20
    f = frame(w, h, 4)
    dst = np.full((h, w, 4), 7, np.uint8)
    out = fo.apply(Effect.FLIP_AXIS, f, out=dst)
    m = min(w, h)
    assert np.array_equal(out[:m, :m], f[:m, :m].transpose(1, 0, 2))
    assert (out[:, m:] == 7).all()
    # in a chain the unwritten texels read as the first frame's zeros
    assert not fo.chain([Effect.FLIP_AXIS], f)[:, m:].any()


def test_chain_is_sequential_application():
    w, h = 61, 45
    f = frame(w, h, 9)
    for effects in ([Effect.GRAYSCALE], [Effect.NOISE, Effect.INVERT_COLOR], [Effect.PIXELATE, Effect.WAVE, Effect.STEP],
                    [Effect.ANAGLYPH, Effect.FLIP_AXIS, Effect.WATERCOLOR, Effect.INTERLACE, Effect.NONE]):
        cur = f
        for e in effects:
            cur = fo.apply(e, cur, 1.25)
        assert np.array_equal(fo.chain(effects, f, 1.25), cur), effects
    assert not fo.chain([], f).any()  # nothing dispatched, output untouched
    # threads only split rows
    assert np.array_equal(fo.chain([Effect.WATERCOLOR, Effect.NOISE], f, 2.0, threads=4),
                          fo.chain([Effect.WATERCOLOR, Effect.NOISE], f, 2.0))


def test_ping_pong_keeps_stale_texels_of_a_reused_texture():
    # [wave, none, flip, none] on a 40x16 frame: flip (index 2) writes intermediate 1, which still holds effect 0's
    # output outside the 16x16 square (post_processor.rs:170-186)
    w, h = 40, 16
    f = frame(w, h, 11)
    out = fo.chain([Effect.WAVE, Effect.NONE, Effect.FLIP_AXIS, Effect.NONE], f)
    wave = fo.apply(Effect.WAVE, f)
    assert np.array_equal(out[:, h:], wave[:, h:])
    assert np.array_equal(out[:, :h], wave[:, :h].transpose(1, 0, 2))


def test_postfx_oracle_pins():
    """Regression pins of the checker's own outputs (generated by tools/make_postfx_pins.py)."""
    pins = json.load(open(os.path.join(ROOT, "tests", "golden", "postfx_pins.json")))
    for p in pins["cases"]:
        f = frame(p["w"], p["h"], p["seed"])
        out = fo.chain(p["effects"], f, p["time"])
        assert hashlib.sha256(out.tobytes()).hexdigest() == p["sha256"], p
