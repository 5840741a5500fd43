"""CPU tests of the fragment-shader checker (oracle/wgsl_rt_oracle.cpp) and of the forma_frag C ABI surface.

The reference holds no golden image of rt.wgsl / rt2.wgsl and cannot run here (wgpu), so the restatement is checked
against an INDEPENDENT float64 numpy evaluation of both shaders (written from the WGSL, vectorised, true pow),
hand-derived values and structural properties; its outputs are then pinned (tests/golden/wgsl_rt_pins.json) so that
it cannot drift silently.
"""
import ctypes
import hashlib
import json
import math
import os
import re

import numpy as np
import pytest

from forma_b200 import _lib
from forma_b200 import frag
from oracle import wgsl_rt_binding as wb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- ABI surface ----------------------------------------------------------------------------------------------------
def test_library_exports_every_forma_frag_symbol():
    hdr = open(os.path.join(ROOT, "include", "forma_frag.h")).read()
    declared = set(re.findall(r"FORMA_FRAG_API[^;]*?\b(forma_frag_\w+)\s*\(", hdr))
    assert len(declared) == 8
    assert declared == set(_lib.SYMBOLS_FRAG), declared ^ set(_lib.SYMBOLS_FRAG)
    L = _lib.lib()
    for name in declared:
        assert getattr(L, name)


def test_frag_create_without_a_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = ctypes.c_void_p()
    rc = _lib.lib().forma_frag_create(0, 8, 8, ctypes.byref(h))
    assert rc == _lib.ECUDA and not h.value
    assert b"no CPU fallback" in _lib.lib().forma_rt_last_error()
    with pytest.raises(_lib.FormaError):
        frag.State(8, 8)


def test_light_direction_host_helper_matches_the_checker():
    # forma_frag_light_direction needs no device; both sides use the host's cos / sin in double, rounded to f32
    for t in (0.0, 1.0, 3.0, 15.70796, 31.4159, 1e3, -7.25, 12345.678):
        a, b = frag.light_direction(t), wb.light2(t)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), t
        assert abs(float(np.linalg.norm(a.astype(np.float64))) - 1.0) < 2e-7
        ang = float(np.float32(t) / np.float32(10.0))
        ref = np.array([math.cos(ang), 0.75, math.sin(ang)])
        assert np.allclose(a, ref / np.linalg.norm(ref), atol=2e-7)


# ---- the fixed pow ---------------------------------------------------------------------------------------------------
def test_log2_exp2_accuracy_and_special_values():
    rng = np.random.default_rng(1)
    xs = np.concatenate([np.exp2(rng.uniform(-126, 127, 4000)), 1 + rng.uniform(-1e-3, 1e-3, 2000),
                         rng.uniform(0, 1, 2000), [1.0, 2.0, 0.5, 1e-40, 3e-45]]).astype(np.float32)
    for x in xs:
        got, ref = wb.log2(float(x)), math.log2(float(x))
        assert abs(got - ref) <= 2.5e-7 * max(1.0, abs(ref)), (x, got, ref)
    assert wb.log2(1.0) == 0.0 and wb.log2(8.0) == 3.0 and wb.log2(0.25) == -2.0
    assert wb.log2(0.0) == -math.inf and wb.log2(-0.0) == -math.inf and wb.log2(math.inf) == math.inf
    assert math.isnan(wb.log2(-1.0)) and math.isnan(wb.log2(math.nan))
    for x in np.concatenate([rng.uniform(-149, 128, 6000), np.arange(-150, 129), [0.5, -0.5, 1e-8]]).astype(np.float32):
        got, ref = wb.exp2(float(x)), 2.0 ** float(x)
        if ref < 2.0 ** -126:   # subnormal results: absolute bound of one subnormal step
            assert abs(got - ref) <= 1.5 * 2.0 ** -149, (x, got, ref)
        elif ref < 3.4e38:
            assert abs(got - ref) <= 4e-7 * ref, (x, got, ref)
    assert wb.exp2(0.0) == 1.0 and wb.exp2(10.0) == 1024.0 and wb.exp2(-3.0) == 0.125
    assert wb.exp2(128.0) == math.inf and wb.exp2(math.inf) == math.inf
    assert wb.exp2(-151.0) == 0.0 and wb.exp2(-math.inf) == 0.0 and math.isnan(wb.exp2(math.nan))


def test_pow_on_the_shaders_exponents():
    rng = np.random.default_rng(2)
    for y in (0.4, 5.0, 250.0, 5000.0, 1 / 2.2, 1 / 2.4):
        yf = float(np.float32(y))
        xs = np.concatenate([rng.uniform(0, 1, 1500), 1 - 10 ** rng.uniform(-7, -1, 1500), rng.uniform(1, 50, 200) if y <= 5 else []])
        for x in xs.astype(np.float32):
            ref = float(x) ** yf
            if 1e-30 < ref < 1e30:
                # the rounding of y * log2(x) is amplified by |y log2 x| <= ~100 here
                assert abs(wb.pow(float(x), yf) - ref) <= 1.2e-5 * ref, (x, y)
        assert wb.pow(0.0, yf) == 0.0 and wb.pow(1.0, yf) == 1.0
        assert math.isnan(wb.pow(-0.5, yf)) and math.isnan(wb.pow(math.nan, yf))


# ---- an independent float64 evaluation of the two shaders ------------------------------------------------------------
def _norm(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def _dot(a, b):
    return np.sum(a * b, axis=-1)


def _rays(w, h, res, aspect_x, aspect_on_uv, origin, z):
    ys, xs = np.mgrid[0:h, 0:w]
    ux = 2.0 * (xs + 0.5) / res[0] - 1.0
    uy = -(2.0 * (ys + 0.5) / res[1] - 1.0)
    d = np.stack([aspect_x * ux, uy, np.full_like(ux, z)], -1).reshape(-1, 3)
    return np.tile(np.asarray(origin, float), (w * h, 1)), _norm(d)


def _sphere_t(o, d, c, r):
    op = c - o
    b = _dot(op, d)
    det = b * b - _dot(op, op) + r * r
    sq = np.sqrt(np.maximum(det, 0))
    t = np.where(b - sq < 0, b + sq, b - sq)
    return np.where((det < 0) | (t < 0), np.inf, t)


def numpy_rt1(w, h, res, cap=40):
    """rt.wgsl in float64; returns (linear rgb (h, w, 3), finished mask)."""
    o, d = _rays(w, h, res, 4.0 / 3.0, False, (0, 1, 4), -1.5)
    n = o.shape[0]
    spheres = [((-2, 1, 0), (1, 0, 0), 0.04), ((0, 1, 0), (0, 1, 0), 0.1), ((2, 1, 0), (0, 0, 1), 0.5)]
    sun = _norm(np.array([1.0, 0.5, 0.0]))
    sun_sky = _norm(np.array([1.0, 0.5, 0.5]))

    def scene(o, d):
        with np.errstate(divide="ignore", invalid="ignore"):
            tp = np.where(d[:, 1] > 0, np.inf, -o[:, 1] / d[:, 1])
        tp = np.where(np.isnan(tp), np.inf, tp)
        t = np.where(tp < 1e10, tp, np.inf)
        nrm = np.tile([0.0, 1.0, 0.0], (len(o), 1))
        col = np.tile([0.5, 0.4, 0.3], (len(o), 1))
        f0 = np.where(np.isfinite(t), 0.04, -1.0)
        for c, k, f in spheres:
            c = np.asarray(c, float)
            ts = _sphere_t(o, d, c, 1.0)
            better = ts < np.where(np.isfinite(t), t, 1e10)
            t = np.where(better, ts, t)
            nrm = np.where(better[:, None], o + np.where(better, ts, 0)[:, None] * d - c, nrm)
            col = np.where(better[:, None], np.asarray(k, float), col)
            f0 = np.where(better, f, f0)
        return t, nrm, col, f0

    accum, att = np.zeros((n, 3)), np.ones((n, 3))
    alive = np.ones(n, bool)
    for _ in range(cap):
        idx = np.nonzero(alive)[0]
        if not len(idx):
            break
        oo, dd = o[idx], d[idx]
        t, nrm, col, f0 = scene(oo, dd)
        hit = f0 >= 0
        # miss -> sky
        m = idx[~hit]
        dm = dd[~hit]
        tr = np.clip((dm[:, 1] - 0.02) / (0.5 - 0.02), 0, 1)
        tr = (tr * tr * (3 - 2 * tr)) ** 0.4
        sky = 2.0 * (np.array([0.52, 0.77, 1.0]) * (1 - tr[:, None]) + np.array([0.12, 0.43, 1.0]) * tr[:, None])
        accum[m] += sky + (np.abs(dm @ sun_sky) ** 5000.0)[:, None]
        alive[m] = False
        # hit
        k = idx[hit]
        oh, dh, th, nh, ch, fh = oo[hit], dd[hit], t[hit], nrm[hit], col[hit], f0[hit]
        f = (1 - np.clip(_dot(nh, -dh), 0, 1)) ** 5 * (1 - fh) + fh
        p = oh + th[:, None] * dh
        _, _, _, sf0 = scene(p + 4e-4 * sun, np.tile(sun, (len(p), 1)))
        inc = np.where(sf0 < 0, np.clip(nh @ sun, 0, 1) * 1e3, 0.0)
        accum[k] += ch + (1 - f)[:, None] * att[k] * ch * inc[:, None]
        att[k] *= f[:, None]
        r = dh - 2 * _dot(nh, dh)[:, None] * nh
        o[k], d[k] = p + 4e-4 * r, r
    return (accum / 1000.0).reshape(h, w, 3), (~alive).reshape(h, w)


def numpy_rt2(w, h, res, time):
    """rt2.wgsl in float64 (pow(l, 2.0) as l * l); returns linear rgb (h, w, 3)."""
    o, d = _rays(w, h, res, res[0] / res[1], True, (0, 2.5, 12), -1.0)
    n = o.shape[0]
    prims = [(2.0, (-4, 3, 0), (1, 0, 0.2), 1.0, 0.001), (3.0, (4, 3, 0), (0, 0.2, 1), 1.0, 0.0),
             (1.0, (6, 1, 4), (0, 1, 0.2), 0.5, 0.1), (1.0, (0.5, 1, 6), (1, 1, 1), 0.5, 0.25)]  # s1, s2, s4, s3
    a = float(np.float32(time) / np.float32(10.0))
    light = _norm(np.array([math.cos(a), 0.75, math.sin(a)]))

    def trace(o, d):
        with np.errstate(divide="ignore", invalid="ignore"):
            ln = -o[:, 1] / d[:, 1]
        hit = ~(ln < 0)
        nrm = np.tile([0.0, 1.0, 0.0], (len(o), 1))
        col = np.ones((len(o), 3))
        dif, spec = np.ones(len(o)), np.zeros(len(o))
        for r, c, k, df, sp in prims:
            c = np.asarray(c, float)
            ts = _sphere_t(o, d, c, r)
            h2 = np.isfinite(ts)
            ln = np.where(h2, ts, ln)
            nrm = np.where(h2[:, None], (o + np.where(h2, ts, 0)[:, None] * d - c) / r, nrm)
            col = np.where(h2[:, None], np.asarray(k, float), col)
            dif, spec = np.where(h2, df, dif), np.where(h2, sp, spec)
            hit |= h2
        return hit, ln, nrm, col, dif, spec

    color, mask = np.zeros((n, 3)), np.ones((n, 3))
    alive = np.ones(n, bool)
    ambient = np.array([0.6, 0.8, 1.0]) * 100.0 / 2.2
    for _ in range(17):
        idx = np.nonzero(alive)[0]
        if not len(idx):
            break
        oo, dd = o[idx], d[idx]
        hit, ln, nrm, col, dif, spec = trace(oo, dd)
        m = idx[~hit]
        spot = 1e6 * np.abs(dd[~hit] @ light) ** 250.0
        color[m] += mask[m] * (ambient + spot[:, None])
        alive[m] = False
        k = idx[hit]
        oh, dh, lh, nh, ch, dfh, sph = oo[hit], dd[hit], ln[hit], nrm[hit], col[hit], dif[hit], spec[hit]
        r0 = ch * sph[:, None]
        hv = np.clip(_dot(nh, -dh), 0, 1)
        fres = r0 + (1 - r0) * ((1 - hv) ** 5.0)[:, None]
        mask[k] *= fres
        p = oh + lh[:, None] * dh
        shit = trace(p + 1e-3 * light, np.tile(light, (len(p), 1)))[0]
        with np.errstate(divide="ignore", invalid="ignore"):
            add = (np.clip(nh @ light, 0, 1) * 100.0 * dfh)[:, None] * ch * (1 - fres) * mask[k] / fres
        color[k] += np.where(shit[:, None], 0.0, add)
        r = dh - 2 * _dot(nh, dh)[:, None] * nh
        o[k], d[k] = p + 1e-3 * r, r
    with np.errstate(invalid="ignore"):
        return ((color * 1e-2) ** (1 / 2.2)).reshape(h, w, 3)


def _agreement(lin32, ref64, ok=None, rtol=2e-3, atol=2e-4):
    good = np.isfinite(ref64).all(-1) & np.isfinite(lin32).all(-1)
    if ok is not None:
        good &= ok
    close = (np.abs(lin32 - ref64) <= atol + rtol * np.abs(ref64)).all(-1)
    return good.mean(), (close & good).sum() / max(1, good.sum())


def test_rt_wgsl_restatement_against_float64_numpy():
    w, h = 200, 150
    _, lin, rays, nb = wb.render(wb.RT, w, h, resolution=(w, h), want_bounces=True)
    ref, done = numpy_rt1(w, h, (w, h))
    covered, agree = _agreement(lin[..., :3].astype(np.float64), ref, done)
    # f32 vs f64 paths part ways only at silhouette / shadow edges and on long mirror chains
    assert covered > 0.99 and agree > 0.985, (covered, agree)
    assert (lin[..., 3] == 1.0).all()
    # rays = loop trips + shadow queries; a path of b bounces that ends on the sky costs (b + 1) + b
    ended = nb < 1000
    assert rays == int((2 * nb.astype(np.int64) + ended).sum())
    assert nb.max() >= 3 and nb.min() == 0


def test_rt2_wgsl_restatement_against_float64_numpy():
    w, h = 200, 150
    for t in (0.0, 3.0, 20.0):
        _, lin, rays, nb = wb.render(wb.RT2, w, h, time=t, resolution=(w, h), want_bounces=True)
        ref = numpy_rt2(w, h, (w, h), t)
        covered, agree = _agreement(lin[..., :3].astype(np.float64), ref, rtol=5e-3)
        assert covered > 0.99 and agree > 0.98, (t, covered, agree)
        ended = nb < 17
        assert rays == int((2 * nb.astype(np.int64) + ended).sum())
        assert nb.max() <= 17


def test_hand_derived_pixels():
    # rt.wgsl, 1600 x 1200 constants: the centre column looks along -z from (0, 1, 4); a pixel just above centre
    # hits the middle sphere (green, f0 0.1) head on: first bounce adds (0, 1, 0) and the mirror ray leaves to the sky
    _, lin, _, nb = wb.render(wb.RT, 1600, 1200, want_bounces=True)
    assert nb[598, 800] >= 1 and lin[598, 800, 1] > lin[598, 800, 0] and lin[598, 800, 1] > lin[598, 800, 2]
    # top rows see only sky: no bounce, and sky = 2 * mix(...) / 1000 with transition 1 well above the horizon
    assert nb[0].max() == 0
    d = np.array([4 / 3 * (2 * 800.5 / 1600 - 1), -(2 * 0.5 / 1200 - 1), -1.5])
    d /= np.linalg.norm(d)
    assert d[1] > 0.5
    assert np.allclose(lin[0, 800, :3], 2 * np.array([0.12, 0.43, 1.0]) / 1000, rtol=1e-5)
    # rt2.wgsl: the top-left pixel misses everything: pow((ambient + spot) * EXPOSURE, 1 / GAMMA)
    _, lin2, _, nb2 = wb.render(wb.RT2, 1600, 1200, time=0.0, want_bounces=True)
    assert nb2[0, 0] == 0
    amb = np.array([0.6, 0.8, 1.0]) * 100 / 2.2
    assert np.allclose(lin2[0, 0, :3], (amb * 1e-2) ** (1 / 2.2), rtol=2e-5)


def test_unorm_and_srgb_targets():
    w, h = 96, 72
    u8, lin, _ = wb.render(wb.RT2, w, h, time=3.0, resolution=(w, h))
    x = lin[..., :3]
    ref = (np.clip(np.where(np.isnan(x), 0, x), 0, 1).astype(np.float32) * np.float32(255) + np.float32(0.5)).astype(np.uint8)
    assert np.array_equal(u8[..., :3], ref) and (u8[..., 3] == 255).all()
    s8, lin_s, _ = wb.render(wb.RT2, w, h, time=3.0, resolution=(w, h), srgb=True)
    assert np.array_equal(lin_s.view(np.uint32), lin.view(np.uint32))
    c = np.clip(np.where(np.isnan(x), 0, x).astype(np.float64), 0, 1)
    curve = np.where(c <= 0.0031308, 12.92 * c, 1.055 * c ** (1 / 2.4) - 0.055)
    assert np.abs(s8[..., :3].astype(int) - np.floor(curve * 255 + 0.5).astype(int)).max() <= 1
    assert (s8[..., :3].astype(int) >= u8[..., :3].astype(int)).all()


def test_crop_semantics_and_threads():
    # a frame smaller than `resolution` is the top-left crop of the full image (fragment positions do not change)
    full = wb.render(wb.RT, 160, 120, resolution=(160, 120))[1]
    crop = wb.render(wb.RT, 50, 30, resolution=(160, 120))[1]
    assert np.array_equal(crop.view(np.uint32), full[:30, :50].view(np.uint32))
    a = wb.render(wb.RT2, 160, 120, time=5.0, threads=1)
    b = wb.render(wb.RT2, 160, 120, time=5.0, threads=5)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)) and a[2] == b[2]


def test_wgsl_rt_oracle_pins():
    pins = json.load(open(os.path.join(ROOT, "tests", "golden", "wgsl_rt_pins.json")))
    assert len(pins["cases"]) >= 8
    for c in pins["cases"]:
        u8, lin, rays = wb.render(c["shader"], c["w"], c["h"], time=c["time"], srgb=c["srgb"],
                                  resolution=tuple(c["resolution"]) if c["resolution"] else None, threads=4)
        assert hashlib.sha256(lin.tobytes()).hexdigest() == c["sha256_linear"], c
        assert hashlib.sha256(u8.tobytes()).hexdigest() == c["sha256_rgba8"], c
        assert rays == c["rays"], c
