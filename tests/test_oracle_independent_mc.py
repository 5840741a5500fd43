"""An INDEPENDENT Monte-Carlo evaluation of the reference tracer for sphere scenes - float64 numpy, vectorised over
paths, its own generator, written from the Rust source (tracer.rs:189-219 get_color, sphere.rs:23-152, utility.rs:4-54,
camera.rs:62-72) and sharing no code with oracle/forma_oracle.cpp - against the oracle's converged image.

The reference holds no image of this path and its RNG is OS-seeded, so nothing can pin the restatement bit for bit;
what this test pins is its MEANING: two independent restatements of the same Rust functions must converge to the same
image.  A slip in Schlick, refract, the metal absorption rule, the depth cut, the sky or the (h - y) row mapping moves
the mean by far more than the Monte-Carlo noise bound asserted here."""
import numpy as np
import pytest

from oracle import binding as ob


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def dot(a, b):
    return np.sum(a * b, axis=-1)


def in_unit_ball(rng, n, dims):
    """utility.rs:4-26: rejection sampling of 2 xi - 1 until p.p < 1 (z = 0 for the disc)."""
    p = np.zeros((n, 3))
    todo = np.arange(n)
    while len(todo):
        q = np.zeros((len(todo), 3))
        q[:, :dims] = 2.0 * rng.random((len(todo), dims)) - 1.0
        ok = dot(q, q) < 1.0
        p[todo[ok]] = q[ok]
        todo = todo[~ok]
    return p


def mc_image(objs, cam, w, h, spp, seed, max_depth=50):
    rng = np.random.default_rng(seed)
    f = lambda name: np.asarray(ob.cam_field(cam, name), dtype=np.float64)  # noqa: E731
    pos, llc, hor, ver, cu, cv = f("position"), f("lower_left_corner"), f("horizontal"), f("vertical"), f("u"), f("v")
    lens = float(f("lens_radius")[0])
    centers = objs["a"].astype(np.float64)
    radii = objs["b"][:, 0].astype(np.float64)
    colors = objs["color"].astype(np.float64)
    mats = objs["material"].astype(np.int64)
    fuzz = objs["fuzz"].astype(np.float64)
    ys, xs = np.mgrid[0:h, 0:w]
    img = np.zeros((h * w, 3))
    for _ in range(spp):
        n = h * w
        u = (xs.ravel() + rng.random(n)) / w                 # tracer.rs:171
        v = ((h - ys.ravel()) + rng.random(n)) / h           # tracer.rs:172: h - y, not h - 1 - y
        rd = lens * in_unit_ball(rng, n, 2)                  # camera.rs:63
        off = rd[:, :1] * cu + rd[:, 1:2] * cv
        o = pos + off
        d = llc + u[:, None] * hor + v[:, None] * ver - pos - off
        thr = np.ones((n, 3))
        alive = np.arange(n)
        col = np.zeros((n, 3))
        for depth in range(max_depth + 1):
            if not len(alive):
                break
            oo, dd = o[alive], d[alive]
            # closest sphere, t in (0.001, inf), strict; first root then second (sphere.rs:23-51)
            oc = oo[:, None, :] - centers[None]
            a = dot(dd, dd)[:, None]
            b = np.sum(oc * dd[:, None, :], -1)
            c = np.sum(oc * oc, -1) - radii[None] ** 2
            disc = b * b - a * c
            sq = np.sqrt(np.maximum(disc, 0.0))
            t1, t2 = (-b - sq) / a, (-b + sq) / a
            t = np.where(t1 > 0.001, t1, np.where(t2 > 0.001, t2, np.inf))
            t = np.where(disc > 0.0, t, np.inf)
            k = np.argmin(t, axis=1)
            tt = t[np.arange(len(alive)), k]
            miss = ~np.isfinite(tt)
            # sky (tracer.rs:213-217)
            m = alive[miss]
            s = 0.5 * (unit(dd[miss])[:, 1] + 1.0)
            col[m] = thr[m] * ((1.0 - s)[:, None] * np.ones(3) + s[:, None] * np.array([0.5, 0.7, 1.0]))
            hit = ~miss
            ids, kk, th = alive[hit], k[hit], tt[hit]
            if depth >= max_depth:      # tracer.rs:205: scatter only while depth < 50, else black
                break
            p = oo[hit] + th[:, None] * dd[hit]
            nrm = (p - centers[kk]) / radii[kk][:, None]
            din = dd[hit]
            mat = mats[kk]
            nd = np.zeros_like(din)
            keep = np.ones(len(ids), bool)
            att = colors[kk].copy()
            lam = (mat != 1) & (mat != 2)   # 0, 3 and anything else: lambertian ray (sphere.rs:60-70,84-89,147-152)
            rs = in_unit_ball(rng, len(ids), 3)
            nd[lam] = nrm[lam] + rs[lam]
            att[mat == 3] = 1.0
            me = mat == 1                   # sphere.rs:91-105
            refl_u = unit(din[me]) - 2.0 * dot(unit(din[me]), nrm[me])[:, None] * nrm[me]
            nd[me] = refl_u + fuzz[kk][me][:, None] * rs[me]
            keep[me] = dot(nd[me], nrm[me]) > 0.0
            di = mat == 2                   # sphere.rs:107-145
            if di.any():
                dv, nn = din[di], nrm[di]
                dn = dot(dv, nn)
                ln = np.linalg.norm(dv, axis=-1)
                outside = dn <= 0.0
                outward = np.where(outside[:, None], nn, -nn)
                ratio = np.where(outside, 1.0 / 1.3, 1.3)
                cosine = np.where(outside, -dn / ln, 1.3 * dn / ln)
                uv = unit(dv)
                dt = dot(uv, outward)
                dsc = 1.0 - ratio * ratio * (1.0 - dt * dt)
                refr = ratio[:, None] * (uv - dt[:, None] * outward) - np.sqrt(np.maximum(dsc, 0.0))[:, None] * outward
                r0 = ((1.0 - 1.3) / (1.0 + 1.3)) ** 2
                prob = np.where(dsc > 0.0, r0 + (1.0 - r0) * (1.0 - cosine) ** 5, 1.0)
                refl = dv - 2.0 * dn[:, None] * nn          # un-normalised direction (sphere.rs:113)
                nd[di] = np.where((rng.random(len(dv)) < prob)[:, None], refl, refr)
            thr[ids] *= att
            o[ids], d[ids] = p, nd
            alive = ids[keep]               # absorbed metal paths stay black (col is 0)
        img += col
    return (img / spp).reshape(h, w, 3)


def fuzzy_metal_scene():
    """One large metal sphere with fuzz 1 (where sphere.rs:104's absorption rule decides a third of the paths) and a
    "light" sphere (material 3: lambertian ray, attenuation one) on a lambertian ground."""
    from oracle.scene_oracle import OBJECT_DTYPE
    o = np.zeros(3, dtype=OBJECT_DTYPE)
    o["kind"] = 0
    o["id"] = np.arange(3)
    o["a"] = [(0.0, 0.0, -1.0), (0.9, -0.2, -0.4), (0.0, -100.5, -1.0)]
    o["b"][:, 0] = [0.5, 0.25, 100.0]
    o["material"] = [1, 3, 0]
    o["fuzz"] = [1.0, 0.0, 0.0]
    o["color"] = [(0.9, 0.6, 0.3), (0.2, 0.9, 0.2), (0.5, 0.5, 0.5)]
    return o


@pytest.mark.parametrize("which", [0, 2, "fuzzy-metal"])
def test_oracle_converges_to_the_independent_estimate(which):
    w, h = 48, 36
    objs = fuzzy_metal_scene() if which == "fuzzy-metal" else ob.builtin_scene(which)
    objs = objs[objs["kind"] == 0]          # spheres (scene 0's plane has orientation (0,0,0): never hit, scenes.rs:8-17)
    cam = ob.camera_new(w, h)
    spp_o, spp_m = 1024, 384
    _, lin, _ = ob.render(objs, cam, w, h, spp_o, seed=11, nthreads=8, want_rgb=False)
    mean_o = lin.astype(np.float64) / spp_o
    mean_m = mc_image(objs, cam, w, h, spp_m, seed=5)
    # per-pixel radiance is in [0, 1]: sigma of one sample <= 0.5, so the difference of the two means has
    # sigma <= 0.5 * sqrt(1/1024 + 1/384) = 0.030 per channel; systematic slips are several times that
    diff = mean_m - mean_o
    assert np.abs(diff).max() < 0.15, np.abs(diff).max()
    assert np.sqrt((diff ** 2).mean()) < 0.02, np.sqrt((diff ** 2).mean())
    # image-wide bias per channel: averaging 1728 pixels leaves sigma <= 0.001
    assert np.abs(diff.mean(axis=(0, 1))).max() < 0.004, diff.mean(axis=(0, 1))
    # and per region, where a material-specific slip would show: split the frame into a 4 x 4 grid
    blocks = diff.reshape(4, h // 4, 4, w // 4, 3).mean(axis=(1, 3))
    assert np.abs(blocks).max() < 0.008, np.abs(blocks).max()
