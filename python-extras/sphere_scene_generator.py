"""Seeded N-sphere scene in the reference's scene schema (the dataclasses of the reference's
python-extras/scene_generator.py:6-52: Vec3 / Quaternion / Camera / Light / SceneObject / Scene),
with the optional per-object `rt` block of DESIGN.md "Scene mapping".

The reference's own generator only emits a 10x10 grid of cubes (scene_generator.py:67-86); BASELINE
config D ("synthetic 100k-sphere scene ... LBVH stress") is authored here: N spheres, centres uniform
in [-50,50] x [0.5,20] x [-50,50], radius U[0.2,0.6], 70 % lambertian / 20 % metal (fuzz U[0,0.5]) /
10 % dielectric, colour U[0.1,0.9]^3, plus a ground sphere of radius 1000; camera (0,10,-80) looking
at the origin, fov 60.
"""
import json
import math
import sys

import numpy as np


def vec3(x, y, z):
    return {"x": float(x), "y": float(y), "z": float(z)}


def quat(x, y, z, w):
    return {"x": float(x), "y": float(y), "z": float(z), "w": float(w)}


def generate(n=100_000, seed=1234, ground=True):
    rng = np.random.default_rng(seed)
    pos = np.stack([rng.uniform(-50, 50, n), rng.uniform(0.5, 20, n), rng.uniform(-50, 50, n)], axis=1)
    rad = rng.uniform(0.2, 0.6, n)
    kind = rng.random(n)
    col = rng.uniform(0.1, 0.9, (n, 3))
    fuzz = rng.uniform(0.0, 0.5, n)
    ident = quat(0, 0, 0, 1)
    objects = []
    if ground:
        objects.append({"mesh": "sphere", "material": "DiffuseColorMaterial", "position": vec3(0, -1000, 0),
                        "rotation": ident, "scale": vec3(2000, 2000, 2000),
                        "rt": {"material": "lambertian", "color": [0.5, 0.5, 0.5], "fuzz": 0.0}})
    for i in range(n):
        m = "lambertian" if kind[i] < 0.7 else ("metal" if kind[i] < 0.9 else "dielectric")
        d = 2.0 * rad[i]
        objects.append({"mesh": "sphere", "material": "DiffuseColorMaterial", "position": vec3(*pos[i]),
                        "rotation": ident, "scale": vec3(d, d, d),
                        "rt": {"material": m, "color": [round(float(c), 6) for c in col[i]],
                               "fuzz": round(float(fuzz[i]), 6) if m == "metal" else 0.0}})
    th = math.atan2(10.0, 80.0)  # pitch down so that R*(0,0,1) points from (0,10,-80) to the origin
    camera = {"position": vec3(0, 10, -80), "rotation": quat(math.sin(th / 2), 0, 0, math.cos(th / 2)), "fov": 60.0}
    light = {"color": vec3(0, 0, 0), "intensity": 1.0, "position": vec3(0, 0, 0), "rotation": quat(0, 0, 0, 0),
             "scale": vec3(0, 0, 0)}
    return {"name": f"spheres_{n}", "camera": camera, "lights": [light], "objects": objects}


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
    out = sys.argv[2] if len(sys.argv) > 2 else f"spheres_{n}.json"
    with open(out, "w") as f:
        json.dump(generate(n), f)
    print(f"Scene saved to {out}")
