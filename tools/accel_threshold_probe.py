"""Where the LBVH starts to beat the linear scan (development aid; kBvhThreshold in context.cu): the sphere-field
generator at n = 8 .. 512 spheres + ground, 1920x1080, 16 spp, ACCEL_BRUTE against ACCEL_LBVH.
python tools/accel_threshold_probe.py  (needs a GPU)"""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200  # noqa: E402
from forma_b200 import _lib  # noqa: E402

spec = importlib.util.spec_from_file_location("g", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
gen = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gen)
for n in (8, 12, 16, 24, 32, 48, 64, 128, 256, 512):
    d = gen.generate(n, seed=7)
    # keep the field's density: shrink its footprint with n so that the camera still sees spheres, not just ground
    s = (n / 100_000.0) ** (1.0 / 3.0)
    for o in d["objects"][1:]:
        for k in ("x", "z"):
            o["position"][k] *= s * 4
        o["position"]["y"] = 0.5 + (o["position"]["y"] - 0.5) * s * 4
    d["camera"]["position"] = {"x": 0.0, "y": 10.0 * s * 4 + 1.0, "z": -80.0 * s * 4 - 2.0}
    txt = json.dumps(d)
    row = []
    for accel in (_lib.ACCEL_BRUTE, _lib.ACCEL_LBVH):
        m = forma_b200.create_model(1920, 1080, seed=0, accel=accel)
        m.scene.load_json(txt)
        m.scene.camera.use_json()
        best = min((m.render(16), m.stats().kernel_ms)[1] for _ in range(3))
        row.append(best)
        rays = m.stats().rays
        m.close()
    print(f"n = {n:4d}: {rays / (1920 * 1080 * 16):.2f} rays/path, scan {row[0]:8.3f} ms, LBVH {row[1]:8.3f} ms  -> {'LBVH' if row[1] < row[0] else 'scan'}", flush=True)
