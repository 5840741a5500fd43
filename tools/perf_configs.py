"""Throughput of BASELINE.json's configs on one GPU (development aid; bench.py is the contract)."""
import importlib.util, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200
from forma_b200 import _lib

def gen(n):
    spec = importlib.util.spec_from_file_location("g", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    return json.dumps(mod.generate(n))

def run(name, txt, w, h, spp, accel, builtin=None, reps=3, orbit=False):
    m = forma_b200.create_model(w, h, seed=0, accel=accel)
    if builtin is not None:
        m.scene.load_builtin(builtin)
        if orbit: m.scene.camera.orbit((0.0, 0.0, 0.0))  # the frontend's first update(): eye to (5, 0, 0) (camera.rs:97-102, Q12)
    else:
        m.scene.load_json(txt); m.scene.camera.use_json()
    best = None
    for _ in range(reps):
        m.render(spp)
        st = m.stats()
        if best is None or st.kernel_ms < best[0]: best = (st.kernel_ms, st.rays)
    print(json.dumps({"case": name, "accel": {0: "auto", 1: "brute", 2: "lbvh"}[accel], "in_use": st.accel_in_use, "w": w, "h": h, "spp": spp,
                      "kernel_ms": round(best[0], 3), "rays": best[1], "rays_per_path": round(best[1] / (w * h * spp), 3),
                      "mrays_s": round(best[1] / best[0] / 1e3, 1)}), flush=True)
    m.close()

S = os.path.join(ROOT, "tests", "golden", "scenes")
AUTO_ONLY = "--auto" in sys.argv  # skip the brute-force runs of A / B (slow, only a reference point)
which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["C", "A", "B", "D", "builtin"]
if "C" in which:
    t = open(os.path.join(S, "scene_09.json")).read()
    for a in (_lib.ACCEL_BRUTE, _lib.ACCEL_AUTO): run("C scene_09 1080p 256spp", t, 1920, 1080, 256, a)
if "A" in which:
    t = open(os.path.join(S, "scene_01.json")).read()
    for a in ((_lib.ACCEL_AUTO,) if AUTO_ONLY else (_lib.ACCEL_BRUTE, _lib.ACCEL_AUTO)): run("A scene_01 600x600 50spp", t, 600, 600, 50, a)
    run("A' scene_01 1080p 16spp", t, 1920, 1080, 16, _lib.ACCEL_AUTO)
if "B" in which:
    t = open(os.path.join(S, "scene_05.json")).read()
    for a in ((_lib.ACCEL_AUTO,) if AUTO_ONLY else (_lib.ACCEL_BRUTE, _lib.ACCEL_AUTO)): run("B' scene_05 1080p 4spp", t, 1920, 1080, 4, a)
if "builtin" in which:
    for b in (0, 2):
        run(f"builtin{b} 1080p 64spp", None, 1920, 1080, 64, _lib.ACCEL_AUTO, builtin=b)
        run(f"builtin{b} 600x600 50spp after update()", None, 600, 600, 50, _lib.ACCEL_AUTO, builtin=b, orbit=True)
if "D" in which:
    t = gen(100_000)
    run("D 100k spheres 1080p 16spp", t, 1920, 1080, 16, _lib.ACCEL_AUTO, reps=2)
