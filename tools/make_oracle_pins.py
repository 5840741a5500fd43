"""Writes tests/golden/oracle_pins.json: SHA-256 of the oracle's outputs on fixed (scene, size, spp, seed) cases.
These are NOT reference outputs (the Rust reference cannot be built here): they pin the oracle
against accidental change, so that a green GPU-vs-oracle suite keeps meaning what it meant when the restatement
was reviewed against the cited reference lines.  Regenerate only together with a reviewed oracle change."""
import hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle import binding as ob
from conftest import load_scene_pair

CASES = [("builtin0", 64, 40, 3, 7), ("builtin1", 48, 32, 2, 8), ("builtin2", 64, 40, 3, 9),
         ("scene_01", 64, 36, 2, 42), ("scene_05", 48, 27, 1, 42), ("scene_09", 80, 45, 4, 42)]


def run(name, w, h, spp, seed):
    if name.startswith("builtin"):
        objs, cam = ob.builtin_scene(int(name[-1])), ob.camera_new(w, h)
    else:
        _, objs, cam = load_scene_pair(name, w, h)
    rgb, lin, rays = ob.render(objs, cam, w, h, spp, seed=seed)
    ids, t = ob.primary_hits(objs, cam, w, h)
    lin = np.where(np.isnan(lin), np.float32(0), lin)  # NaN payloads are platform-specific
    return {"rgb": hashlib.sha256(rgb.tobytes()).hexdigest(), "linear": hashlib.sha256(lin.tobytes()).hexdigest(),
            "ids": hashlib.sha256(ids.tobytes()).hexdigest(), "t": hashlib.sha256(t.tobytes()).hexdigest(), "rays": rays}


if __name__ == "__main__":
    pins = {f"{n} {w}x{h} {spp}spp seed{seed}": run(n, w, h, spp, seed) for n, w, h, spp, seed in CASES}
    out = os.path.join(ROOT, "tests", "golden", "oracle_pins.json")
    json.dump(pins, open(out, "w"), indent=1)
    print("wrote", out, len(pins), "cases")
