"""the -fmad=true build against the oracle on config B (ids must match exactly, t within 1e-4 relative)
and its throughput.  Run with the FMA library copied over fo-rma_b200/libforma_rt.so (development aid; the
shipped build is the exact one)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import forma_b200
from oracle import binding as ob
from conftest import load_scene_pair

w, h = 1920, 1080
txt, objs, cam = load_scene_pair("scene_05", w, h)
m = forma_b200.create_model(w, h, seed=0)
m.scene.load_json(txt); m.scene.camera.use_json()
ids, t = m.primary_hits()
ids_o, t_o = ob.primary_hits(objs, cam, w, h)
hit = ids_o >= 0
rel = np.abs(t[hit] - t_o[hit]) / np.abs(t_o[hit])
out = {"config_B_ids_equal": bool(np.array_equal(ids, ids_o)), "config_B_id_mismatches": int((ids != ids_o).sum()),
       "config_B_t_max_rel": float(rel.max()), "config_B_t_bit_equal_frac": float((t[hit].view(np.uint32) == t_o[hit].view(np.uint32)).mean())}
ws, hs, spp = 320, 180, 8
txt, objs, cam = load_scene_pair("scene_01", ws, hs)
m = forma_b200.create_model(ws, hs, seed=42)
m.scene.load_json(txt); m.scene.camera.use_json()
rgb, lin = m.render(spp, want_linear=True)
rgb_o, lin_o, rays_o = ob.render(objs, cam, ws, hs, spp, seed=42, nthreads=os.cpu_count())
out["scene_01_320x180x8_rgb_equal_frac"] = float((rgb == rgb_o).all(axis=2).mean())
out["scene_01_rays"] = [int(m.stats().rays), int(rays_o)]
txt = open(os.path.join(ROOT, "tests", "golden", "scenes", "scene_09.json")).read()
m = forma_b200.create_model(w, h, seed=0)
m.scene.load_json(txt); m.scene.camera.use_json()
best = min((m.render(256), m.stats().kernel_ms)[1] for _ in range(3))
out["config_C_kernel_ms"] = round(best, 3)
for which in (0, 2):
    m.scene.load_builtin(which); m.scene.camera.reset()
    best = min((m.render(64), m.stats().kernel_ms)[1] for _ in range(3))
    out[f"builtin{which}_1080p_64spp_kernel_ms"] = round(best, 3)
print(json.dumps(out))
