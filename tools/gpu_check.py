"""Quick GPU-vs-oracle check used during development (the real parity suite is tests/ -m gpu)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import forma_b200  # noqa: E402
from oracle import binding as ob  # noqa: E402

SCENES = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes")


def compare(name, m, objs, cam, w, h, spp, depth=50, seed=0):
    t0 = time.time()
    rgb_o, lin_o, rays_o = ob.render(objs, cam, w, h, spp, max_depth=depth, seed=seed, nthreads=os.cpu_count())
    t_cpu = time.time() - t0
    rgb_g, lin_g = m.render(spp, max_depth=depth, want_linear=True)
    st = m.stats()
    same_lin = np.array_equal(lin_o.view(np.uint32), lin_g[..., :3].view(np.uint32))
    nan_ok = np.array_equal(np.isnan(lin_o), np.isnan(lin_g[..., :3]))
    md = float(np.nanmax(np.abs(lin_o - lin_g[..., :3])))
    print(json.dumps({"case": name, "w": w, "h": h, "spp": spp, "rays_oracle": rays_o, "rays_gpu": int(st.rays),
                      "linear_bit_exact": bool(same_lin), "nan_pattern_equal": bool(nan_ok), "max_abs_diff": md,
                      "rgb_equal": bool(np.array_equal(rgb_o, rgb_g)), "kernel_ms": round(st.kernel_ms, 3),
                      "gpu_mrays_s": round(st.rays / st.kernel_ms / 1e3, 1), "cpu_mrays_s": round(rays_o / t_cpu / 1e6, 2)}),
          flush=True)


def main():
    w, h = 320, 180
    for which in (0, 2):
        m = forma_b200.create_model(w, h, seed=7)
        m.scene.load_builtin(which)
        objs = ob.builtin_scene(which)
        cam = ob.camera_new(w, h)
        compare(f"builtin{which}", m, objs, cam, w, h, 8, seed=7)
        if which == 0:  # the frontend's per-frame rotate() makes the plane live (macroquad.rs:12,67)
            m.scene.objects[0].rotate((-1.0, 0.0, 0.0))
            objs["b"][0] = (-1.0, 0.0, 0.0)
            m.scene.camera.orbit((0.3, 0.2, 0.0))
            ob.camera_orbit(cam, (0.3, 0.2, 0.0))
            compare("builtin0+plane+orbit", m, objs, cam, w, h, 8, seed=7)
        m.close()
    from oracle import scene_oracle as so
    for sc in ("scene_09", "scene_05", "scene_01"):
        path = os.path.join(SCENES, sc + ".json")
        if not os.path.exists(path):
            continue
        txt = open(path).read()
        d = so.load(path)
        objs = so.flatten(d)
        fr, at, up, fov = so.json_camera(d)
        cam = ob.camera_look(fr, at, up, fov, w / h)
        m = forma_b200.create_model(w, h, seed=3)
        m.scene.load_json(txt)
        m.scene.camera.use_json()
        compare(sc, m, objs, cam, w, h, 4, seed=3)
        ids_o, t_o = ob.primary_hits(objs, cam, w, h)
        ids_g, t_g = m.primary_hits()
        print(json.dumps({"case": sc + " primary", "ids_equal": bool(np.array_equal(ids_o, ids_g)),
                          "t_bit_exact": bool(np.array_equal(t_o.view(np.uint32), t_g.view(np.uint32))),
                          "hits": int((ids_g >= 0).sum())}), flush=True)
        m.close()
    # headline config C timing: scene_09 1080p 256 spp
    path = os.path.join(SCENES, "scene_09.json")
    m = forma_b200.create_model(1920, 1080, seed=0)
    m.scene.load_json(open(path).read())
    m.scene.camera.use_json()
    for spp in (16, 256, 256):
        m.render(spp)
        st = m.stats()
        print(json.dumps({"case": "scene_09 1080p", "spp": spp, "kernel_ms": round(st.kernel_ms, 3),
                          "total_ms": round(st.total_ms, 3), "rays": int(st.rays),
                          "mrays_s": round(st.rays / st.kernel_ms / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    main()
