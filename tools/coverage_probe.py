"""How the two scan kernels compare as an object fills more of the frame (development aid): scenes/scene_09.json's
tetrahedron from closer and closer, 1920x1080, 64 spp; coverage = fraction of primary rays that hit.
python tools/coverage_probe.py  (needs a GPU; FORMA_RT_SCAN_QUEUE=0 / 1 forces a kernel, unset = the library's choice)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200  # noqa: E402

txt = open(os.path.join(ROOT, "tests", "golden", "scenes", "scene_09.json")).read()
for dist in (40.0, 12.0, 6.0, 3.5, 2.5, 1.8):
    row = []
    for force in ("0", "1", None):
        if force is None:
            os.environ.pop("FORMA_RT_SCAN_QUEUE", None)
        else:
            os.environ["FORMA_RT_SCAN_QUEUE"] = force
        m = forma_b200.create_model(1920, 1080, seed=0)
        m.scene.load_json(txt)
        m.scene.camera.look((10.0 - dist * 0.6, 0.3, 10.0 - dist * 0.8), (10.0, 0.0, 10.0), (0.0, 1.0, 0.0), 60.0)
        ids, _ = m.primary_hits()
        best = min((m.render(64), m.stats().kernel_ms)[1] for _ in range(3))
        row.append(best)
        cov = float((ids >= 0).mean())
        rays = m.stats().rays
        m.close()
    print(f"distance {dist:5.1f}: coverage {100 * cov:5.1f} %, {rays / (1920 * 1080 * 64):.2f} rays/path, render_kernel {row[0]:7.3f} ms, "
          f"render_queue_kernel {row[1]:7.3f} ms, library's choice {row[2]:7.3f} ms", flush=True)
