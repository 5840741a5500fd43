"""Latency of the reference's interactive frame (tracer::update: orbit + one 1-spp frame into a host buffer) through
the C ABI, with and without the per-frame objects[0] translate / rotate of frontend/macroquad.rs:66-67.
python tools/update_latency.py [width height]  (needs a GPU)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import forma_b200  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (600, 600)
for move in (False, True):
    m = forma_b200.create_model(w, h, seed=0)
    wall, kern, total = [], [], []
    for f in range(40):
        t0 = time.perf_counter()
        if move:
            m.scene.objects[0].translate((-1.0, 0.0, 0.0))
            m.scene.objects[0].rotate((-1.0, 0.01 * f, 0.0))
        forma_b200.update(m, 0, 0.016)
        wall.append((time.perf_counter() - t0) * 1e3)
        st = m.stats()
        kern.append(st.kernel_ms)
        total.append(st.total_ms)
    print(f"{w}x{h} objects[0] moved per frame: {move}: wall {np.median(wall[5:]):.3f} ms, device total "
          f"{np.median(total[5:]):.3f} ms, kernel {np.median(kern[5:]):.3f} ms")
