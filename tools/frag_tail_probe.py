"""Where rt.wgsl's frame time goes: kernel time of top-left crops of the 1600 x 1200 image (development aid).

The crops that contain one of the three pixels whose path runs all 1000 loop trips, (1131, 656), (723, 785) and
(876, 785), cost ~0.45 ms whatever their size; the others cost what their pixel count predicts.
python tools/frag_tail_probe.py  (needs a GPU)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from forma_b200.frag import Shader, State  # noqa: E402

for (w, h) in ((1600, 600), (1600, 1200), (1140, 660), (32, 800), (730, 790), (16, 8)):
    d = torch.empty((h, w, 4), dtype=torch.uint8, device="cuda:0")
    with State(w, h, shader=Shader.RT) as st:
        ts = []
        for i in range(8):
            st.render_device(d.data_ptr(), None, 0.0)
            st.synchronize()
            ts.append(st.stats().kernel_ms)
        print(f"{w}x{h}: kernel_ms {min(ts):.4f}, rays {st.stats().rays}")
