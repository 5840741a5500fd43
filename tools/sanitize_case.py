"""The smallest case that runs every kernel of the tracer path once (for compute-sanitizer --tool memcheck):
render_kernel (sky + partition), render_queue_kernel, render_tree_kernel, accumulate / tonemap / reduce_tonemap,
primary_hits, the no-allocation entry point.  python tools/sanitize_case.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200  # noqa: E402
from forma_b200 import _lib  # noqa: E402

S = os.path.join(ROOT, "tests", "golden", "scenes")
w, h = 67, 21
m = forma_b200.create_model(w, h, seed=1)
m.scene.load_json(open(os.path.join(S, "scene_09.json")).read()); m.scene.camera.use_json()
m.render(3, want_linear=True)                       # render_kernel, ragged tiles
m.set_partition(1, 3); m.render(2); m.set_partition(0, 1)
m.scene.load_builtin(2)
rgb, lin = m.render(5, want_linear=True)            # render_queue_kernel + accumulate_kernel
m.primary_hits()
m.accumulate(4); m.accumulate(2)                    # progressive: prior sums
m.scene.load_json(open(os.path.join(S, "scene_01.json")).read()); m.scene.camera.use_json()
m.render(3, want_linear=True)                       # LBVH build + render_tree_kernel
m.primary_hits()
import torch
dev = torch.device("cuda", 0)
a = torch.zeros((w * h, 4), dtype=torch.float32, device=dev); b = torch.zeros_like(a)
out = torch.zeros(w * h * 3 + 8, dtype=torch.uint8, device=dev)
m.render_device(2, 0, 50, None, a.data_ptr(), 0); m.render_device(2, 2, 50, None, b.data_ptr(), 0)
m.reduce_tonemap_device([a.data_ptr(), b.data_ptr()], 4, 3, w * h - 2, out.data_ptr() + 1, None, 0)
m.tonemap_device(a.data_ptr(), 2, out.data_ptr(), 0, w * h, 0)
m.render_device_noalloc(3, 0, 50, out.data_ptr(), a.data_ptr(), 0)
m.synchronize()
torch.cuda.synchronize()
print("sanitize_case ok", int(rgb.sum()), m.stats().rays)
