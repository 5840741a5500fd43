"""The reference's interactive frame for ncu: python tools/perf_update.py [frames] [max_depth] - 600x600, get_simple_scene after
the first update() (eye at (5, 0, 0)), 1 spp per frame through forma_rt_render (sample index = frame)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import forma_b200
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 50
m = forma_b200.create_model(600, 600, seed=0)
m.scene.camera.orbit((0.0, 0.0, 0.0))
ks = []
for f in range(frames):
    m.render(1, sample_offset=f, max_depth=depth)
    ks.append(m.stats().kernel_ms)
print("depth", depth, "kernel_ms", [round(k, 4) for k in ks], "rays", m.stats().rays)
