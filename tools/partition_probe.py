"""Per-ray cost of config D's tree kernel when ONE GPU renders only the tiles of rank 0 of an N-way tile-interleaved
split (development aid): separates what the split itself costs (locality of the rays in flight) from multi-GPU
effects.  python tools/partition_probe.py [spp]  (needs a GPU)"""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 32
spec = importlib.util.spec_from_file_location("g", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
mod = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mod)
m = forma_b200.create_model(1920, 1080, seed=0)
m.scene.load_json(json.dumps(mod.generate(100_000)))
m.scene.camera.use_json()
for world in (1, 2, 4, 8):
    m.set_partition(0, world)
    best = None
    for _ in range(2):
        m.render(spp)
        st = m.stats()
        if best is None or st.kernel_ms < best[0]:
            best = (st.kernel_ms, st.rays)
    print(f"rank 0 of {world}: kernel {best[0]:.2f} ms, {best[1] / 1e6:.1f} Mrays, {best[1] / best[0] / 1e3:.0f} Mrays/s", flush=True)
