"""Config C (scene_09, 1080p, 256 spp) when ONE GPU renders only the tiles of rank r of an N-way tile-interleaved split
into LOCAL memory (development aid): separates what the partition costs render_kernel from remote-store effects.
python tools/partition_probe_c.py  (needs a GPU)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200  # noqa: E402

m = forma_b200.create_model(1920, 1080, seed=0)
m.scene.load_json(open(os.path.join(ROOT, "tests", "golden", "scenes", "scene_09.json")).read())
m.scene.camera.use_json()
for world in (1, 2, 4, 8):
    for rank in sorted({0, world - 1}):
        m.set_partition(rank, world)
        best = None
        for _ in range(3):
            m.render(256)
            st = m.stats()
            if best is None or st.kernel_ms < best[0]:
                best = (st.kernel_ms, st.rays)
        print(f"rank {rank} of {world}: kernel {best[0]:.3f} ms (ideal {3.87 / world:.3f}), {best[1] / 1e6:.1f} Mrays", flush=True)
