"""One render of one case (for ncu): python tools/perf_one.py A|B|D [spp]"""
import os, sys
sys.argv = [sys.argv[0]] + sys.argv[1:]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import forma_b200
which = sys.argv[1]
S = os.path.join(ROOT, "tests", "golden", "scenes")
name, w, h, spp = {"A": ("scene_01", 600, 600, 50), "B": ("scene_05", 1920, 1080, 4), "C": ("scene_09", 1920, 1080, 256),
                   "D": (None, 1920, 1080, 16), "S0": ("builtin0", 1920, 1080, 64), "S2": ("builtin2", 1920, 1080, 64),
                   "S0r": ("builtin0", 600, 600, 50), "S2r": ("builtin2", 600, 600, 50)}[which]
if len(sys.argv) > 2: spp = int(sys.argv[2])
m = forma_b200.create_model(w, h, seed=0)
if name and name.startswith("builtin"):
    m.scene.load_builtin(int(name[-1]))
    m.scene.camera.orbit((0.0, 0.0, 0.0))  # the frontend's first update(): eye to (5, 0, 0) (camera.rs:97-102, Q12)
    txt = None
elif name:
    txt = open(os.path.join(S, name + ".json")).read()
else:
    import importlib.util, json
    spec = importlib.util.spec_from_file_location("g", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    txt = json.dumps(mod.generate(100_000))
if txt: m.scene.load_json(txt); m.scene.camera.use_json()
for _ in range(2):
    m.render(spp)
    st = m.stats()
print(which, st.kernel_ms, st.rays)
