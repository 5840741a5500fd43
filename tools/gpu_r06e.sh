#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 900 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r06e_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r06e_pytest.log
rm -f gpurun_out/r06e_perf.txt
bash tools/gpu_variants.sh r06e default nonear
python - <<'PY' | tee gpurun_out/r06e_walkstats.txt
import json, importlib.util, os, sys
sys.path.insert(0, ".")
import forma_b200
spec = importlib.util.spec_from_file_location("g", "python-extras/sphere_scene_generator.py"); g = importlib.util.module_from_spec(spec); spec.loader.exec_module(g)
txt = json.dumps(g.generate(100000))
for lib in (None, "fo-rma_b200/variants/nonear.so"):
    pass
m = forma_b200.create_model(1920, 1080, seed=0)
m.scene.load_json(txt); m.scene.camera.use_json()
ws = m.walk_stats(16)
print("with the near bound: visits per ray %.2f, leaf tests per ray %.2f" % (ws.node_visits / ws.rays, ws.leaf_tests / ws.rays))
PY
FORMA_RT_LIB=$PWD/fo-rma_b200/variants/nonear.so python - <<'PY' | tee -a gpurun_out/r06e_walkstats.txt
import json, importlib.util, os, sys
sys.path.insert(0, ".")
import forma_b200
spec = importlib.util.spec_from_file_location("g", "python-extras/sphere_scene_generator.py"); g = importlib.util.module_from_spec(spec); spec.loader.exec_module(g)
m = forma_b200.create_model(1920, 1080, seed=0)
m.scene.load_json(json.dumps(g.generate(100000))); m.scene.camera.use_json()
ws = m.walk_stats(16)
print("without the near bound: visits per ray %.2f, leaf tests per ray %.2f" % (ws.node_visits / ws.rays, ws.leaf_tests / ws.rays))
PY
