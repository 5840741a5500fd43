#!/bin/bash
cd "$GRAFT_REPO_ROOT"
for d in 50 10 3 0; do python tools/perf_update.py 6 $d; done > gpurun_out/r05s_update_depth.txt 2>&1
for q in 0 1; do FORMA_RT_SCAN_QUEUE=$q python tools/perf_update.py 6 50; done >> gpurun_out/r05s_update_depth.txt 2>&1
cat gpurun_out/r05s_update_depth.txt
python tools/perf_update.py 6 50 > /dev/null 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_ -s 3 -c 1 -o gpurun_out/r05s_update -f python tools/perf_update.py 6 50 > gpurun_out/r05s_ncu.log 2>&1
python - <<'PY'
import json, time, sys
sys.path.insert(0, ".")
import forma_b200
txt = open("/dev/stdin").read() if False else None
PY
python -c "
import json, time, importlib.util, sys
sys.path.insert(0, '.')
import forma_b200
spec = importlib.util.spec_from_file_location('g', 'python-extras/sphere_scene_generator.py'); g = importlib.util.module_from_spec(spec); spec.loader.exec_module(g)
txt = json.dumps(g.generate(100000)).encode()
m = forma_b200.create_model(1920, 1080, seed=0)
for i in range(3):
    t0 = time.perf_counter(); m.scene.load_json(txt); m.scene.camera.use_json(); t1 = time.perf_counter(); m.render(1); t2 = time.perf_counter()
    print('D scene: parse+mapping %.1f ms, first frame (pack, H2D, LBVH build, 1 spp) %.1f ms' % ((t1-t0)*1e3, (t2-t1)*1e3))
" > gpurun_out/r05s_sceneload.txt 2>&1; cat gpurun_out/r05s_sceneload.txt
