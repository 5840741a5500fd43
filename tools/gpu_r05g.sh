#!/bin/bash
# ncu captures: the queue scan kernel on the reference's built-in scene, the tree kernel on D and A
cd "$GRAFT_REPO_ROOT"
for c in S0 S0r D A; do
  python tools/perf_one.py $c > gpurun_out/r05g_plain_$c.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_ -s 1 -c 1 -o gpurun_out/r05g_$c -f python tools/perf_one.py $c > gpurun_out/r05g_ncu_$c.log 2>&1
  cat gpurun_out/r05g_plain_$c.log
done
