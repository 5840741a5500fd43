#!/bin/bash
cd "$GRAFT_REPO_ROOT"
python bench.py --steps 20 --warmup 3 > gpurun_out/r05i_bench1.json 2> gpurun_out/r05i_bench1.err; echo "bench rc=$?"
tail -3 gpurun_out/r05i_bench1.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r05i_ref.json 2> gpurun_out/r05i_ref.err; echo "ref rc=$?"
python tools/perf_one.py C > gpurun_out/r05i_plain_C.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_ -s 1 -c 1 -o gpurun_out/r05i_C -f python tools/perf_one.py C > gpurun_out/r05i_ncu_C.log 2>&1
cat gpurun_out/r05i_plain_C.log
