#!/bin/bash
# the contract's commands on the final code of the round: bench (N = 1), reference arm, launch list of the same bench command
cd "$GRAFT_REPO_ROOT"
tag=$1
python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench1.json 2> gpurun_out/${tag}_bench1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --no-other-configs --no-cpu-baseline > gpurun_out/${tag}_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-other-configs --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1
python tools/perf_one.py C > gpurun_out/${tag}_plain_C.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_ -s 1 -c 1 -o gpurun_out/${tag}_C -f python tools/perf_one.py C > gpurun_out/${tag}_ncu_C.log 2>&1
cat gpurun_out/${tag}_plain_C.log
