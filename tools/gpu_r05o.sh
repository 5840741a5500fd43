#!/bin/bash
cd "$GRAFT_REPO_ROOT"
for v in default q_b7 q_b8; do
  echo "== $v" >> gpurun_out/r05o_queue_occ.txt
  if [ "$v" = default ]; then unset FORMA_RT_LIB; else export FORMA_RT_LIB=$PWD/fo-rma_b200/variants/$v.so; fi
  python tools/perf_configs.py builtin >> gpurun_out/r05o_queue_occ.txt 2>&1
done
unset FORMA_RT_LIB
cat gpurun_out/r05o_queue_occ.txt
python bench.py --steps 2 --warmup 3 --no-other-configs --no-cpu-baseline > gpurun_out/r05o_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r05o_launches.csv python bench.py --steps 2 --warmup 3 --no-other-configs --no-cpu-baseline > gpurun_out/r05o_ncu_launches.log 2>&1
for c in C D; do
  python tools/perf_one.py $c > gpurun_out/r05o_plain_$c.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_ -s 1 -c 1 -o gpurun_out/r05o_$c -f python tools/perf_one.py $c > gpurun_out/r05o_ncu_$c.log 2>&1
  cat gpurun_out/r05o_plain_$c.log
done
