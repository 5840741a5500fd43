#!/bin/bash
cd "$GRAFT_REPO_ROOT"
V=$PWD/fo-rma_b200/variants
FORMA_RT_LIB=$V/t_pl.so timeout 900 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r05r_pytest_t_pl.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r05r_pytest_t_pl.log
rm -f gpurun_out/r05r_perf.txt
bash tools/gpu_variants.sh r05r default t_pl > /dev/null
for lb in 4 12 16; do for rf in 16 20 24; do
  echo "== t_pl leaf_batch $lb refill $rf" >> gpurun_out/r05r_perf.txt
  FORMA_RT_LIB=$V/t_pl.so FORMA_RT_LEAF_BATCH=$lb FORMA_RT_REFILL=$rf python tools/perf_configs.py --auto A D 2>/dev/null | grep -v "A' " >> gpurun_out/r05r_perf.txt
done; done
cat gpurun_out/r05r_perf.txt
