#!/bin/bash
cd "$GRAFT_REPO_ROOT"
V=$PWD/fo-rma_b200/variants
FORMA_RT_LIB=$V/t_top255.so timeout 900 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r05t_pytest_top255.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r05t_pytest_top255.log
rm -f gpurun_out/r05t_perf.txt
bash tools/gpu_variants.sh r05t default t_top63 t_top255 t_top511
