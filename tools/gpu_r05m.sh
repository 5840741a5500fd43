#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r05m_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r05m_pytest.log
echo "== default (phased, 2 passes)" > gpurun_out/r05m_perf.txt
python tools/perf_configs.py builtin >> gpurun_out/r05m_perf.txt 2>&1
for v in q_simple q_p1 q_p3 q_p2b5; do
  echo "== $v" >> gpurun_out/r05m_perf.txt
  FORMA_RT_LIB=$PWD/fo-rma_b200/variants/$v.so python tools/perf_configs.py builtin >> gpurun_out/r05m_perf.txt 2>&1
done
cat gpurun_out/r05m_perf.txt
