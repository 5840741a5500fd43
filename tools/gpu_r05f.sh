#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r05f_noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -4 gpurun_out/r05f_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r05f_pytest.log 2>&1; echo "all rc=$?"
tail -4 gpurun_out/r05f_pytest.log
