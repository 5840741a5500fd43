#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r05l_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r05l_pytest.log
python tools/perf_configs.py --auto C A B D builtin > gpurun_out/r05l_perf.jsonl 2> gpurun_out/r05l_perf.err
cat gpurun_out/r05l_perf.jsonl
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r05l_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r05l_smoke.log
