#!/bin/bash
# round 2, after the machine loss: parity of the rewritten kernels, every config's time, scan-kernel A/B
cd "$GRAFT_REPO_ROOT"
(rustc --version; cargo --version) > gpurun_out/r05e_rust_probe.txt 2>&1
nproc >> gpurun_out/r05e_rust_probe.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/r05e_rust_probe.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r05e_pytest.log 2>&1
tail -5 gpurun_out/r05e_pytest.log
python tools/perf_configs.py --auto A B D > gpurun_out/r05e_perf.jsonl 2> gpurun_out/r05e_perf.err
for q in 0 1; do
  echo "== FORMA_RT_SCAN_QUEUE=$q" >> gpurun_out/r05e_scan_ab.txt
  FORMA_RT_SCAN_QUEUE=$q python tools/perf_configs.py builtin C 2>&1 | grep -v brute >> gpurun_out/r05e_scan_ab.txt
  FORMA_RT_SCAN_QUEUE=$q python tools/update_latency.py >> gpurun_out/r05e_scan_ab.txt 2>&1
done
echo "== heuristic (no override)" >> gpurun_out/r05e_scan_ab.txt
python tools/perf_configs.py builtin C 2>&1 | grep -v brute >> gpurun_out/r05e_scan_ab.txt
python tools/update_latency.py >> gpurun_out/r05e_scan_ab.txt 2>&1
cat gpurun_out/r05e_perf.jsonl gpurun_out/r05e_scan_ab.txt
