#!/bin/bash
cd "$GRAFT_REPO_ROOT"
for v in default bs1 bs4 bs8 smb6 smb8; do
  echo "== $v" >> gpurun_out/r06g_knobs.txt
  if [ "$v" = default ]; then unset FORMA_RT_LIB; else export FORMA_RT_LIB=$PWD/fo-rma_b200/variants/$v.so; fi
  python tools/perf_configs.py --auto C A D builtin 2>/dev/null | grep -v "A' \|brute" >> gpurun_out/r06g_knobs.txt
done
cat gpurun_out/r06g_knobs.txt
