#!/bin/bash
# usage: tools/gpu_variants.sh <tag> <variant> [<variant> ...]   (variants = fo-rma_b200/variants/<name>.so; "default" = the built library)
# the tree-path configs per variant library
cd "$GRAFT_REPO_ROOT"
tag=$1; shift
for v in "$@"; do
  echo "== $v" >> gpurun_out/${tag}_perf.txt
  if [ "$v" = default ]; then unset FORMA_RT_LIB; else export FORMA_RT_LIB=$PWD/fo-rma_b200/variants/$v.so; fi
  python tools/perf_configs.py --auto A B D 2>> gpurun_out/${tag}_perf.err | grep -v "A' " >> gpurun_out/${tag}_perf.txt
done
cat gpurun_out/${tag}_perf.txt
