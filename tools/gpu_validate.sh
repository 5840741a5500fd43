#!/bin/bash
# usage: tools/gpu_validate.sh <tag>: the round-end sequence on one GPU - pytest -m gpu, smoke, bench, reference arm
cd "$GRAFT_REPO_ROOT"
tag=$1
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${tag}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench1.json 2> gpurun_out/${tag}_bench1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
