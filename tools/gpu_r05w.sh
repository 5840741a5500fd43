#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r05w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r05w_pytest.log
python tools/perf_configs.py C builtin 2>&1 | grep -v brute > gpurun_out/r05w_perf.jsonl; cat gpurun_out/r05w_perf.jsonl
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r05w_bench2.json 2> gpurun_out/r05w_bench2.err; echo "bench rc=$?"
