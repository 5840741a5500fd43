#!/bin/bash
# usage: tools/gpu_multi.sh <tag> <N>: the multi-rank parity test and bench.py on N GPUs of one box
cd "$GRAFT_REPO_ROOT"
tag=$1; N=$2
nvidia-smi -L | head -8
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu > gpurun_out/${tag}_multirank_pytest_$N.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${tag}_multirank_pytest_$N.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${tag}_bench$N.json 2> gpurun_out/${tag}_bench$N.err; echo "bench rc=$?"
tail -5 gpurun_out/${tag}_bench$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/${tag}_ref$N.json 2>> gpurun_out/${tag}_bench$N.err; echo "ref rc=$?"
