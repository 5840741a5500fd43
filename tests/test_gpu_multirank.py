"""-m gpu: the multi-GPU exchange on REAL ranks - one process per GPU under torchrun, CUDA-IPC peer memory, flags.

Needs at least two GPUs (ranks that wait on each other's flags must not share a GPU); on a one-GPU box the test is
skipped and the CPU `gloo` test (tests/test_multirank_gloo.py) covers the host-side arithmetic.  What runs is
tools/exchange_check.py: 24 frames of uneven, skewed work per split;
  * sample-split: rank 0's frame after the flag-ordered fused reduce + tone-map must equal ONE GPU's composition of
    the same per-rank sample ranges summed in rank order, byte for byte;
  * tile-split: rank 0's frame must equal one GPU's render of the whole frame, byte for byte (same sample order)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_exchange_on_all_gpus_of_the_box():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU: ranks that spin on each other's flags need a GPU each")
    n = min(n, 8)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                          "--master-addr", "127.0.0.1", "--master-port", "29631",
                          os.path.join(ROOT, "tools", "exchange_check.py"), "24"],
                         capture_output=True, text=True, env=env, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert f"exchange_check: 24 frames on {n} GPUs, 0 bad" in out.stdout
    assert f"exchange_check (tiles): 24 frames on {n} GPUs, 0 bad" in out.stdout
