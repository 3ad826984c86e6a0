"""Tensor-parallel parity on real GPUs (needs >= 2): the NVLink peer-memory exchange kernel (b200_allreduce_rmsnorm) against
NCCL + b200_rmsnorm bit for bit, and the decoder at TP=2 against TP=1.  Runs tools/tp_check.py under torchrun."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tp2_exchange_and_model_parity():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "tp_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=420)
    assert r.returncode == 0 and "tp_check ok" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
