"""Data-parallel training step on >= 2 GPUs (skipped on a 1-GPU box): the overlapped, staged gradient all-reduce inside
``loss.backward()`` (train._GradAllReduce behind skg_head_backward_staged's stage events) leaves the rank mean of the
single-rank gradients in every ``p.grad``.  Runs tools/dist_train_check.py under torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_staged_allreduce_equals_rank_mean():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tools", "dist_train_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "\nOK" in out.stdout and "TRAIN_OK" in out.stdout, out.stdout[-2000:]
