"""Data-parallel gradient exchange captured inside the backward graph (SURVEY §8 a15 / e; reference
train_ddpm_cond_celebhq_multi_gpu.py:257-263,362), on two real GPUs over NCCL: runs tools/check_overlap.py under
torchrun — the bucketed in-graph all-reduce (flat gradient buffer in gradient-arrival order) must give bit-identical
gradients to one collective over the whole buffer, on every rank. Skips on a one-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_in_graph_bucketed_exchange_is_bit_identical_on_two_ranks():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29537", os.path.join(ROOT, "tools", "check_overlap.py"), "--steps", "3"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "gradients identical: True  ranks agree: True" in r.stdout, r.stdout[-2000:]
    assert "reduced_in_graph=True" in r.stdout
