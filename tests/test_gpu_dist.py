"""Two-GPU NCCL test (skipped with fewer than 2 GPUs): the row-sharded training step's all-reduced gradients equal the single-GPU
gradients on the concatenated batch, eagerly and through the captured graph with the all-reduce inside (scripts/dist_grad_check.py).
The host-side sharding logic is covered on CPU with gloo in tests/test_dist_cpu.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_train_step_equals_single_gpu_step():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "scripts", "dist_grad_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "OK" in r.stdout
