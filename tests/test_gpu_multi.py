"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): sharded rollout + advantage over NCCL equals the
single-process reference goldens.  Launches tests/dist_rollout_check.py under torchrun, one rank per GPU."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_rollout_and_advantage_nccl():
    n = min(torch.cuda.device_count(), 4)
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29671", os.path.join(here, "dist_rollout_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("dist check ok") == 3
