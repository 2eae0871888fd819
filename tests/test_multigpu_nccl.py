"""Multi-GPU parity on real GPUs (skipped with fewer than 2): NCCL row-partitioned build == single GPU."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_row_partitioned_nccl_equals_single_gpu():
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if ngpu < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "nccl_rowpart_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=560)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert "NCCL_ROWPART_OK" in res.stdout
