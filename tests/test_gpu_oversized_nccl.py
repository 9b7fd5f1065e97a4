"""The NCCL leg of the oversized pile (`tools.compute_obs_obs_similarity_top` over the ranks of
`torch.distributed`): two ranks on two GPUs, launched from here with torchrun, must gather the very matrix one GPU computes
alone.  Skipped on a box with one GPU (the one-GPU split is covered by tests/test_gpu_oversized.py)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_ranks_gather_the_unsharded_result():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 29600 + os.getpid() % 300
    command = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.join(ROOT, "scripts", "run_oversized_pile.py"), "--cells", "9000", "--genes", "600",
               "--top", "70"]
    result = subprocess.run(command, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert result.returncode == 0, result.stdout[-2000:] + result.stderr[-2000:]
    assert "world 2:" in result.stdout and "identical: True" in result.stdout, result.stdout[-2000:]
