"""GPU test of the multi-GPU path: launches tests/multigpu_check.py under torchrun on every GPU of
the box (2..8).  Skipped on a single-GPU box; the host-side logic is covered on CPU by
tests/test_exchange.py / tests/test_sharding.py (gloo, world size 2)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_sharded_step_matches_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(n, 8)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(ROOT, "tests", "multigpu_check.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0 and "multigpu_check ok" in r.stdout, r.stdout[-3000:]
