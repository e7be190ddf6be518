"""Two-GPU checks of the symmetric-memory kernels (csrc/ws_collective.cu), run when the box has >= 2 GPUs:
the fused exchange + Adam step must leave bit-identical parameters to all-reduce + fused Adam on every rank."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")
def test_sharded_adam_matches_allreduce_plus_adam_on_two_gpus():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29577", os.path.join(ROOT, "tools", "sharded_adam_check.py"), "200000"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "bit-identical on all ranks: True" in r.stdout
