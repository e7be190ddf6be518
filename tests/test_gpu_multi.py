"""Multi-GPU checks of the symmetric-memory kernels (csrc/ws_collective.cu), run when the box has >= 2 GPUs (and
again at 4 ranks when it has >= 4): tests/multi_rank_checks.py under torchrun.  bench.py repeats the all-reduce
check in every N > 1 line (`exchange_check`), so the driver's scaling runs carry the evidence too."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_rank_checks.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "FAIL" not in r.stdout, r.stdout[-3000:]
    assert r.stdout.count("OK ") >= 10, r.stdout[-3000:]
    return r.stdout


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")
def test_collectives_on_two_gpus():
    out = _run(2, 29577)
    assert "OK sharded_adam" in out and "OK gather_frames" in out


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs four GPUs (multimem path)")
def test_collectives_on_four_gpus():
    out = _run(4, 29579)
    assert "OK allreduce[multimem]" in out
