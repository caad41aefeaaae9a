"""Runs tests/multigpu_check.py under torchrun on every GPU of the box (2 ranks, and all of them up to 8; skipped on a
single-GPU box): fused == split == staged statistics exchange == NCCL path == one process, uneven shards, the global
mixup bank == the unsharded run, lookahead order == ordinary order, NormalizeBatch(sync), a late rank."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_GPUS = torch.cuda.device_count() if torch.cuda.is_available() else 0
RANKS = sorted({2, min(N_GPUS, 8)}) if N_GPUS >= 2 else [2]


@pytest.mark.skipif(N_GPUS < 2, reason="needs two GPUs on one node")
@pytest.mark.parametrize("ranks", RANKS)
def test_multigpu_check(ranks):
    port = 29600 + (os.getpid() + 7 * ranks) % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(ranks), "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multigpu_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert f"multigpu_check: OK on {ranks} GPUs" in res.stdout
