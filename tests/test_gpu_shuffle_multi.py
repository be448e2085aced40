"""GPU, 2 ranks: the in-box shuffle (RepartitionExec with a communicator) over NVLink peer memory and over the NCCL
send/recv fallback.  Needs two GPUs on the box; skipped otherwise.  The worker (shuffle_check_worker.py) asserts
that the multiset of full rows is preserved bit-exactly, equal keys land on one rank, and the (source rank, source
row) order inside each output partition is kept."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_two_rank_shuffle_paths():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29531", os.path.join(HERE, "shuffle_check_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "shuffle_check: all ok" in p.stdout


def test_two_rank_q3_with_shuffle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29532", os.path.join(HERE, "q3_dist_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "q3_dist: all ok" in p.stdout
