"""Peer-memory exchange (multiview.PeerExchange, gsr_backward_early_rgb_peers) -- needs two GPUs on the box; skipped
on a single-GPU box. Launches tests/mg_peer_check.py under torchrun (rendezvous on 127.0.0.1)."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_peer_exchange_matches_nccl_paths_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(ROOT / "tests" / "mg_peer_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert "PEER_CHECK_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
