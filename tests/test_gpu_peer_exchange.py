"""Peer-memory exchange (multiview.PeerExchange, gsr_backward_early_rgb_peers) -- needs two GPUs on the box; skipped
on a single-GPU box. Launches tests/mg_peer_check.py under torchrun (rendezvous on 127.0.0.1)."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_exchange_matches_plain_allreduce(world):
    """rtol 1e-4 (+ the plain path's own atomics spread) against one all_reduce per gradient tensor, at every world
    size the box has GPUs for: at 8 ranks every multicast address has 8 writers."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(29533 + world), str(ROOT / "tests" / "mg_peer_check.py")],
                       capture_output=True, text=True, timeout=900)
    assert "PEER_CHECK_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
