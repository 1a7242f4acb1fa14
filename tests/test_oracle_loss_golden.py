"""CPU: the loss oracle (oracle/loss_oracle.c) against tests/golden/l*.npz -- outputs and autograd gradients of the
reference's own gs3dgs/utils/loss_utils.py run on CPU fp32 (tools/make_golden_loss.py). This pins the oracle."""
import glob
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
CASES = sorted(glob.glob(str(ROOT / "tests" / "golden" / "l?.npz")))

# The reference evaluates sigma = E[x^2] - mu^2 in fp32; on smooth images that cancellation leaves ~1e-5 relative
# noise in ITS map (case l4), so scalars are compared to 2e-5 and gradients to 1e-3 of the largest element.
SCALAR_ATOL = 2e-5
GRAD_RTOL_OF_MAX = 1e-3


def grad_close(a, b, n):
    return float(np.abs(a - b).max()) <= GRAD_RTOL_OF_MAX * float(np.abs(b).max()) + 1e-4 / n


def test_golden_files_present():
    assert len(CASES) == 5


@pytest.mark.parametrize("path", CASES, ids=lambda p: Path(p).stem)
def test_loss_oracle_matches_reference_module(path):
    from oracle import gs_oracle as O
    g = np.load(path)
    lam = float(g["lambda_dssim"])
    r = O.l1_ssim(g["img"], g["gt"], lam)
    n = g["img"].size
    for k in ("l1", "ssim", "loss"):
        assert abs(r[k] - float(g[k])) <= SCALAR_ATOL, (k, r[k], float(g[k]))
    for k in ("dl1", "dssim", "dloss"):
        assert grad_close(r[k], g[k], n), (k, float(np.abs(r[k] - g[k]).max()), float(np.abs(g[k]).max()))


def test_loss_oracle_properties():
    from oracle import gs_oracle as O
    rng = np.random.default_rng(3)
    a = rng.random((3, 30, 45), dtype=np.float32)
    r = O.l1_ssim(a, a.copy(), 0.2)
    assert r["l1"] == 0.0 and abs(r["ssim"] - 1.0) < 1e-12 and abs(r["loss"]) < 1e-12
    assert float(np.abs(r["dssim"]).max()) < 1e-12 and float(np.abs(r["dl1"]).max()) == 0.0
    # finite-difference check of the analytic gradient (double accumulators make this meaningful)
    b = np.clip(a + 0.2 * rng.standard_normal(a.shape).astype(np.float32), 0, 1)
    r = O.l1_ssim(a, b, 0.2)
    for (c, y, x) in ((0, 0, 0), (1, 15, 22), (2, 29, 44)):
        e = 1e-3
        ap = a.copy(); ap[c, y, x] += e
        am = a.copy(); am[c, y, x] -= e
        fd = (O.l1_ssim(ap, b, 0.2, grads=False)["ssim"] - O.l1_ssim(am, b, 0.2, grads=False)["ssim"]) / (2 * e)
        assert abs(fd - r["dssim"][c, y, x]) <= 2e-3 * abs(fd) + 1e-9, (fd, r["dssim"][c, y, x])
