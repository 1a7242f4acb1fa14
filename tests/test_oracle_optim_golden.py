"""CPU: the optimizer oracle (oracle/optim_oracle.py) against tests/golden/adam.npz -- torch.optim.Adam configured as
GaussianModel.training_setup does and the reference's densification-statistics expressions, run on CPU fp32
(tools/make_golden_optim.py)."""
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def test_adam_oracle_matches_torch_adam():
    from oracle import optim_oracle as O
    g = np.load(ROOT / "tests" / "golden" / "adam.npz")
    lrs = list(g["lrs"])
    worst = 0.0
    for gi, n in enumerate(NAMES):
        p = g[f"p0_{n}"]; m = np.zeros_like(p); v = np.zeros_like(p)
        for t in range(1, int(g["steps"]) + 1):
            lr = 0.00012 if (n == "xyz" and t >= 3) else lrs[gi]
            p, m, v = O.adam_step(p, g[f"g{t}_{n}"], m, v, lr, t)
            for name, ours, ref in (("p", p, g[f"p{t}_{n}"]), ("m", m, g[f"m{t}_{n}"]), ("v", v, g[f"v{t}_{n}"])):
                # torch's CPU kernels may or may not contract to fma: agree to 2 ulp, not necessarily bit for bit
                # (relative to the element, plus a few ulp of the tensor's largest element: a parameter near zero
                # inherits the rounding of its much larger update term)
                tol = 3e-7 * np.abs(ref) + 3e-7 * float(np.abs(ref).max())
                err = np.abs(ours.astype(np.float64) - ref) / tol
                worst = max(worst, float(err.max()))
                assert float(err.max()) <= 1.0, (n, t, name, float(err.max()))
    assert worst <= 1.0


def test_densification_stats_oracle_matches_reference_expressions():
    from oracle import optim_oracle as O
    g = np.load(ROOT / "tests" / "golden" / "adam.npz")
    a, d, m = O.densification_stats(g["ds_accum0"][:, 0], g["ds_denom0"][:, 0], g["ds_maxr0"], g["ds_vgrad"], g["ds_radii"])
    np.testing.assert_allclose(a, g["ds_accum1"][:, 0], rtol=2e-7, atol=0)
    assert np.array_equal(d, g["ds_denom1"][:, 0]) and np.array_equal(m, g["ds_maxr1"])
    inv = g["ds_radii"] <= 0
    assert np.array_equal(a[inv], g["ds_accum0"][inv, 0])
