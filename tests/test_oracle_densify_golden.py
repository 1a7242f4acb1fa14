"""The numpy restatement of densify_and_prune (oracle/optim_oracle.py) against tests/golden/densify.npz, which holds
the reference's own GaussianModel.densify_and_prune (gs3dgs and gs2dgs) run on CPU (tools/make_golden_densify.py)."""
import os

import numpy as np
import pytest

from oracle import optim_oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify.npz"))


@pytest.mark.parametrize("c", [str(x) for x in G["cases"]])
def test_oracle_matches_reference_model(c):
    params = {n: G[f"{c}_in_{n}"] for n in O.GROUPS}
    moments = {n: (G[f"{c}_in_m_{n}"], G[f"{c}_in_v_{n}"]) for n in O.GROUPS if f"{c}_in_m_{n}" in G}
    max_grad, min_opacity, extent, mss, percent_dense = (float(x) for x in G[f"{c}_cfg"])
    par, mom = O.densify_and_prune(params, moments, G[f"{c}_in_accum"], G[f"{c}_in_denom"], max_grad, min_opacity, extent,
                                   mss or None, percent_dense, G[f"{c}_normals"])
    rows = int(G[f"{c}_out_stats_rows"][0])
    P = params["xyz"].shape[0]
    assert rows != P and rows > P // 2          # the case densifies and prunes for real
    for n in O.GROUPS:
        ref = G[f"{c}_out_{n}"]
        assert par[n].shape == ref.shape, (n, par[n].shape, ref.shape)
        if n in ("xyz", "scaling"):             # recomputed for split children: numpy exp/log vs torch's, bmm order
            np.testing.assert_allclose(par[n], ref, rtol=2e-6, atol=2e-6)
            assert (par[n] == ref).mean() > 0.5          # copied rows are exact
        else:
            assert np.array_equal(par[n], ref), n
        if f"{c}_out_m_{n}" in G:
            assert np.array_equal(mom[n][0], G[f"{c}_out_m_{n}"]) and np.array_equal(mom[n][1], G[f"{c}_out_v_{n}"]), n
        else:
            assert n not in mom


def test_reset_opacity_oracle_matches_reference_model():
    """GaussianModel.reset_opacity run on CPU (tools/make_golden_densify.py:reset_case)"""
    got = O.reset_opacity(G["reset_in_opacity"])
    ref = G["reset_out_opacity"]
    np.testing.assert_allclose(got, ref, rtol=2e-6, atol=2e-6)        # numpy exp/log vs torch's CPU kernels
    assert float(ref.max()) <= np.log(0.01 / 0.99) + 1e-5 and (ref < ref.max() - 1e-3).any()   # capped and not all capped
