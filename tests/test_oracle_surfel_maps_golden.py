"""The numpy restatement of the 2DGS caller-side maps (oracle/surfel_maps_oracle.py; the row DESIGN.md section 10 names
next, no product kernel yet) against tests/golden/surfel_maps.npz: the reference's own depth_to_normal /
depths_to_points run on CPU, and the torch replay of render()'s inline statements and the two regularisers."""
import os

import numpy as np
import pytest

from oracle import surfel_maps_oracle as SM

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "surfel_maps.npz"))


def _close(a, b, what):
    # small float32 matrix inverses differ between numpy (LAPACK) and torch in the last bits
    scale = float(np.abs(b).max()) + 1e-30
    assert a.shape == b.shape, what
    assert float(np.abs(a - b).max()) <= 2e-5 * scale, (what, float(np.abs(a - b).max()), scale)


@pytest.mark.parametrize("c", [str(x) for x in G["cases"]])
def test_maps_match_reference(c):
    W, H = (int(v) for v in G[f"{c}_WH"])
    wvt, fpt, allmap = G[f"{c}_wvt"], G[f"{c}_fpt"], G[f"{c}_allmap"]
    ratio = float(G[f"{c}_depth_ratio"])
    m = SM.surfel_maps(allmap, wvt, fpt, W, H, ratio)
    assert np.array_equal(m["render_alpha"], allmap[1:2]) and np.array_equal(m["render_dist"], allmap[6:7])
    _close(m["render_normal"], G[f"{c}_render_normal"], "render_normal")
    assert np.isfinite(m["render_depth"]).all()                                   # nan_to_num removed 0/0 and the NaN medians
    _close(m["render_depth"], G[f"{c}_render_depth"], "surf_depth")
    _close(SM.depths_to_points(wvt, fpt, W, H, G[f"{c}_render_depth"]), G[f"{c}_points"], "points")
    dn = SM.depth_to_normal(wvt, fpt, W, H, G[f"{c}_render_depth"])
    ref = G[f"{c}_depth_to_normal"]
    assert np.all(dn[0] == 0) and np.all(dn[:, 0] == 0) and np.all(dn[-1] == 0) and np.all(dn[:, -1] == 0)
    # unit normals from float32 finite differences of points ~6 units away: compare directions
    cos = (dn * ref).sum(-1)[1:-1, 1:-1]
    assert float(cos.min()) > 1 - 1e-3 and float(np.median(cos)) > 1 - 1e-6, (float(cos.min()), float(np.median(cos)))
    ln, ld = SM.regularisers(G[f"{c}_render_normal"], G[f"{c}_surf_normal"], allmap[6:7], *G[f"{c}_lambdas"])
    assert abs(ln - G[f"{c}_losses"][0]) <= 1e-6 * abs(G[f"{c}_losses"][0]) + 1e-9
    assert abs(ld - G[f"{c}_losses"][1]) <= 1e-6 * abs(G[f"{c}_losses"][1]) + 1e-9
