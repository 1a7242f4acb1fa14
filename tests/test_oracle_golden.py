"""CPU-only: pins the C oracle (oracle/gs_oracle.c) against golden vectors produced by the UNMODIFIED
reference CUDA build on a B200 (tools/make_golden.py -> tests/golden/*.npz)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import gs_oracle as O

GOLD = sorted((Path(__file__).resolve().parent / "golden").glob("g[0-9]*.npz"))


def _run(d):
    P, D, W, H, N = [int(x) for x in d["meta"]]
    pre = "colors_precomp" in d.files
    f = O.forward(d["bg"], d["means3D"], d["colors_precomp"] if pre else None, d["opacities"],
                  None if pre else d["scales"], None if pre else d["rotations"], 1.0,
                  d["cov3D_precomp"] if pre else None, d["viewmatrix"], d["projmatrix"],
                  float(d["tanfov"][0]), float(d["tanfov"][1]), H, W, None if pre else d["shs"], D, d["campos"])
    return f, (P, D, W, H, N)


@pytest.mark.parametrize("path", GOLD, ids=[p.stem for p in GOLD])
def test_forward_matches_reference_golden(path):
    d = np.load(path)
    f, (P, D, W, H, N) = _run(d)
    # integers: identical except fp-boundary cases (the CPU has no FMA contraction / CUDA expf)
    rad_mis = int((f.radii != d["radii"]).sum())
    assert rad_mis <= max(1, P // 1000), f"radii differ at {rad_mis} of {P}"
    same = f.radii == d["radii"]
    vis = (d["radii"] > 0) & same
    assert abs(f.num_rendered - N) <= 8 * max(1, rad_mis)
    np.testing.assert_allclose(f.inter["means2D"][vis], d["i_means2D"][vis], rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(f.inter["depths"][vis], d["i_depths"][vis], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(f.inter["conic_opacity"][vis], d["i_conic_opacity"][vis], rtol=2e-4, atol=1e-6)
    if "colors_precomp" not in d.files:
        np.testing.assert_allclose(f.inter["rgb"][vis], d["i_rgb"][vis], rtol=1e-5, atol=1e-6)
        np.testing.assert_array_equal(f.inter["clamped"][vis], d["i_clamped"][vis])
        np.testing.assert_allclose(f.inter["cov3D"][vis], d["i_cov3D"][vis], rtol=1e-5,
                                   atol=1e-6 * float(np.abs(d["i_cov3D"]).max()))
    assert (f.inter["tiles_touched"][vis] != d["i_tiles_touched"][vis].astype(np.uint32)).sum() == 0
    if rad_mis == 0 and f.num_rendered == N:
        keys_ref = d["i_point_list_keys"].astype(np.uint64)
        # tile ids identical everywhere; depth bits may differ by an ulp, the ORDER must be the reference's
        assert ((f.inter["point_list_keys"] >> np.uint64(32)) == (keys_ref >> np.uint64(32))).all()
        mism = (f.inter["point_list"] != d["i_point_list"].astype(np.uint32)).mean()
        assert mism < 5e-3, f"sorted list differs at {mism:.2%}"
        np.testing.assert_array_equal(f.inter["ranges"], d["i_ranges"].astype(np.uint32))
        nc = (f.inter["n_contrib"] != d["i_n_contrib"].astype(np.uint32)).mean()
        assert nc < 5e-3
    for k in ("color", "depth", "alpha"):
        np.testing.assert_allclose(getattr(f, k), d[k], rtol=1e-4, atol=2e-5, err_msg=k)


@pytest.mark.parametrize("path", GOLD, ids=[p.stem for p in GOLD])
def test_backward_matches_reference_golden(path):
    d = np.load(path)
    f, (P, D, W, H, N) = _run(d)
    g = O.backward(f, d["dL_dcolor"], d["dL_ddepth"], d["dL_dalpha"])
    same = (f.radii == d["radii"])
    pre = "colors_precomp" in d.files
    names = ["means2D", "opacity", "colors", "means3D", "cov3D"] + ([] if pre else ["sh", "scales", "rotations"])
    for k in names:
        ref = d["g_" + k].reshape(P, -1)[same]
        spread = float(d["s_" + k].max())
        got = g[k].reshape(P, -1)[same].astype(np.float64)
        atol = 1e-5 * np.abs(ref).max() + 2 * spread
        bad = np.abs(got - ref) > atol + 2e-3 * np.abs(ref)
        assert bad.mean() < 2e-3, f"{k}: {bad.sum()} of {bad.size} beyond tolerance"


def test_oracle_linearity_in_upstream_gradient():
    """Backward is linear in the upstream gradients: g(a*x + b*y) = a*g(x) + b*g(y)."""
    d = np.load(GOLD[0])
    f, (P, D, W, H, N) = _run(d)
    rng = np.random.default_rng(0)
    x = [rng.standard_normal(d[k].shape).astype(np.float32) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    y = [rng.standard_normal(d[k].shape).astype(np.float32) for k in ("dL_dcolor", "dL_ddepth", "dL_dalpha")]
    gx, gy = O.backward(f, *x), O.backward(f, *y)
    gz = O.backward(f, *[2.0 * a - 0.5 * b for a, b in zip(x, y)])
    for k in ("means3D", "sh", "opacity", "scales", "rotations"):
        ref = 2.0 * gx[k].astype(np.float64) - 0.5 * gy[k]
        np.testing.assert_allclose(gz[k], ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max())


def test_oracle_edge_cases():
    d = np.load(GOLD[0])
    P, D, W, H, N = [int(x) for x in d["meta"]]
    # empty input
    f = O.forward(d["bg"], np.zeros((0, 3), np.float32), None, np.zeros((0, 1), np.float32), np.zeros((0, 3), np.float32),
                  np.zeros((0, 4), np.float32), 1.0, None, d["viewmatrix"], d["projmatrix"], 0.6, 0.4, 32, 48,
                  np.zeros((0, 1, 3), np.float32), 0, d["campos"])
    assert f.num_rendered == 0 and float(np.abs(f.color).max()) == 0.0     # rasterize_points.cu:83
    # everything behind the camera: culled, background only
    m = d["means3D"].copy(); m[:, 2] = -3.0
    f = O.forward(d["bg"], m, None, d["opacities"], d["scales"], d["rotations"], 1.0, None, d["viewmatrix"],
                  d["projmatrix"], float(d["tanfov"][0]), float(d["tanfov"][1]), H, W, d["shs"], D, d["campos"])
    assert f.num_rendered == 0 and (f.radii == 0).all() and float(np.abs(f.alpha).max()) == 0.0
    vis = O.mark_visible(d["means3D"], d["viewmatrix"], d["projmatrix"])
    assert vis.dtype == bool and ((d["radii"] > 0) <= vis).all()
