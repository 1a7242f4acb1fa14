"""CPU-only: pins the surfel and knn C oracles (oracle/surfel_oracle.c, oracle/knn_oracle.c) against golden
vectors produced by the UNMODIFIED reference CUDA build on a B200 (tools/make_golden.py ->
tests/golden/s*.npz, k*.npz)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import gs_oracle as O

GDIR = Path(__file__).resolve().parent / "golden"
SGOLD = sorted(GDIR.glob("s[0-9]*.npz"))
KGOLD = sorted(GDIR.glob("k[0-9]*.npz"))


def test_golden_families_present():
    assert len(SGOLD) >= 4 and len(KGOLD) >= 4, "surfel / knn golden vectors missing (tools/make_golden.py)"


def _run(d):
    P, D, W, H, N = [int(x) for x in d["meta"]]
    pre = "colors_precomp" in d.files
    f = O.surfel_forward(d["bg"], d["means3D"], d["colors_precomp"] if pre else None, d["opacities"],
                         None if pre else d["scales"], None if pre else d["rotations"], 1.0,
                         d["transMat_precomp"] if pre else None, d["viewmatrix"], d["projmatrix"],
                         float(d["tanfov"][0]), float(d["tanfov"][1]), H, W, None if pre else d["shs"], D, d["campos"])
    return f, (P, D, W, H, N), pre


@pytest.mark.parametrize("path", SGOLD, ids=[p.stem for p in SGOLD])
def test_surfel_forward_matches_reference_golden(path):
    d = np.load(path)
    f, (P, D, W, H, N), pre = _run(d)
    rad_mis = int((f.radii != d["radii"]).sum())
    assert rad_mis <= max(1, P // 1000), f"radii differ at {rad_mis} of {P}"
    same = f.radii == d["radii"]
    vis = (d["radii"] > 0) & same
    assert abs(f.num_rendered - N) <= 8 * max(1, rad_mis)
    np.testing.assert_allclose(f.inter["depths"][vis], d["i_depths"][vis], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(f.inter["means2D"][vis], d["i_means2D"][vis], rtol=1e-4, atol=2e-3)
    np.testing.assert_allclose(f.inter["normal_opacity"][vis], d["i_normal_opacity"][vis], rtol=1e-5, atol=1e-6)
    if not pre:
        tscale = float(np.abs(d["i_transMat"][vis]).max())
        np.testing.assert_allclose(f.inter["transMat"][vis], d["i_transMat"][vis], rtol=1e-5, atol=1e-6 * tscale)
        np.testing.assert_allclose(f.inter["rgb"][vis], d["i_rgb"][vis], rtol=1e-5, atol=1e-6)
        np.testing.assert_array_equal(f.inter["clamped"][vis], d["i_clamped"][vis])
    assert (f.inter["tiles_touched"][vis] != d["i_tiles_touched"][vis].astype(np.uint32)).sum() == 0
    if rad_mis == 0 and f.num_rendered == N:
        keys_ref = d["i_point_list_keys"].astype(np.uint64)
        assert ((f.inter["point_list_keys"] >> np.uint64(32)) == (keys_ref >> np.uint64(32))).all()
        assert (f.inter["point_list"] != d["i_point_list"].astype(np.uint32)).mean() < 5e-3
        np.testing.assert_array_equal(f.inter["ranges"], d["i_ranges"].astype(np.uint32))
        nc = d["i_n_contrib"].astype(np.uint32)
        assert (f.inter["n_contrib"] != nc[0]).mean() < 5e-3
        assert (f.inter["median_contrib"] != nc[1]).mean() < 5e-3
        acc = d["i_accum"]
        for k, ref in (("final_T", acc[0]), ("M1", acc[1]), ("M2", acc[2])):
            bad = np.abs(f.inter[k] - ref) > 2e-5 + 1e-4 * np.abs(ref)
            assert bad.mean() < 2e-3, k
    bad = np.abs(f.color - d["color"]) > 2e-5 + 1e-4 * np.abs(d["color"])
    assert bad.mean() < 1e-3, "color"
    # channel-wise scale: depth/median depth are O(10), distortion O(1), alpha/normal O(1)
    for ch in range(7):
        ref = d["others"][ch]
        bad = np.abs(f.others[ch] - ref) > 1e-5 * max(1.0, float(np.abs(ref).max())) + 1e-4 * np.abs(ref)
        assert bad.mean() < 2e-3, f"others[{ch}]"      # a median-depth flip at T == 0.5 boundaries is a whole-depth jump


@pytest.mark.parametrize("path", SGOLD, ids=[p.stem for p in SGOLD])
def test_surfel_backward_matches_reference_golden(path):
    d = np.load(path)
    f, (P, D, W, H, N), pre = _run(d)
    g = O.surfel_backward(f, d["dL_dcolor"], d["dL_dothers"])
    same = (f.radii == d["radii"])
    names = ["means2D", "opacity", "colors", "means3D", "transMat"] + ([] if pre else ["sh", "scales", "rotations"])
    for k in names:
        ref = d["g_" + k].reshape(P, -1)[same]
        spread = float(d["s_" + k].max())
        got = g[k].reshape(P, -1)[same].astype(np.float64)
        atol = 1e-5 * np.abs(ref).max() + 2 * spread
        bad = np.abs(got - ref) > atol + 2e-3 * np.abs(ref)
        assert bad.mean() < 3e-3, f"{k}: {bad.sum()} of {bad.size} beyond tolerance"


def test_surfel_oracle_linearity_and_edges():
    d = np.load(SGOLD[0])
    f, (P, D, W, H, N), pre = _run(d)
    rng = np.random.default_rng(0)
    x = [rng.standard_normal(d[k].shape).astype(np.float32) for k in ("dL_dcolor", "dL_dothers")]
    y = [rng.standard_normal(d[k].shape).astype(np.float32) for k in ("dL_dcolor", "dL_dothers")]
    gx, gy = O.surfel_backward(f, *x), O.surfel_backward(f, *y)
    gz = O.surfel_backward(f, *[2.0 * a - 0.5 * b for a, b in zip(x, y)])
    for k in ("means3D", "sh", "opacity", "scales", "rotations", "transMat"):
        ref = 2.0 * gx[k].astype(np.float64) - 0.5 * gy[k]
        np.testing.assert_allclose(gz[k], ref, rtol=1e-3, atol=1e-4 * np.abs(ref).max())
    # alpha channel is 1 - final_T; the median contributor never comes after the last one
    np.testing.assert_array_equal(f.others[1], 1 - f.inter["final_T"])
    assert (f.inter["median_contrib"] <= f.inter["n_contrib"]).all()
    # everything behind the camera: background only, zero maps
    m = d["means3D"].copy(); m[:, 2] = -3.0
    f2 = O.surfel_forward(d["bg"], m, None, d["opacities"], d["scales"], d["rotations"], 1.0, None, d["viewmatrix"],
                          d["projmatrix"], float(d["tanfov"][0]), float(d["tanfov"][1]), H, W, d["shs"], D, d["campos"])
    assert f2.num_rendered == 0 and (f2.radii == 0).all() and float(np.abs(f2.others).max()) == 0.0
    f3 = O.surfel_forward(d["bg"], np.zeros((0, 3), np.float32), None, np.zeros((0, 1), np.float32),
                          np.zeros((0, 2), np.float32), np.zeros((0, 4), np.float32), 1.0, None, d["viewmatrix"],
                          d["projmatrix"], 0.6, 0.4, 32, 48, np.zeros((0, 1, 3), np.float32), 0, d["campos"])
    assert f3.num_rendered == 0 and float(np.abs(f3.color).max()) == 0.0


@pytest.mark.parametrize("path", KGOLD, ids=[p.stem for p in KGOLD])
def test_knn_oracle_matches_reference_golden(path):
    d = np.load(path)
    pts, ref = d["points"], d["dist2"]
    got = O.knn_dist2(pts)
    brute = O.knn_dist2(pts, brute=True)
    np.testing.assert_array_equal(got, brute)            # the reference search is exact
    # CPU (no FMA contraction) vs the CUDA build: a few ulp on each squared distance
    fin = np.isfinite(ref)
    np.testing.assert_array_equal(np.isfinite(got), fin)
    np.testing.assert_allclose(got[fin], ref[fin], rtol=2e-6, atol=1e-12)


def test_knn_oracle_permutation_invariance():
    rng = np.random.default_rng(3)
    pts = rng.random((3000, 3)).astype(np.float32)
    perm = rng.permutation(3000)
    np.testing.assert_array_equal(O.knn_dist2(pts)[perm], O.knn_dist2(pts[perm]))
