"""GPU parity of the 3DGS path: product (libgsr_b200.so through its C ABI) vs the reference CUDA
build on the same GPU (oracle/_ref) -- integers and forward images bit-exact, gradients within
rtol 1e-4 (atol 1e-6*max|ref|, widened by the reference's own run-to-run atomics noise) -- and vs
the CPU oracle within fp32 tolerance."""
import numpy as np
import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu

RTOL = 1e-4          # north_star: images and gradients within rtol 1e-4
ATOL_SCALE = 1e-6    # atol = 1e-6 * max|reference tensor|


def _need_ref():
    from oracle import ref_cuda
    if not ref_cuda.available("dgr"):
        pytest.skip("oracle/_ref/libref_dgr.so not built (make -C oracle ref needs /root/reference)")
    return ref_cuda


def _bits(t):
    return t.contiguous().view(torch.int32) if t.dtype.is_floating_point else t


def _assert_exact(name, a, b):
    a, b = _bits(a), _bits(b)
    assert a.shape == b.shape, f"{name}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    n = int((a != b).sum())
    assert n == 0, f"{name}: {n} of {a.numel()} elements differ"


def _grad_ok(name, ours, ref_runs):
    """ours within tolerance of the mean of the reference runs, widened by their own spread."""
    st = torch.stack([r.double().reshape(-1) for r in ref_runs])
    mean = st.mean(0)
    spread = float((st.max(0).values - st.min(0).values).max())
    atol = ATOL_SCALE * float(mean.abs().max()) + 2.0 * spread
    err = (ours.double().reshape(-1) - mean).abs()
    tol = atol + RTOL * mean.abs()
    bad = int((err > tol).sum())
    assert bad == 0, f"{name}: {bad} elements beyond tolerance, worst ratio {float((err / tol).max()):.3g}"


CASES = [
    ("tiny_ragged", "C1", dict(P=5000, W=200, H=136)),
    ("C1", "C1", dict()),
    ("C5_proxy", "C5", dict()),
    ("C2_100k", "C2", dict(P=100_000)),
    ("sh1", "C2", dict(P=20_000, W=320, H=200, sh_degree=1)),
    ("sh2_bg", "C2", dict(P=20_000, W=320, H=200, sh_degree=2, random_bg=True)),
    ("dense_long_lists", "C1", dict(P=200_000, W=96, H=64)),
    ("dense_mid_lists", "C1", dict(P=60_000, W=96, H=64)),
]


def test_sort_size_classes_are_covered():
    """binning.cu sorts tile segments with three kernels chosen by length (<= 1024, <= 8192, longer); the parity
    cases above must reach all three, otherwise their bit-exact list checks prove less than they seem to."""
    longest = {}
    for name, kw in (("C1", dict()), ("mid", dict(P=60_000, W=96, H=64)), ("long", dict(P=200_000, W=96, H=64))):
        sc = h.make("C1", **kw)
        rg = h.ours_intermediates(h.ours_forward(sc, sc.cameras[0]))["ranges"].long()
        longest[name] = int((rg[:, 1] - rg[:, 0]).max())
    assert longest["C1"] <= 1024 < longest["mid"] <= 8192 < longest["long"], longest


@pytest.mark.parametrize("name,cfg,kw", CASES, ids=[c[0] for c in CASES])
def test_forward_bit_exact_and_backward_close(name, cfg, kw):
    ref_cuda = _need_ref()
    sc = h.make(cfg, **kw); cam = sc.cameras[0]
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    _assert_exact("radii", fo["radii"], fr["radii"])
    vis = fr["radii"] > 0
    for k in ("means2D", "depths", "conic_opacity", "rgb", "cov3D", "tiles_touched"):
        _assert_exact(k, io[k][vis], ir[k][vis])
    cl_ref = ir["clamped"][vis].to(torch.int32)
    cl_ours = io["clamped"][vis].to(torch.int32)
    _assert_exact("clamped", cl_ours, cl_ref[:, 0] + 2 * cl_ref[:, 1] + 4 * cl_ref[:, 2])
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    _assert_exact("ranges", io["ranges"], ir["ranges"])
    _assert_exact("n_contrib", io["n_contrib"], ir["n_contrib"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5)]
    go = h.ours_backward(sc, cam, fo, grads)
    runs = [h.ref_backward(sc, cam, fr, grads) for _ in range(4)]
    for k in ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"):
        _grad_ok("dL_d" + k, go[k], [r[k] for r in runs])


def test_full_size_c2_bit_exact():
    """BASELINE.json config[1] at full size: 1M Gaussians, SH3, 1297x840."""
    ref_cuda = _need_ref()
    sc = h.make("C2"); cam = sc.cameras[0]
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    _assert_exact("radii", fo["radii"], fr["radii"])
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    _assert_exact("ranges", io["ranges"], ir["ranges"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])
    # size-independent properties: sorted keys, ranges partition the list, every instance's tile matches
    keys = io["point_list_keys"]
    assert bool((keys[1:] >= keys[:-1]).all())
    rg = io["ranges"].long()
    assert int((rg[:, 1] - rg[:, 0]).sum()) == fo["num_rendered"]
    assert int(io["tiles_touched"].long().sum()) == fo["num_rendered"]
    # and the gradients at full size (rtol 1e-4 + the reference's own atomics spread)
    _full_size_gradients(sc, cam, fo, fr)


def _full_size_gradients(sc, cam, fo, fr):
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5)]
    go = h.ours_backward(sc, cam, fo, grads)
    runs = [h.ref_backward(sc, cam, fr, grads) for _ in range(4)]
    for k in ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"):
        _grad_ok("dL_d" + k, go[k], [r[k] for r in runs])
    # invisible Gaussians get exact zeros, as the reference's zero-filled tensors keep
    inv = fr["radii"] == 0
    for k in ("means3D", "sh", "opacity", "scales", "rotations", "means2D"):
        assert float(go[k][inv].abs().max()) == 0.0, k


def test_full_size_c3_shape_bit_exact_and_gradients():
    """BASELINE.json config[2]'s shape on one GPU: 6M Gaussians, SH3, 1920x1080 (one view of the 8)."""
    ref_cuda = _need_ref()
    sc = h.make("C3"); cam = sc.cameras[0]
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    _assert_exact("radii", fo["radii"], fr["radii"])
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    _assert_exact("ranges", io["ranges"], ir["ranges"])
    _assert_exact("n_contrib", io["n_contrib"], ir["n_contrib"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])
    rg = io["ranges"].long()
    assert int((rg[:, 1] - rg[:, 0]).max()) > 1024          # the crowded-tile sort path is the one exercised here
    _full_size_gradients(sc, cam, fo, fr)


def test_precomputed_colors_and_covariance():
    _need_ref()
    sc = h.make("C1", P=20_000, W=256, H=192, random_bg=True); cam = sc.cameras[0]
    g = torch.Generator().manual_seed(3)
    colors = torch.rand(sc.P, 3, generator=g).cuda()
    fr0 = h.ref_forward(sc, cam)
    from oracle import ref_cuda
    cov = ref_cuda.dgr_intermediates(fr0)["cov3D"].clone()
    fo = h.ours_forward(sc, cam, colors_precomp=colors, cov3D_precomp=cov)
    fr = h.ref_forward(sc, cam, colors_precomp=colors, cov3D_precomp=cov)
    assert fo["num_rendered"] == fr["num_rendered"]
    for k in ("color", "depth", "alpha", "radii"):
        _assert_exact(k, fo[k], fr[k])
    grads = [x.cuda() for x in synth.upstream_grads(cam.W, cam.H, 6)]
    go = h.ours_backward(sc, cam, fo, grads)
    runs = [h.ref_backward(sc, cam, fr, grads) for _ in range(4)]
    for k in ("means2D", "colors", "opacity", "means3D", "cov3D"):
        _grad_ok("dL_d" + k, go[k], [r[k] for r in runs])
    assert float(go["scales"].abs().max()) == 0.0 and float(go["rotations"].abs().max()) == 0.0


def test_other_views_and_scale_modifier():
    _need_ref()
    sc = h.make("C2", P=50_000, W=400, H=260, n_views=4)
    for v in (1, 2, 3):
        cam = sc.cameras[v]
        fo = h.ours_forward(sc, cam, scale_modifier=1.7); fr = h.ref_forward(sc, cam, scale_modifier=1.7)
        assert fo["num_rendered"] == fr["num_rendered"]
        for k in ("color", "depth", "alpha", "radii"):
            _assert_exact(f"view{v}.{k}", fo[k], fr[k])


def test_mark_visible_matches_reference():
    ref_cuda = _need_ref()
    from scorp_b200.diff_gaussian_rasterization import _C
    sc = h.make("C1", P=30_000)
    cam = synth.make_camera(512, 512, T=(0.0, 0.0, -5.0))      # view z = world z - 5: part of the cube is behind
    cam.viewmatrix = cam.viewmatrix.cuda(); cam.projmatrix = cam.projmatrix.cuda()
    ours = _C.mark_visible(sc.means3D, cam.viewmatrix, cam.projmatrix)
    ref = ref_cuda.dgr_mark_visible(sc.means3D, cam.viewmatrix, cam.projmatrix)
    assert torch.equal(ours, ref) and 0 < int(ours.sum()) < sc.P


def test_against_cpu_oracle():
    """Same seeded inputs through the CPU restatement (oracle/gs_oracle.c)."""
    from oracle import gs_oracle as O
    sc_cpu = synth.make_scene("C1", P=8000, W=176, H=120, random_bg=True); cam_c = sc_cpu.cameras[0]
    f = O.forward(sc_cpu.bg, sc_cpu.means3D, None, sc_cpu.opacities, sc_cpu.scales, sc_cpu.rotations, 1.0, None,
                  cam_c.viewmatrix, cam_c.projmatrix, cam_c.tanfovx, cam_c.tanfovy, cam_c.H, cam_c.W, sc_cpu.shs,
                  sc_cpu.sh_degree, cam_c.campos)
    gr = synth.upstream_grads(cam_c.W, cam_c.H, 9)
    g = O.backward(f, *gr)
    sc = h.make("C1", P=8000, W=176, H=120, random_bg=True); cam = sc.cameras[0]
    fo = h.ours_forward(sc, cam)
    go = h.ours_backward(sc, cam, fo, [x.cuda() for x in gr])
    radii = fo["radii"].cpu().numpy()
    assert (radii != f.radii).mean() < 1e-3
    for k in ("color", "depth", "alpha"):
        np.testing.assert_allclose(fo[k].cpu().numpy(), getattr(f, k), rtol=1e-3, atol=2e-4)
    for k in ("means2D", "opacity", "means3D", "sh", "scales", "rotations"):
        a = go[k].cpu().numpy().reshape(-1); b = g[k].reshape(-1)
        scale = np.abs(b).max()
        assert np.abs(a - b).max() <= 2e-3 * scale + 1e-12, k


@pytest.mark.parametrize("name,P,W,H,plane", [
    ("clones_small_lists", 40_000, 320, 208, False),     # lists of a few hundred entries
    ("clones_mid_lists", 60_000, 96, 64, False),         # 1024 < L <= 8192: the radix classes
    ("one_depth_plane", 12_000, 96, 64, True),           # EVERY Gaussian at the same depth: runs as long as the lists
])
def test_equal_depths_come_out_in_index_order(name, P, W, H, plane):
    """The reference's order inside a tile is (depth bits, Gaussian index): its radix sort is stable and
    duplicateWithKeys emits instances in index order (rasterizer_impl.cu:98-109,301-309). Cloned Gaussians
    (densify_and_clone copies xyz) share their depth bit for bit, so ties are a real input; here half of the Gaussians
    are clones of the other half at shuffled indices, or all of them lie in one plane facing the camera."""
    ref_cuda = _need_ref()
    sc = h.make("C1", P=P, W=W, H=H, random_bg=True); cam = sc.cameras[0]
    g = torch.Generator().manual_seed(12)
    if plane:
        sc.means3D[:, 2] = 6.0                           # camera 0 is the identity: depth = z exactly
    else:
        perm = torch.randperm(P, generator=g).cuda()
        src, dst = perm[: P // 2], perm[P // 2: 2 * (P // 2)]
        sc.means3D[dst] = sc.means3D[src]                # same position, different scale / rotation / colour
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    keys = ir["point_list_keys"]
    ties = int((keys[1:] == keys[:-1]).sum())
    assert ties > (fo["num_rendered"] // 4 if not plane else fo["num_rendered"] // 2), ties
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    _assert_exact("n_contrib", io["n_contrib"], ir["n_contrib"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])


@pytest.mark.parametrize("name,P,W,H,width", [
    ("cluster_two_pass", 60_000, 96, 64, 2e-2),          # buckets overflow, the 16-bit map still separates the keys
    ("cluster_four_pass", 60_000, 96, 64, 2e-6),         # runs of equal 16-bit keys too: the full sort on the depth bits
    ("cluster_small_lists", 40_000, 320, 208, 2e-4),     # the same in the small size class (lists of a few hundred)
])
def test_clustered_depths_take_the_fallback_sorts(name, P, W, H, width):
    """The per-tile sort first tries ONE counting pass over buckets of the tile's depth range (binning.cu,
    tile_bucket_sort); 97 % of the Gaussians squeezed into a thin slab with the rest spread over the whole range
    overfills its buckets, so the tile goes to the 16-bit two-pass sort, and for a thinner slab to the four-pass sort.
    Every level must give the reference's lists bit for bit."""
    ref_cuda = _need_ref()
    sc = h.make("C1", P=P, W=W, H=H, random_bg=True); cam = sc.cameras[0]
    g = torch.Generator().manual_seed(5)
    z = sc.means3D[:, 2].clone()
    slab = (torch.rand(P, generator=g) < 0.97).cuda()
    z[slab] = 6.0 + width * torch.rand(P, generator=g).cuda()[slab]         # camera 0 is the identity: depth = z
    sc.means3D[:, 2] = z
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    _assert_exact("n_contrib", io["n_contrib"], ir["n_contrib"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])


def test_screen_filling_splats_on_a_large_tile_grid():
    """64 Gaussians scaled up until each covers the whole 4096x4096 image (65536 tiles): a warp's 32 rectangles hold
    ~2^21 tiles, beyond the 2^20 from which the packed rectangle walk (common.cuh, warp_rects_packed) divides with integers
    instead of a float reciprocal, and every tile's list holds all 64. Lists, ranges and image bit-exact as always."""
    ref_cuda = _need_ref()
    sc = h.make("C1", P=64, W=4096, H=4096, random_bg=True); cam = sc.cameras[0]
    sc.scales[:] = 60.0
    sc.opacities[:] = 0.05
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    tiles = (4096 // 16) * (4096 // 16)
    full = int((fr["radii"] > 0).sum()) * tiles
    assert fr["num_rendered"] > 0.9 * full and fr["num_rendered"] > 40 * tiles, (fr["num_rendered"], full)      # each warp: > 2^20
    assert fo["num_rendered"] == fr["num_rendered"]
    io, ir = h.ours_intermediates(fo), ref_cuda.dgr_intermediates(fr)
    _assert_exact("ranges", io["ranges"], ir["ranges"])
    _assert_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
    _assert_exact("point_list", io["point_list"], ir["point_list"])
    for k in ("color", "depth", "alpha"):
        _assert_exact(k, fo[k], fr[k])
