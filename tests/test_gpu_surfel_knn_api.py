"""The reference-facing Python API of the 2DGS package and simple_knn on the GPU, used the way
gs2dgs/gaussian_renderer/__init__.py:24-150 and gs2dgs/scene/gaussian_model.py:139 use them."""
import numpy as np
import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu


def _settings(sc, cam, mod):
    return mod.GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=sc.bg,
        scale_modifier=1.0, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, sh_degree=sc.sh_degree,
        campos=cam.campos, prefiltered=False, debug=False)


def test_render_like_gs2dgs_and_autograd_matches_C_level():
    import scorp_b200.diff_surfel_rasterization as dsr
    sc = h.make("C4", P=20_000, W=256, H=192); cam = sc.cameras[0]
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    means3D, shs, opac, scales, rots = params
    screenspace = torch.zeros_like(means3D, requires_grad=True) + 0
    screenspace.retain_grad()
    rast = dsr.GaussianRasterizer(raster_settings=_settings(sc, cam, dsr))
    color, radii, allmap = rast(means3D=means3D, means2D=screenspace, shs=shs, colors_precomp=None, opacities=opac,
                                scales=scales, rotations=rots, cov3D_precomp=None)
    assert color.shape == (3, cam.H, cam.W) and allmap.shape == (7, cam.H, cam.W)
    assert radii.dtype == torch.int32 and radii.shape == (sc.P,)
    # the post-processing of gs2dgs render(): alpha, normal, median depth, expected depth, distortion
    render_alpha = allmap[1:2]; render_normal = allmap[2:5]; render_depth_expected = allmap[0:1] / render_alpha.clamp_min(1e-8)
    assert bool(torch.isfinite(render_depth_expected).all()) and bool(torch.isfinite(render_normal).all())
    gc, go7 = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 21, channels=(3, 7))]
    loss = (color * gc).sum() + (allmap * go7).sum()
    loss.backward(retain_graph=True)
    fo = h.ours_surfel_forward(sc, cam)
    runs = [h.ours_surfel_backward(sc, cam, fo, [gc, go7]) for _ in range(4)]
    assert torch.equal(color.detach(), fo["color"]) and torch.equal(allmap.detach(), fo["others"]) and torch.equal(radii, fo["radii"])

    def same(a, k):      # separate backward runs: within the spread of the four C-level runs (order of float atomics)
        return h.close_to_runs(a, [r[k] for r in runs], rtol=1e-3, atol_scale=1e-5)[0]
    for p, k in zip(params, ("means3D", "sh", "opacity", "scales", "rotations")):
        assert same(p.grad, k), k
    assert same(screenspace.grad, "means2D")
    assert float(screenspace.grad[:, 2].abs().max()) == 0.0
    first = means3D.grad.clone()
    loss.backward()
    assert same(means3D.grad - first, "means3D")


def test_surfel_validation_empty_and_dropin_names():
    import scorp_b200
    scorp_b200.install()
    import diff_surfel_rasterization as d2
    import scorp_b200.diff_surfel_rasterization as d1
    from simple_knn._C import distCUDA2 as k2
    from scorp_b200.simple_knn._C import distCUDA2 as k1
    assert d2.GaussianRasterizer is d1.GaussianRasterizer and k1 is k2
    sc = h.make("C4", P=100, W=64, H=48, random_bg=True); cam = sc.cameras[0]
    rast = d2.GaussianRasterizer(_settings(sc, cam, d2))
    z = torch.zeros_like(sc.means3D)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(sc.means3D, z, sc.opacities, scales=sc.scales, rotations=sc.rotations)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        rast(sc.means3D, z, sc.opacities, shs=sc.shs, scales=sc.scales)
    e = torch.zeros(0, 3, device="cuda")
    color, radii, allmap = rast(e, e, torch.zeros(0, 1, device="cuda"), shs=torch.zeros(0, 16, 3, device="cuda"),
                                scales=torch.zeros(0, 2, device="cuda"), rotations=torch.zeros(0, 4, device="cuda"))
    assert radii.numel() == 0 and float(color.abs().max()) == 0 and float(allmap.abs().max()) == 0
    far = sc.means3D.clone(); far[:, 2] = -5.0
    color, radii, allmap = rast(far, z, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    assert int(radii.abs().sum()) == 0 and float(allmap.abs().max()) == 0
    assert torch.equal(color, sc.bg.view(3, 1, 1).expand(3, 48, 64))


def _brute_knn(pts):
    p = pts.double().cpu().numpy()
    out = np.empty(len(p), dtype=np.float64)
    for s in range(0, len(p), 2048):
        d = ((p[s:s + 2048, None, :] - p[None, :, :]) ** 2).sum(-1)
        d[np.arange(d.shape[0]), np.arange(s, s + d.shape[0])] = np.inf
        d.partition(2, axis=1)
        out[s:s + 2048] = d[:, :3].sum(1) / 3.0
    return out


KNN_CASES = {
    "uniform": lambda g: torch.rand(20_000, 3, generator=g) * 8 - 4,
    "clustered": lambda g: torch.cat([torch.randn(5_000, 3, generator=g) * s + c for s, c in
                                      ((0.05, 0.0), (0.5, 3.0), (2.0, -5.0), (0.001, 1.0))]),
    "planar": lambda g: torch.cat([torch.rand(10_000, 2, generator=g), torch.zeros(10_000, 1)], 1),
    "line": lambda g: torch.cat([torch.rand(5_000, 1, generator=g), torch.zeros(5_000, 2)], 1),
    "duplicates": lambda g: torch.rand(2_000, 3, generator=g).repeat(4, 1),
    "far_offset": lambda g: torch.rand(20_000, 3, generator=g) * 2 + 1000.0,
    "four_points": lambda g: torch.rand(4, 3, generator=g),
}


@pytest.mark.parametrize("name", list(KNN_CASES))
def test_distCUDA2_exact(name):
    from scorp_b200.simple_knn._C import distCUDA2
    from oracle import ref_cuda
    pts = KNN_CASES[name](torch.Generator().manual_seed(11)).cuda().contiguous()
    ours = distCUDA2(pts)
    assert ours.shape == (pts.shape[0],) and ours.dtype == torch.float32
    np.testing.assert_allclose(ours.cpu().numpy(), _brute_knn(pts), rtol=2e-5, atol=1e-12)
    if ref_cuda.available("knn"):
        ref = ref_cuda.knn_dist2(pts)
        assert torch.equal(ours.view(torch.int32), ref.view(torch.int32)), "not bit-identical to the reference"


def test_distCUDA2_full_size_and_model_init_use():
    """1M points (the C2/C4 scene size); used like gaussian_model.create_from_pcd: clamp_min + log(sqrt)."""
    from scorp_b200.simple_knn._C import distCUDA2
    from oracle import ref_cuda
    g = torch.Generator().manual_seed(5)
    pts = (torch.rand(1_000_000, 3, generator=g) * 8 - 4).cuda()
    d2 = torch.clamp_min(distCUDA2(pts), 1e-7)
    scales = torch.log(torch.sqrt(d2))[..., None].repeat(1, 2)
    assert bool(torch.isfinite(scales).all())
    if ref_cuda.available("knn"):
        assert torch.equal(distCUDA2(pts), ref_cuda.knn_dist2(pts))
    # order independence: a permutation of the input permutes the output
    perm = torch.randperm(pts.shape[0], generator=g).cuda()
    assert torch.equal(distCUDA2(pts[perm]), distCUDA2(pts)[perm])
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).numel() == 0
    with pytest.raises(Exception):
        distCUDA2(torch.zeros(4, 3))
