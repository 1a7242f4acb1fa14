"""The reference-facing Python API on the GPU: GaussianRasterizer / GaussianRasterizationSettings used
exactly like gs3dgs/gaussian_renderer/__init__.py:39-114 does, incl. autograd, repeated backward
(utils/mask.py:52-89), validation errors and empty input."""
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


def test_render_like_gs3dgs_and_autograd_matches_C_level():
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make("C1", P=20_000, W=256, H=192); cam = sc.cameras[0]
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    means3D, shs, opac, scales, rots = params
    screenspace = torch.zeros_like(means3D, requires_grad=True) + 0
    screenspace.retain_grad()
    rast = dgr.GaussianRasterizer(raster_settings=_settings(sc, cam, dgr))
    color, radii, depth, alpha = rast(means3D=means3D, means2D=screenspace, shs=shs, colors_precomp=None,
                                      opacities=opac, scales=scales, rotations=rots, cov3D_precomp=None)
    assert color.shape == (3, cam.H, cam.W) and depth.shape == (1, cam.H, cam.W) and alpha.shape == (1, cam.H, cam.W)
    assert radii.dtype == torch.int32 and radii.shape == (sc.P,)
    gc, gd, ga = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 21)]
    loss = (color * gc).sum() + (depth * gd).sum() + (alpha * ga).sum()
    loss.backward(retain_graph=True)
    fo = h.ours_forward(sc, cam)
    runs = [h.ours_backward(sc, cam, fo, [gc, gd, ga]) for _ in range(4)]
    assert torch.equal(color.detach(), fo["color"]) and torch.equal(radii, fo["radii"])
    # separate backward runs: equal up to the order of float atomics, i.e. within the spread of the four C-level runs
    def same(a, k):
        return h.close_to_runs(a, [r[k] for r in runs], rtol=1e-3, atol_scale=1e-5)[0]
    for p, k in zip(params, ("means3D", "sh", "opacity", "scales", "rotations")):
        assert same(p.grad, k), k
    assert same(screenspace.grad, "means2D")
    assert float(screenspace.grad[:, 2].abs().max()) == 0.0
    # backward is re-entrant: a second pass over the same graph ADDS its gradients to the accumulated ones
    first = means3D.grad.clone()
    loss.backward()
    assert same(means3D.grad - first, "means3D")


def test_override_color_gradient_channels_equal():
    """utils/mask.py:48-69: constant colours -> the three colour-gradient channels are equal."""
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make("C1", P=10_000, W=128, H=128); cam = sc.cameras[0]
    col = torch.ones(sc.P, 3, device="cuda", requires_grad=True)
    rast = dgr.GaussianRasterizer(_settings(sc, cam, dgr))
    color, radii, depth, alpha = rast(sc.means3D, torch.zeros_like(sc.means3D), sc.opacities, colors_precomp=col,
                                      scales=sc.scales, rotations=sc.rotations)
    w = torch.rand(1, cam.H, cam.W, device="cuda")
    (color * w).sum().backward(retain_graph=True)
    g = col.grad
    assert torch.allclose(g[:, 0], g[:, 1], rtol=1e-5, atol=1e-9) and torch.allclose(g[:, 0], g[:, 2], rtol=1e-5, atol=1e-9)
    assert float(g.abs().max()) > 0


def test_argument_validation_and_empty():
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make("C1", P=100, W=64, H=48); cam = sc.cameras[0]
    rast = dgr.GaussianRasterizer(_settings(sc, cam, dgr))
    z = torch.zeros_like(sc.means3D)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(sc.means3D, z, sc.opacities, scales=sc.scales, rotations=sc.rotations)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(sc.means3D, z, sc.opacities, shs=sc.shs, colors_precomp=torch.ones(100, 3).cuda(), scales=sc.scales,
             rotations=sc.rotations)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        rast(sc.means3D, z, sc.opacities, shs=sc.shs, scales=sc.scales)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        rast(sc.means3D, z, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
             cov3D_precomp=torch.ones(100, 6).cuda())
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rast(sc.means3D[:, :2], z, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    # P == 0: zero images (rasterize_points.cu:83)
    e = torch.zeros(0, 3, device="cuda")
    color, radii, depth, alpha = rast(e, e, torch.zeros(0, 1, device="cuda"), shs=torch.zeros(0, 1, 3, device="cuda"),
                                      scales=e, rotations=torch.zeros(0, 4, device="cuda"))
    assert radii.numel() == 0 and float(color.abs().max()) == 0 and float(alpha.abs().max()) == 0
    # nothing visible: background image
    far = sc.means3D.clone(); far[:, 2] = -5.0
    sc2 = h.make("C1", P=100, W=64, H=48, random_bg=True)
    rast2 = dgr.GaussianRasterizer(_settings(sc2, sc2.cameras[0], dgr))
    color, radii, depth, alpha = rast2(far, z, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    assert int(radii.abs().sum()) == 0
    assert torch.equal(color, sc2.bg.view(3, 1, 1).expand(3, 48, 64))


def test_non_contiguous_and_dropin_import_name():
    import scorp_b200
    scorp_b200.install()
    import diff_gaussian_rasterization as d2
    import scorp_b200.diff_gaussian_rasterization as d1
    assert d2.GaussianRasterizer is d1.GaussianRasterizer
    sc = h.make("C1", P=5000, W=96, H=80); cam = sc.cameras[0]
    rast = d2.GaussianRasterizer(_settings(sc, cam, d2))
    wide = torch.zeros(sc.P, 6, device="cuda"); wide[:, ::2] = sc.means3D
    a = rast(wide[:, ::2], torch.zeros_like(sc.means3D), sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    b = rast(sc.means3D, torch.zeros_like(sc.means3D), sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    vis = rast.markVisible(sc.means3D)
    assert vis.dtype == torch.bool and vis.shape == (sc.P,)


@pytest.mark.parametrize("surfel", [False, True])
def test_no_reference_cycles_hold_device_memory(surfel):
    """With the cyclic garbage collector off, repeated forward+backward through the public API must not grow the
    allocated device memory: every scratch buffer has to die with the autograd graph (a bound-method ctypes callback
    once kept 140 MB per step alive until the next gc pass and produced multi-ms hiccups)."""
    import gc
    import scorp_b200.diff_gaussian_rasterization as dgr
    import scorp_b200.diff_surfel_rasterization as dsr
    mod = dsr if surfel else dgr
    sc = h.make("C4" if surfel else "C1", P=50_000, W=320, H=208); cam = sc.cameras[0]
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]

    def step():
        for p in params:
            p.grad = None
        rast = mod.GaussianRasterizer(raster_settings=_settings(sc, cam, mod))
        outs = rast(means3D=params[0], means2D=torch.zeros_like(params[0], requires_grad=True), shs=params[1],
                    opacities=params[2], scales=params[3], rotations=params[4])
        (outs[0].sum() + outs[2].sum()).backward()

    step(); step()
    gc.collect(); gc.disable()
    try:
        torch.cuda.synchronize(); base = torch.cuda.memory_allocated()
        for _ in range(8):
            step()
        torch.cuda.synchronize(); grown = torch.cuda.memory_allocated() - base
    finally:
        gc.enable()
    assert grown < 4 * 1024 * 1024, f"{grown / 1e6:.1f} MB kept alive by reference cycles"
