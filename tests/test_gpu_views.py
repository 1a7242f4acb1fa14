"""Batched multi-view forward on the GPU: one gsr_rasterize_forward_views call over V cameras
must equal, bit for bit, V forwards of the REFERENCE CUDA (oracle/_ref/libref_dgr.so) run as the Python loop of
align_3dgs_clpe_9dof.py:146-160, and V single-view forwards of this library."""
import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu


def _stack(cams):
    return (torch.stack([c.viewmatrix for c in cams]), torch.stack([c.projmatrix for c in cams]),
            torch.stack([c.campos for c in cams]))


def _views(sc, cams, colors_precomp=None, scale_modifier=1.0):
    from scorp_b200.diff_gaussian_rasterization import _C
    E = h.empty()
    vm, pm, cp = _stack(cams)
    return _C.rasterize_gaussians_views(sc.bg, sc.means3D, E if colors_precomp is None else colors_precomp, sc.opacities,
                                        sc.scales, sc.rotations, scale_modifier, E, vm, pm, cams[0].tanfovx, cams[0].tanfovy,
                                        cams[0].H, cams[0].W, sc.shs if colors_precomp is None else E, sc.sh_degree, cp,
                                        False, False)


@pytest.mark.parametrize("name,kw,V", [
    ("C5", dict(P=40_000, W=256, H=256), 7),        # SH0 pose sweep with random rotations
    ("C2", dict(P=50_000, W=333, H=190), 4),        # SH3: view-dependent colours, ragged image size
    ("C1", dict(P=5_000, W=64, H=48), 1),           # V = 1 degenerates to the single-view path
])
def test_batched_views_equal_single_view_loop(name, kw, V):
    rots = synth.random_rotations(V, seed=128)
    sc = synth.make_scene(name, n_views=V, rotations_npz=rots, random_bg=True, **kw).to("cuda")
    n, color, depth, alpha, radii = _views(sc, sc.cameras)
    assert color.shape == (V, 3, kw["H"], kw["W"]) and radii.shape == (V, sc.P)
    from oracle import ref_cuda
    if not ref_cuda.available("dgr"):
        pytest.skip("oracle/_ref/libref_dgr.so not built (make -C oracle ref needs /root/reference)")
    total = 0
    for v, cam in enumerate(sc.cameras):
        fr = h.ref_forward(sc, cam)                     # the reference CUDA, one call per view
        assert torch.equal(radii[v], fr["radii"]), f"view {v}: radii vs reference"
        for nm, mine in (("color", color), ("depth", depth), ("alpha", alpha)):
            assert torch.equal(mine[v], fr[nm]), f"view {v}: {nm} vs reference"
        fo = h.ours_forward(sc, cam)
        assert fo["num_rendered"] == fr["num_rendered"]
        total += fo["num_rendered"]
        assert torch.equal(radii[v], fo["radii"]), f"view {v}: radii"
        assert torch.equal(color[v], fo["color"]), f"view {v}: colour"
        assert torch.equal(depth[v], fo["depth"]) and torch.equal(alpha[v], fo["alpha"]), f"view {v}: depth/alpha"
    assert n == total
    assert len({int(r.sum()) for r in radii}) > 1 or V == 1      # the views really differ


def test_batched_views_shared_precomputed_colors_and_public_api():
    import scorp_b200.diff_gaussian_rasterization as dgr
    V = 3
    sc = synth.make_scene("C1", P=20_000, W=200, H=120, n_views=V).to("cuda")
    col = torch.rand(sc.P, 3, device="cuda")
    n, color, depth, alpha, radii = _views(sc, sc.cameras, colors_precomp=col, scale_modifier=0.8)
    from oracle import ref_cuda
    for v, cam in enumerate(sc.cameras):
        fo = h.ours_forward(sc, cam, colors_precomp=col, scale_modifier=0.8)
        assert torch.equal(color[v], fo["color"]) and torch.equal(radii[v], fo["radii"])
        if ref_cuda.available("dgr"):
            fr = h.ref_forward(sc, cam, colors_precomp=col, scale_modifier=0.8)
            assert torch.equal(color[v], fr["color"]) and torch.equal(radii[v], fr["radii"]) and torch.equal(depth[v], fr["depth"])
    cam = sc.cameras[0]
    rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, sc.sh_degree, cam.campos, False, False)
    rast = dgr.GaussianRasterizer(rs)
    vm, pm, cp = _stack(sc.cameras)
    c2, r2, d2, a2 = rast.forward_views(vm, pm, cp, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales,
                                        rotations=sc.rotations)
    c1, r1, d1, a1 = rast(sc.means3D, torch.zeros_like(sc.means3D), sc.opacities, shs=sc.shs, scales=sc.scales,
                          rotations=sc.rotations)
    assert torch.equal(c2[0], c1) and torch.equal(r2[0], r1) and torch.equal(d2[0], d1) and torch.equal(a2[0], a1)
    with pytest.raises(RuntimeError, match="viewmatrices"):
        rast.forward_views(vm[0], pm, cp, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations)
    # no Gaussians: zero images, as the single-view path (rasterize_points.cu:83)
    z = lambda *s: torch.zeros(*s, device="cuda")
    c0, r0, _, _ = rast.forward_views(vm, pm, cp, z(0, 3), z(0, 1), shs=z(0, 1, 3), scales=z(0, 3), rotations=z(0, 4))
    assert c0.shape == (V, 3, cam.H, cam.W) and float(c0.abs().max()) == 0.0 and r0.shape == (V, 0)


def test_batched_views_with_the_reference_rotations_equal_reference_cuda():
    """C5 as named: poses from rotation_matrices/rotations_128.npz (tests/golden/), every 8th of the 128."""
    import numpy as np
    from pathlib import Path
    from oracle import ref_cuda
    if not ref_cuda.available("dgr"):
        pytest.skip("oracle/_ref/libref_dgr.so not built")
    rots = np.load(Path(__file__).resolve().parent / "golden" / "rotations_128.npz")["rotations"][::8]
    cams = [synth.orbit_camera(512, 512, R) for R in rots]
    sc = synth.make_scene("C5", cameras=cams).to("cuda")
    n, color, depth, alpha, radii = _views(sc, sc.cameras)
    total = 0
    for v, cam in enumerate(sc.cameras):
        fr = h.ref_forward(sc, cam)
        total += fr["num_rendered"]
        assert torch.equal(radii[v], fr["radii"]) and torch.equal(color[v], fr["color"]), f"pose {8 * v}"
        assert torch.equal(depth[v], fr["depth"]) and torch.equal(alpha[v], fr["alpha"]), f"pose {8 * v}"
    assert n == total
