"""Model-space entry points on the GPU: rasterize_model on the tensors GaussianModel
stores must equal what the reference's caller computes -- torch.cat(features_dc, features_rest), sigmoid, exp,
normalize (gs3dgs/scene/gaussian_model.py:126-146) followed by the stock rasterizer -- in the forward, and the
autograd chain of those torch ops in the backward."""
import json
import os
from types import SimpleNamespace

import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu


def _stored(sc, seed=5):
    """Stored parameters whose activations reproduce the scene (rotation left un-normalised on purpose)."""
    g = torch.Generator().manual_seed(seed)
    dev = sc.means3D.device
    op = sc.opacities.clamp(1e-4, 1 - 1e-4)
    stored = dict(
        xyz=sc.means3D.clone(),
        features_dc=sc.shs[:, :1, :].contiguous().clone(),
        features_rest=sc.shs[:, 1:, :].contiguous().clone(),
        opacity=torch.log(op / (1 - op)),
        scaling=torch.log(sc.scales),
        rotation=sc.rotations * (0.25 + 3.0 * torch.rand(sc.P, 1, generator=g)).to(dev),
    )
    return {k: v.requires_grad_(True) for k, v in stored.items()}


def _settings(sc, cam, mod, scale_modifier=1.0):
    return mod.GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=sc.bg,
        scale_modifier=scale_modifier, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, sh_degree=sc.sh_degree,
        campos=cam.campos, prefiltered=False, debug=False)


def _stock(st, rs):
    """the reference caller's staging (torch kernels) + the stock drop-in rasterizer"""
    import scorp_b200.diff_gaussian_rasterization as dgr
    shs = torch.cat((st["features_dc"], st["features_rest"]), dim=1)
    m2d = torch.zeros_like(st["xyz"], requires_grad=True) + 0
    m2d.retain_grad()
    out = dgr.GaussianRasterizer(rs)(means3D=st["xyz"], means2D=m2d, opacities=torch.sigmoid(st["opacity"]), shs=shs,
                                     scales=torch.exp(st["scaling"]),
                                     rotations=torch.nn.functional.normalize(st["rotation"]))
    return out, m2d


def _model(st, rs):
    from scorp_b200 import model_space as ms
    m2d = torch.zeros_like(st["xyz"], requires_grad=True) + 0
    m2d.retain_grad()
    return ms.rasterize_model(st["xyz"], m2d, st["features_dc"], st["features_rest"], st["opacity"], st["scaling"],
                              st["rotation"], rs), m2d


def _note(name, **kw):
    """exact-equality observations for profiles/ (not assertions)"""
    os.makedirs(os.path.join(h.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(h.ROOT, "gpurun_out", "model_space_diag.jsonl"), "a") as f:
        f.write(json.dumps(dict(case=name, **kw)) + "\n")


@pytest.mark.parametrize("name,kw,smod", [
    ("C1", dict(P=30_000, W=256, H=192), 1.0),           # SH0: features_rest is empty
    ("C2", dict(P=60_000, W=320, H=208), 1.0),           # SH3
    ("C2", dict(P=20_000, W=200, H=120), 0.7),           # scale modifier
])
def test_forward_and_backward_match_torch_staging(name, kw, smod):
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make(name, **kw); cam = sc.cameras[0]
    rs = _settings(sc, cam, dgr, smod)
    sa, sb = _stored(sc), _stored(sc)
    (ca, ra, da, aa), m2a = _stock(sa, rs)
    (cb, rb, db, ab), m2b = _model(sb, rs)
    exact = dict(color=torch.equal(ca, cb), depth=torch.equal(da, db), alpha=torch.equal(aa, ab),
                 radii=torch.equal(ra, rb))
    _note(f"{name}-{kw['P']}-{smod}", **exact, radii_mismatch=int((ra != rb).sum()))
    # the activations are written as the float sequences torch's kernels execute, so everything is bit-identical
    assert torch.equal(ra, rb), f"{int((ra != rb).sum())} radii differ"
    assert torch.equal(ca, cb) and torch.equal(da, db) and torch.equal(aa, ab), exact

    gc, gd, ga = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 33)]
    ((cb * gc).sum() + (db * gd).sum() + (ab * ga).sum()).backward()
    # the stock side four times: two backward runs differ by the order of float atomics in the blend backward, and the
    # spread of these runs widens the tolerance (helpers.close_to_runs)
    loss_a = (ca * gc).sum() + (da * gd).sum() + (aa * ga).sum()
    runs_a, runs_m2a = [], []
    for _ in range(4):
        for t in sa.values():
            t.grad = None
        m2a.grad = None
        loss_a.backward(retain_graph=True)
        runs_a.append({k: (sa[k].grad.clone() if sa[k].grad is not None else None) for k in sa})
        runs_m2a.append(m2a.grad.clone())

    def same(x, ys):
        return h.close_to_runs(x, ys, rtol=1e-3, atol_scale=1e-5)[0]
    for k in sa:
        if sa[k].numel() == 0:
            assert sb[k].grad is None or sb[k].grad.numel() == 0
            continue
        assert sb[k].grad is not None and sb[k].grad.shape == sa[k].shape, k
        assert same(sb[k].grad, [r[k] for r in runs_a]), (k, float((sb[k].grad - sa[k].grad).abs().max()), float(sa[k].grad.abs().max()))
    assert same(m2b.grad, runs_m2a)
    # invisible Gaussians get exact zeros in every gradient row (every row is written, none left uninitialised)
    inv = rb == 0
    for k in sb:
        if sb[k].numel():
            assert float(sb[k].grad[inv].abs().max() if inv.any() else 0.0) == 0.0, k


def test_render_mirror_of_gaussian_renderer():
    """model_space.render() with a GaussianModel-shaped object returns gaussian_renderer.render()'s dict."""
    import math
    from scorp_b200 import model_space as ms
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make("C2", P=20_000, W=256, H=160); cam = sc.cameras[0]
    st = _stored(sc)
    pc = SimpleNamespace(get_xyz=st["xyz"], _features_dc=st["features_dc"], _features_rest=st["features_rest"],
                         _opacity=st["opacity"], _scaling=st["scaling"], _rotation=st["rotation"],
                         active_sh_degree=sc.sh_degree, max_sh_degree=3)
    vc = SimpleNamespace(FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy), resolution=(cam.W, cam.H),
                         world_view_transform=cam.viewmatrix, full_proj_transform=cam.projmatrix,
                         camera_center=cam.campos)
    pkg = ms.render(vc, pc, SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False), sc.bg)
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii", "render_depth", "render_alpha"}
    assert pkg["render"].shape == (3, cam.H, cam.W) and pkg["visibility_filter"].dtype == torch.bool
    (c, r, d, a), _ = _stock(_stored(sc), _settings(sc, cam, dgr))
    assert torch.allclose(pkg["render"], c, rtol=1e-4, atol=1e-6) and torch.equal(pkg["radii"], r)
    pkg["render"].sum().backward()
    assert pkg["viewspace_points"].grad is not None and st["features_rest"].grad is not None
    assert float(st["opacity"].grad.abs().max()) > 0


def test_empty_and_errors():
    from scorp_b200 import model_space as ms, _lib
    import scorp_b200.diff_gaussian_rasterization as dgr
    sc = h.make("C1", P=1_000, W=64, H=64); cam = sc.cameras[0]
    rs = _settings(sc, cam, dgr)
    z = lambda *s: torch.zeros(*s, device="cuda")
    c, r, d, a = ms.rasterize_model(z(0, 3), z(0, 3), z(0, 1, 3), z(0, 0, 3), z(0, 1), z(0, 3), z(0, 4), rs)
    assert r.numel() == 0 and float(c.abs().max()) == 0.0
    with pytest.raises(_lib.GsrError):
        ms.rasterize_model(torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(4, 1, 3), torch.zeros(4, 0, 3),
                           torch.zeros(4, 1), torch.zeros(4, 3), torch.zeros(4, 4), rs)
    with pytest.raises(RuntimeError, match="features_dc"):
        ms.rasterize_model(z(4, 3), z(4, 3), z(3, 1, 3), z(4, 0, 3), z(4, 1), z(4, 3), z(4, 4), rs)


@pytest.mark.parametrize("kw,smod", [(dict(P=40_000, W=300, H=200), 1.0), (dict(P=15_000, W=160, H=120, sh_degree=1), 0.8)])
def test_surfel_model_space_matches_torch_staging(kw, smod):
    """2DGS: rasterize_surfel_model == torch staging + the stock surfel rasterizer (gs2dgs render path)."""
    import scorp_b200.diff_surfel_rasterization as dsr
    from scorp_b200 import model_space as ms
    sc = h.make("C4", **kw); cam = sc.cameras[0]
    rs = _settings(sc, cam, dsr, smod)
    sa, sb = _stored(sc), _stored(sc)
    m2a = torch.zeros_like(sa["xyz"], requires_grad=True) + 0; m2a.retain_grad()
    m2b = torch.zeros_like(sb["xyz"], requires_grad=True) + 0; m2b.retain_grad()
    ca, ra, oa = dsr.GaussianRasterizer(rs)(
        means3D=sa["xyz"], means2D=m2a, opacities=torch.sigmoid(sa["opacity"]),
        shs=torch.cat((sa["features_dc"], sa["features_rest"]), dim=1), scales=torch.exp(sa["scaling"]),
        rotations=torch.nn.functional.normalize(sa["rotation"]))
    cb, rb, ob = ms.rasterize_surfel_model(sb["xyz"], m2b, sb["features_dc"], sb["features_rest"], sb["opacity"],
                                           sb["scaling"], sb["rotation"], rs)
    assert torch.equal(ra, rb), f"{int((ra != rb).sum())} radii differ"
    assert torch.equal(ca, cb) and torch.equal(oa, ob)
    gc, go = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 17, channels=(3, 7))]
    ((cb * gc).sum() + (ob * go).sum()).backward()
    loss_a = (ca * gc).sum() + (oa * go).sum()
    runs_a, runs_m2a = [], []          # the stock side four times: the spread of the float atomics widens the tolerance
    for _ in range(4):
        for t in sa.values():
            t.grad = None
        m2a.grad = None
        loss_a.backward(retain_graph=True)
        runs_a.append({k: (sa[k].grad.clone() if sa[k].grad is not None else None) for k in sa})
        runs_m2a.append(m2a.grad.clone())

    def same(x, ys):
        return h.close_to_runs(x, ys, rtol=1e-3, atol_scale=1e-5)[0]
    for k in sa:
        if sa[k].numel():
            assert sb[k].grad is not None and same(sb[k].grad, [r[k] for r in runs_a]), \
                (k, float((sb[k].grad - sa[k].grad).abs().max()), float(sa[k].grad.abs().max()))
    assert same(m2b.grad, runs_m2a)
    inv = rb == 0
    for k in sb:
        if sb[k].numel() and inv.any():
            assert float(sb[k].grad[inv].abs().max()) == 0.0, k
