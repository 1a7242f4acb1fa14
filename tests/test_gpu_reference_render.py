"""The drop-in boundary, exercised by the reference's OWN caller: the unmodified `render()` of
gs3dgs/gaussian_renderer/__init__.py:24-133 and gs2dgs/gaussian_renderer/__init__.py:24-160 (copied, git-ignored, to
baseline/_ref/ by tools/populate_baseline_ref.py) is executed twice on the same stub model and camera --
  (a) on top of scorp_b200.install(): `import diff_gaussian_rasterization` / `diff_surfel_rasterization` resolve to
      this repo's packages, exactly as they would inside SCORP;
  (b) on top of the reference CUDA (tests/ref_modules.py over oracle/_ref)
-- and every entry of the returned dict must be bit-identical (the 2DGS render() post-processes the 7-channel map into
alpha, normals, median / expected depth, distortion and the pseudo surface normal, which is the acceptance test of that
layout), the gradients of a scalar loss over all outputs within rtol 1e-4 (+ the reference's atomics spread).
The model / camera / pipeline objects are stubs exposing what render() reads (get_xyz, get_opacity, get_scaling,
get_rotation, get_features, active_sh_degree; FoVx, FoVy, resolution, world_view_transform, full_proj_transform,
camera_center, znear, zfar; debug, compute_cov3D_python, convert_SHs_python, depth_ratio)."""
import importlib.util
import math
import sys
import types
from pathlib import Path

import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
REF = ROOT / "baseline" / "_ref"


class _Model:
    def __init__(self, sc, surfel):
        self.p = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
        self.active_sh_degree = self.max_sh_degree = sc.sh_degree
    get_xyz = property(lambda s: s.p[0])
    get_features = property(lambda s: s.p[1])
    get_opacity = property(lambda s: s.p[2])
    get_scaling = property(lambda s: s.p[3])
    get_rotation = property(lambda s: s.p[4])


class _Camera:
    def __init__(self, cam):
        self.FoVx, self.FoVy = 2.0 * math.atan(cam.tanfovx), 2.0 * math.atan(cam.tanfovy)
        self.resolution = (cam.W, cam.H)
        self.world_view_transform, self.full_proj_transform, self.camera_center = cam.viewmatrix, cam.projmatrix, cam.campos
        self.znear, self.zfar = 0.01, 100.0


class _Pipe:
    debug = False
    compute_cov3D_python = False
    convert_SHs_python = False
    depth_ratio = 0.0


def _load_render(pkg, raster_name, raster_module, tag):
    """Execute baseline/_ref/<pkg>/gaussian_renderer/__init__.py as `<tag>.gaussian_renderer` with `raster_name` bound to
    `raster_module`; the sibling modules render() imports but this test does not need are stubs, utils come from the copies."""
    src = REF / pkg / "gaussian_renderer" / "__init__.py"
    if not src.is_file():
        pytest.skip("baseline/_ref not populated (tools/populate_baseline_ref.py needs /root/reference)")
    for name in ("cv2", "matplotlib", "matplotlib.pyplot"):        # point_utils imports them at module level, render() does not use them
        try:
            __import__(name)
        except Exception:
            sys.modules[name] = types.ModuleType(name)
    def pkg_mod(name, path=None):
        m = types.ModuleType(name)
        m.__path__ = [str(path)] if path else []
        sys.modules[name] = m
        return m
    pkg_mod(tag, REF / pkg); pkg_mod(f"{tag}.scene"); pkg_mod(f"{tag}.utils", REF / pkg / "utils")
    for name, attr in ((f"{tag}.scene.cameras", "Camera"), (f"{tag}.arguments", "PipelineParams"),
                       (f"{tag}.scene.gaussian_model", "GaussianModel")):
        m = types.ModuleType(name); setattr(m, attr, object); sys.modules[name] = m
    saved = sys.modules.get(raster_name)
    sys.modules[raster_name] = raster_module
    try:
        spec = importlib.util.spec_from_file_location(f"{tag}.gaussian_renderer", src,
                                                      submodule_search_locations=[str(src.parent)])
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        spec.loader.exec_module(mod)
    finally:
        if saved is not None:
            sys.modules[raster_name] = saved
        else:
            sys.modules.pop(raster_name, None)
    return mod.render


def _grad_ok(name, ours, ref_runs):
    st = torch.stack([r.double().reshape(-1) for r in ref_runs])
    mean = st.mean(0)
    spread = float((st.max(0).values - st.min(0).values).max())
    tol = 1e-6 * float(mean.abs().max()) + 2.0 * spread + 1e-4 * mean.abs()
    err = (ours.double().reshape(-1) - mean).abs()
    assert not bool((err > tol).any()), f"{name}: worst ratio {float((err / tol).max()):.3g}"


def _loss_and_grads(render, sc, cam, surfel, pipe, weights):
    pc = _Model(sc, surfel)
    out = render(_Camera(cam), pc, pipe, sc.bg)
    loss = sum((out[k] * w).sum() for k, w in weights.items())
    loss.backward()
    return out, [p.grad.clone() for p in pc.p] + [out["viewspace_points"].grad.clone()]


@pytest.mark.parametrize("pkg,surfel,depth_ratio", [("gs3dgs", False, 0.0), ("gs2dgs", True, 0.0), ("gs2dgs", True, 1.0)])
def test_reference_render_on_the_dropin_equals_reference_cuda(pkg, surfel, depth_ratio):
    from oracle import ref_cuda
    if not ref_cuda.available("dsr" if surfel else "dgr"):
        pytest.skip("oracle/_ref not built")
    import scorp_b200
    import ref_modules
    scorp_b200.install()
    raster_name = "diff_surfel_rasterization" if surfel else "diff_gaussian_rasterization"
    sys.modules.pop(raster_name, None)
    ours_mod = __import__(raster_name)                              # what SCORP's import statement finds after install()
    assert "scorp_b200" in ours_mod.GaussianRasterizer.__module__
    render_ours = _load_render(pkg, raster_name, ours_mod, f"_{pkg}_ours")
    render_ref = _load_render(pkg, raster_name, ref_modules.reference_backed_module(surfel), f"_{pkg}_ref")

    sc = h.make("C4" if surfel else "C2", P=60_000, W=400, H=260, random_bg=True, n_views=2); cam = sc.cameras[1]
    pipe = _Pipe(); pipe.depth_ratio = depth_ratio
    keys = (["render", "render_alpha", "render_normal", "render_dist", "render_depth", "surf_normal"] if surfel
            else ["render", "render_depth", "render_alpha"])
    g = torch.Generator().manual_seed(17)
    weights = None
    out_o = out_r = None
    runs = []
    for i in range(4):                                             # reference four times: atomics spread
        if weights is None:
            pc = _Model(sc, surfel)
            probe = render_ref(_Camera(cam), pc, pipe, sc.bg)
            weights = {k: (torch.randn(probe[k].shape, generator=g) / probe[k].numel()).cuda() for k in keys}
        out_r, gr = _loss_and_grads(render_ref, sc, cam, surfel, pipe, weights)
        runs.append(gr)
    out_o, go = _loss_and_grads(render_ours, sc, cam, surfel, pipe, weights)
    assert set(out_o) == set(out_r)
    for k in out_r:
        a, b = out_o[k].detach(), out_r[k].detach()
        assert a.shape == b.shape and a.dtype == b.dtype, k
        if k == "viewspace_points":
            continue
        assert torch.equal(a, b), f"{k}: {int((a != b).sum())} of {a.numel()} elements differ"
    assert int(out_o["visibility_filter"].sum()) > 1000
    for i, name in enumerate(("xyz", "features", "opacity", "scaling", "rotation", "viewspace_points")):
        _grad_ok(name, go[i], [r[i] for r in runs])
