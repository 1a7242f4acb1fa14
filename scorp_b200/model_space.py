"""Model-space rasterizer entry (caller-side parameter staging), opt-in.

What the reference's caller does right before the rasterizer boundary, every iteration
(gs3dgs/gaussian_renderer/__init__.py:67-90 through the GaussianModel properties,
gs3dgs/scene/gaussian_model.py:126-146):

    shs       = torch.cat((pc._features_dc, pc._features_rest), dim=1)     # 192 MB read + 192 MB write at C2
    opacity   = torch.sigmoid(pc._opacity)
    scales    = torch.exp(pc._scaling)
    rotations = torch.nn.functional.normalize(pc._rotation)

and autograd mirrors each of them in the backward (the cat becomes two slicing copies of the 192 MB SH gradient).
Here the rasterizer takes the STORED tensors, applies the activations inside its preprocess kernel, and its backward
writes the gradients of the stored tensors directly (features_dc / features_rest rows separately, sigmoid / exp /
normalize chained in registers): gsr_rasterize_forward_raw / gsr_rasterize_backward_raw in include/gsr_b200.h.

`rasterize_model(...)` is the functional form; `render(viewpoint_camera, pc, pipe, bg_color, ...)` mirrors
gaussian_renderer.render() for a GaussianModel-like object (same returned dict keys) and falls back to the stock
drop-in path for the option combinations the model-space kernels do not cover (override_color, compute_cov3D_python,
convert_SHs_python) -- that is the stock CUDA path of this library, not a CPU path.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import ptr
from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from .diff_gaussian_rasterization import _C as _dgr

__all__ = ["rasterize_model", "rasterize_surfel_model", "render", "model_forward", "model_backward",
           "surfel_model_forward", "surfel_model_backward", "MODEL_GRAD_FLOATS"]


def MODEL_GRAD_FLOATS(M: int) -> int:
    """floats per Gaussian at the head of the gradient arena: xyz, features_dc, features_rest, opacity, scaling,
    rotation (the same 59 at SH3 as the stock arena, so the multi-GPU exchange of multiview.py applies unchanged)."""
    return 3 + 3 + 3 * (M - 1) + 1 + 3 + 4


def _params(xyz, fdc, frest, op, sc, rot):
    p = _lib.GsrModelParams()
    p.xyz, p.features_dc, p.features_rest = ptr(xyz), ptr(fdc), ptr(frest)
    p.opacity, p.scaling, p.rotation = ptr(op), ptr(sc), ptr(rot)
    return p


def model_forward(xyz, features_dc, features_rest, opacity, scaling, rotation, rs):
    """`_C`-level forward on detached tensors -> (num_rendered, color, depth, alpha, radii, geom, binning, image,
    inputs) where `inputs` are the contiguous fp32 tensors the kernels read (kept for backward)."""
    if xyz.dim() != 2 or xyz.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not xyz.is_cuda:
        raise _lib.GsrError("xyz must be a CUDA tensor: this rasterizer has no CPU path")
    lib = _lib.lib()
    dev = xyz.device
    f = lambda t: _dgr._f32c(t.detach(), dev)
    xyz_, fdc, frest, op, sc, rot = map(f, (xyz, features_dc, features_rest, opacity, scaling, rotation))
    P, H, W = xyz_.size(0), int(rs.image_height), int(rs.image_width)
    M = 1 + (frest.size(1) if frest.numel() else 0)
    if fdc.numel() != P * 3 or frest.numel() != P * (M - 1) * 3:
        raise RuntimeError("features_dc must be (P, 1, 3) and features_rest (P, M-1, 3)")
    cam = tuple(_dgr._f32c(t, dev) for t in (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
    color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _dgr._Arena(dev)
    s = _dgr._settings(*cam, rs.tanfovx, rs.tanfovy, H, W, rs.scale_modifier, rs.sh_degree, rs.prefiltered, rs.debug)
    mp = _params(xyz_, fdc, frest, op, sc, rot)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        rc = lib.gsr_rasterize_forward_raw(C.byref(s), P, M, C.byref(mp), color.data_ptr(), depth.data_ptr(),
                                           alpha.data_ptr(), ptr(radii), arena.fn, None, C.byref(n), _dgr._stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_rasterize_forward_raw")
    return (n.value, color, depth, alpha, radii, arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING),
            arena.get(_lib.GSR_BUF_IMAGE), (xyz_, fdc, frest, op, sc, rot) + cam)


def model_backward(inputs, rs, num_rendered, radii, geom, binning, img, alpha, g_color, g_depth, g_alpha):
    """`_C`-level backward -> (g_xyz, g_means2D, g_dc, g_rest, g_opacity, g_scaling, g_rotation), flat views of one
    arena whose first P * MODEL_GRAD_FLOATS(M) floats are the parameter gradients (one all-reduce in a view-sharded
    trainer)."""
    (xyz, fdc, frest, op, sc, rot, bg, view, proj, campos) = inputs
    lib = _lib.lib()
    dev = xyz.device
    P, H, W = xyz.size(0), int(rs.image_height), int(rs.image_width)
    M = 1 + (frest.size(1) if frest.numel() else 0)
    z = lambda c: torch.zeros((c, H, W), dtype=torch.float32, device=dev)
    g_color = z(3) if g_color is None else _dgr._f32c(g_color, dev)
    g_depth = z(1) if g_depth is None else _dgr._f32c(g_depth, dev)
    g_alpha = z(1) if g_alpha is None else _dgr._f32c(g_alpha, dev)
    widths = (3, 3, 3 * (M - 1), 1, 3, 4, 3)      # xyz, dc, rest, opacity, scaling, rotation | means2D
    flat = torch.empty((P * sum(widths),), dtype=torch.float32, device=dev)
    outs, o = [], 0
    for w in widths:
        outs.append(flat[o:o + P * w]); o += P * w
    g_xyz, g_dc, g_rest, g_op, g_sc, g_rot, g_m2d = outs
    if P != 0:
        ws_bytes = lib.gsr_backward_workspace_bytes(P)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        s = _dgr._settings(bg, view, proj, campos, rs.tanfovx, rs.tanfovy, H, W, rs.scale_modifier, rs.sh_degree,
                           False, rs.debug)
        mp = _params(xyz, fdc, frest, op, sc, rot)
        mg = _lib.GsrModelGrads()
        mg.xyz, mg.features_dc, mg.features_rest = ptr(g_xyz), ptr(g_dc), ptr(g_rest)
        mg.opacity, mg.scaling, mg.rotation = ptr(g_op), ptr(g_sc), ptr(g_rot)
        with torch.cuda.device(dev):
            rc = lib.gsr_rasterize_backward_raw(
                C.byref(s), P, M, int(num_rendered), C.byref(mp), ptr(radii), ptr(geom), ptr(binning), ptr(img),
                ptr(alpha), ptr(g_color), ptr(g_depth), ptr(g_alpha), ptr(g_m2d), C.byref(mg), ws.data_ptr(),
                ws_bytes, _dgr._stream(dev))
        _lib.check(rc, "gsr_rasterize_backward_raw")
        if _dgr._backward_hook[0] is not None:     # multiview.install_backward_allreduce
            _dgr._backward_hook[0](flat[: P * MODEL_GRAD_FLOATS(M)])
    return g_xyz, g_m2d, g_dc, g_rest, g_op, g_sc, g_rot


class _RasterizeModel(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings):
        (n, color, depth, alpha, radii, geom, binning, img, inputs) = model_forward(
            xyz, features_dc, features_rest, opacity, scaling, rotation, raster_settings)
        ctx.rs, ctx.num_rendered = raster_settings, n
        ctx.shapes = (features_dc.shape, features_rest.shape, opacity.shape)
        ctx.save_for_backward(radii, geom, binning, img, alpha, *inputs)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, g_color, _g_radii, g_depth, g_alpha):
        radii, geom, binning, img, alpha, *inputs = ctx.saved_tensors
        P = inputs[0].size(0)
        g_xyz, g_m2d, g_dc, g_rest, g_op, g_sc, g_rot = model_backward(
            tuple(inputs), ctx.rs, ctx.num_rendered, radii, geom, binning, img, alpha, g_color, g_depth, g_alpha)
        dc_shape, rest_shape, op_shape = ctx.shapes
        return (g_xyz.view(P, 3), g_m2d.view(P, 3), g_dc.view(dc_shape), g_rest.view(rest_shape), g_op.view(op_shape),
                g_sc.view(P, 3), g_rot.view(P, 4), None)


# ---- 2DGS surfels: the same staging lives in gs2dgs/scene/gaussian_model.py (scaling is [P, 2]) ------------------

def surfel_model_forward(xyz, features_dc, features_rest, opacity, scaling, rotation, rs):
    """`_C`-level forward of the surfel rasterizer on the stored tensors -> (num_rendered, color[3,H,W],
    others[7,H,W], radii, geom, binning, image, inputs)."""
    if xyz.dim() != 2 or xyz.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not xyz.is_cuda:
        raise _lib.GsrError("xyz must be a CUDA tensor: this rasterizer has no CPU path")
    lib = _lib.lib()
    dev = xyz.device
    f = lambda t: _dgr._f32c(t.detach(), dev)
    xyz_, fdc, frest, op, sc, rot = map(f, (xyz, features_dc, features_rest, opacity, scaling, rotation))
    P, H, W = xyz_.size(0), int(rs.image_height), int(rs.image_width)
    M = 1 + (frest.size(1) if frest.numel() else 0)
    if fdc.numel() != P * 3 or frest.numel() != P * (M - 1) * 3 or sc.numel() != P * 2:
        raise RuntimeError("features_dc must be (P, 1, 3), features_rest (P, M-1, 3) and scaling (P, 2)")
    cam = tuple(_dgr._f32c(t, dev) for t in (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
    color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    others = torch.empty((7, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _dgr._Arena(dev)
    s = _dgr._settings(*cam, rs.tanfovx, rs.tanfovy, H, W, rs.scale_modifier, rs.sh_degree, rs.prefiltered, rs.debug)
    mp = _params(xyz_, fdc, frest, op, sc, rot)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        rc = lib.gsr_surfel_forward_raw(C.byref(s), P, M, C.byref(mp), color.data_ptr(), others.data_ptr(), ptr(radii),
                                        arena.fn, None, C.byref(n), _dgr._stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_surfel_forward_raw")
    return (n.value, color, others, radii, arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING),
            arena.get(_lib.GSR_BUF_IMAGE), (xyz_, fdc, frest, op, sc, rot) + cam)


def surfel_model_backward(inputs, rs, num_rendered, radii, geom, binning, img, g_color, g_others):
    """-> (g_xyz, g_means2D, g_dc, g_rest, g_opacity, g_scaling, g_rotation), parameter gradients first in one arena."""
    (xyz, fdc, frest, op, sc, rot, bg, view, proj, campos) = inputs
    lib = _lib.lib()
    dev = xyz.device
    P, H, W = xyz.size(0), int(rs.image_height), int(rs.image_width)
    M = 1 + (frest.size(1) if frest.numel() else 0)
    z = lambda c: torch.zeros((c, H, W), dtype=torch.float32, device=dev)
    g_color = z(3) if g_color is None else _dgr._f32c(g_color, dev)
    g_others = z(7) if g_others is None else _dgr._f32c(g_others, dev)
    widths = (3, 3, 3 * (M - 1), 1, 2, 4, 3)      # xyz, dc, rest, opacity, scaling, rotation | means2D
    flat = torch.empty((P * sum(widths),), dtype=torch.float32, device=dev)
    outs, o = [], 0
    for w in widths:
        outs.append(flat[o:o + P * w]); o += P * w
    g_xyz, g_dc, g_rest, g_op, g_sc, g_rot, g_m2d = outs
    if P != 0:
        ws_bytes = lib.gsr_surfel_backward_workspace_bytes(P)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        s = _dgr._settings(bg, view, proj, campos, rs.tanfovx, rs.tanfovy, H, W, rs.scale_modifier, rs.sh_degree,
                           False, rs.debug)
        mp = _params(xyz, fdc, frest, op, sc, rot)
        mg = _lib.GsrModelGrads()
        mg.xyz, mg.features_dc, mg.features_rest = ptr(g_xyz), ptr(g_dc), ptr(g_rest)
        mg.opacity, mg.scaling, mg.rotation = ptr(g_op), ptr(g_sc), ptr(g_rot)
        with torch.cuda.device(dev):
            rc = lib.gsr_surfel_backward_raw(
                C.byref(s), P, M, int(num_rendered), C.byref(mp), ptr(radii), ptr(geom), ptr(binning), ptr(img),
                ptr(g_color), ptr(g_others), ptr(g_m2d), C.byref(mg), ws.data_ptr(), ws_bytes, _dgr._stream(dev))
        _lib.check(rc, "gsr_surfel_backward_raw")
        from .diff_surfel_rasterization import _C as _dsr
        if _dsr._backward_hook[0] is not None:
            _dsr._backward_hook[0](flat[: P * (MODEL_GRAD_FLOATS(M) - 1)])
    return g_xyz, g_m2d, g_dc, g_rest, g_op, g_sc, g_rot


class _RasterizeSurfelModel(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings):
        (n, color, others, radii, geom, binning, img, inputs) = surfel_model_forward(
            xyz, features_dc, features_rest, opacity, scaling, rotation, raster_settings)
        ctx.rs, ctx.num_rendered = raster_settings, n
        ctx.shapes = (features_dc.shape, features_rest.shape, opacity.shape)
        ctx.save_for_backward(radii, geom, binning, img, *inputs)
        ctx.mark_non_differentiable(radii)
        return color, radii, others

    @staticmethod
    def backward(ctx, g_color, _g_radii, g_others):
        radii, geom, binning, img, *inputs = ctx.saved_tensors
        P = inputs[0].size(0)
        g_xyz, g_m2d, g_dc, g_rest, g_op, g_sc, g_rot = surfel_model_backward(
            tuple(inputs), ctx.rs, ctx.num_rendered, radii, geom, binning, img, g_color, g_others)
        dc_shape, rest_shape, op_shape = ctx.shapes
        return (g_xyz.view(P, 3), g_m2d.view(P, 3), g_dc.view(dc_shape), g_rest.view(rest_shape), g_op.view(op_shape),
                g_sc.view(P, 2), g_rot.view(P, 4), None)


def rasterize_surfel_model(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings):
    """2DGS: -> (color[3,H,W], radii[P], allmap[7,H,W]); equals diff_surfel_rasterization.GaussianRasterizer(rs)(xyz,
    means2D, sigmoid(opacity), shs=cat(features_dc, features_rest), scales=exp(scaling), rotations=normalize(rotation))
    -- the call of gs2dgs/gaussian_renderer/__init__.py:112-121 -- without the torch staging kernels."""
    return _RasterizeSurfelModel.apply(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation,
                                       raster_settings)


def rasterize_model(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings):
    """-> (color[3,H,W], radii[P] int32, depth[1,H,W], alpha[1,H,W]); equals
    GaussianRasterizer(raster_settings)(xyz, means2D, sigmoid(opacity), shs=cat(features_dc, features_rest),
    scales=exp(scaling), rotations=normalize(rotation))."""
    return _RasterizeModel.apply(xyz, means2D, features_dc, features_rest, opacity, scaling, rotation, raster_settings)


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None):
    """gaussian_renderer.render() (gs3dgs/gaussian_renderer/__init__.py:24-135) reading the model's stored tensors.
    Returns the same dict: render, viewspace_points, visibility_filter, radii, render_depth, render_alpha."""
    xyz = pc.get_xyz
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    w, h = viewpoint_camera.resolution
    rs = GaussianRasterizationSettings(
        image_height=int(h), image_width=int(w),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
        bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=bool(getattr(pipe, "debug", False)))
    stock = (override_color is not None or getattr(pipe, "compute_cov3D_python", False)
             or getattr(pipe, "convert_SHs_python", False))
    if stock:
        cov = pc.get_covariance(scaling_modifier) if getattr(pipe, "compute_cov3D_python", False) else None
        if override_color is None and getattr(pipe, "convert_SHs_python", False):
            raise NotImplementedError("convert_SHs_python: use the stock gaussian_renderer.render()")
        image, radii, depth, alpha = GaussianRasterizer(rs)(
            means3D=xyz, means2D=screenspace_points, opacities=pc.get_opacity,
            shs=pc.get_features if override_color is None else None, colors_precomp=override_color,
            scales=None if cov is not None else pc.get_scaling, rotations=None if cov is not None else pc.get_rotation,
            cov3D_precomp=cov)
    else:
        image, radii, depth, alpha = rasterize_model(xyz, screenspace_points, pc._features_dc, pc._features_rest,
                                                     pc._opacity, pc._scaling, pc._rotation, rs)
    depth = torch.nan_to_num(depth / alpha, 0, 0)
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii,
            "render_depth": depth, "render_alpha": alpha}
