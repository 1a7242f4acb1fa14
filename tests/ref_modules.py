"""TEST INFRASTRUCTURE. Modules shaped like `diff_gaussian_rasterization` / `diff_surfel_rasterization` but backed by
the UNMODIFIED reference CUDA (oracle/_ref via oracle/ref_cuda.py), so the reference's own gaussian_renderer.render()
can be executed on top of the reference kernels on the GPU box (the reference's pybind glue is not built here). The
autograd mapping follows DGR diff_gaussian_rasterization/__init__.py:44-158 and DSR :44-156: gradients are returned in
input order, `means2D` receives the screen-space gradient."""
from __future__ import annotations

import types
from typing import NamedTuple

import torch
import torch.nn as nn

from oracle import ref_cuda


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _n(t):
    return None if t is None or t.numel() == 0 else t.contiguous()


def _make(surfel: bool):
    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, sh, colors, opacities, scales, rotations, cov, rs):
            fwd = (ref_cuda.dsr_forward if surfel else ref_cuda.dgr_forward)(
                rs.bg.contiguous(), means3D.contiguous(), _n(colors), opacities.contiguous(), _n(scales), _n(rotations),
                rs.scale_modifier, _n(cov), rs.viewmatrix.contiguous(), rs.projmatrix.contiguous(), rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, _n(sh), rs.sh_degree, rs.campos.contiguous())
            ctx.rs, ctx.fwd = rs, fwd
            ctx.args = (means3D.contiguous(), _n(colors), _n(scales), _n(rotations), _n(cov), _n(sh))
            ctx.mark_non_differentiable(fwd["radii"])
            if surfel:
                return fwd["color"], fwd["radii"], fwd["others"]
            return fwd["color"], fwd["radii"], fwd["depth"], fwd["alpha"]

        @staticmethod
        def backward(ctx, g_color, g_radii, *g_rest):
            rs, fwd = ctx.rs, ctx.fwd
            means3D, colors, scales, rotations, cov, sh = ctx.args
            H, W = rs.image_height, rs.image_width
            z = lambda c: torch.zeros(c, H, W, device=means3D.device)
            g_color = z(3) if g_color is None else g_color.contiguous()
            if surfel:
                g_o = z(7) if g_rest[0] is None else g_rest[0].contiguous()
                g = ref_cuda.dsr_backward(fwd, rs.bg.contiguous(), means3D, colors, scales, rotations, rs.scale_modifier, cov,
                                          rs.viewmatrix.contiguous(), rs.projmatrix.contiguous(), rs.tanfovx, rs.tanfovy,
                                          g_color, g_o, sh, rs.sh_degree, rs.campos.contiguous())
                g_cov = g["transMat"]
            else:
                g_d = z(1) if g_rest[0] is None else g_rest[0].contiguous()
                g_a = z(1) if g_rest[1] is None else g_rest[1].contiguous()
                g = ref_cuda.dgr_backward(fwd, rs.bg.contiguous(), means3D, colors, scales, rotations, rs.scale_modifier, cov,
                                          rs.viewmatrix.contiguous(), rs.projmatrix.contiguous(), rs.tanfovx, rs.tanfovy,
                                          g_color, g_d, g_a, sh, rs.sh_degree, rs.campos.contiguous())
                g_cov = g["cov3D"]
            return (g["means3D"], g["means2D"], g["sh"], g["colors"], g["opacity"], g["scales"], g["rotations"], g_cov, None)

    class GaussianRasterizer(nn.Module):
        def __init__(self, raster_settings):
            super().__init__()
            self.raster_settings = raster_settings

        def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                    cov3D_precomp=None):
            e = torch.empty(0, device=means3D.device)
            pick = lambda t: e if t is None else t
            return _Fn.apply(means3D, means2D, pick(shs), pick(colors_precomp), opacities, pick(scales), pick(rotations),
                             pick(cov3D_precomp), self.raster_settings)

    m = types.ModuleType("diff_surfel_rasterization" if surfel else "diff_gaussian_rasterization")
    m.GaussianRasterizationSettings = GaussianRasterizationSettings
    m.GaussianRasterizer = GaussianRasterizer
    return m


def reference_backed_module(surfel: bool):
    return _make(surfel)
